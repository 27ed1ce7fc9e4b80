set -x
V=gpurun_variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest34.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest34.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r34_c2.json 2>> gpurun_out/r34.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r34_c1.json 2>> gpurun_out/r34.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r34_c5.json 2>> gpurun_out/r34.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r34_mesh.json 2>> gpurun_out/r34.err
for v in node4 node6 node8; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r34_mesh_$v.json 2>> gpurun_out/r34.err
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r34_c5_$v.json 2>> gpurun_out/r34.err
done
