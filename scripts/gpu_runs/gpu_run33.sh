set -x
V=gpurun_variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest33.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest33.log
for v in node2 node3; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r33_mesh_$v.json 2>> gpurun_out/r33.err
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r33_c5_$v.json 2>> gpurun_out/r33.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r33_mesh_node1.json 2>> gpurun_out/r33.err
