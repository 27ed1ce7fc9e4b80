set -x
V=gpurun_variants
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r28_mesh_pops3.json 2>> gpurun_out/r28.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r28_c5_pops3.json 2>> gpurun_out/r28.err
for v in pops2 pops6 pops64; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r28_mesh_$v.json 2>> gpurun_out/r28.err
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r28_c5_$v.json 2>> gpurun_out/r28.err
done
