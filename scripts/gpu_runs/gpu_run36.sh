set -x
V=gpurun_variants
for v in np4 np6 np8; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r36_mesh_$v.json 2>> gpurun_out/r36.err
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r36_c5_$v.json 2>> gpurun_out/r36.err
done
