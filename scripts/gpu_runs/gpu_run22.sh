set -x
V=gpurun_variants
for v in refill4 refill16 leaf2 leaf8; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r22_mesh_$v.json 2>> gpurun_out/r22.err
done
for v in refill4 refill16; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r22_c5_$v.json 2>> gpurun_out/r22.err
done
for v in shade3 shade4; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r22_c2_$v.json 2>> gpurun_out/r22.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r22_mesh_base.json 2>> gpurun_out/r22.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r22_c5_base.json 2>> gpurun_out/r22.err
