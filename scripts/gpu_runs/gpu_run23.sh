set -x
V=gpurun_variants
for v in leaf1 leaf2r16 leaf3 refill24 refill32; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r23_mesh_$v.json 2>> gpurun_out/r23.err
done
for v in refill24 refill32; do
MCPT_LIBRARY=$PWD/$V/libmcpt_$v.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r23_c5_$v.json 2>> gpurun_out/r23.err
done
