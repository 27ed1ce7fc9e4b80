set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r16_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r16_prof_traverse $CMD > gpurun_out/r16_ncu2.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
$M > gpurun_out/r16_mesh_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r16_prof_mesh $M > gpurun_out/r16_ncu3.log 2>&1
