set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r38b_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r38_prof_traverse $CMD > gpurun_out/r38_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 3 -o gpurun_out/r38_prof_shade $CMD > gpurun_out/r38_ncu3.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
$M > gpurun_out/r38_mesh_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 3 -o gpurun_out/r38_prof_mesh $M > gpurun_out/r38_ncu4.log 2>&1
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
$C5 > gpurun_out/r38_c5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 2 -o gpurun_out/r38_prof_c5 $C5 > gpurun_out/r38_ncu5.log 2>&1
ls -la gpurun_out
