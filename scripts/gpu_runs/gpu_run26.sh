set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r26_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r26_launches.csv $CMD > gpurun_out/r26_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 6 -o gpurun_out/r26_prof_traverse $CMD > gpurun_out/r26_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 6 -o gpurun_out/r26_prof_shade $CMD > gpurun_out/r26_ncu3.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
$M > gpurun_out/r26_mesh_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r26_prof_mesh $M > gpurun_out/r26_ncu4.log 2>&1
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
$C5 > gpurun_out/r26_c5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r26_prof_c5 $C5 > gpurun_out/r26_ncu5.log 2>&1
