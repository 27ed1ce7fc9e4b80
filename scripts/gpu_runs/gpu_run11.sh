set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r11_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r11_launches.csv $CMD > gpurun_out/r11_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 6 -o gpurun_out/r11_prof_traverse $CMD > gpurun_out/r11_ncu2.log 2>&1
M="python scripts/mesh_probe.py 16 16 1"
$M > gpurun_out/r11_mesh_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 3 -o gpurun_out/r11_prof_mesh $M > gpurun_out/r11_ncu3.log 2>&1
