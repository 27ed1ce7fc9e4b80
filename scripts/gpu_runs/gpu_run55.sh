set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r55_prof_traverse $CMD > gpurun_out/r55_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 3 -o gpurun_out/r55_prof_shade $CMD > gpurun_out/r55_ncu3.log 2>&1
