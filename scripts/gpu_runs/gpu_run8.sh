set -x
python -m pytest tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest8.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest8.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r8_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 10 -o gpurun_out/r8_prof_traverse $CMD > gpurun_out/r8_ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 6 -o gpurun_out/r8_prof_shade $CMD > gpurun_out/r8_ncu_full2.log 2>&1
