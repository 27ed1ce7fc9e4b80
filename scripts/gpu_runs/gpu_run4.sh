set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest4.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest4.log
python bench.py > gpurun_out/r4_bench.json 2> gpurun_out/r4_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r4_ref.json 2> gpurun_out/r4_ref.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r4_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r4_launches.csv $CMD > gpurun_out/r4_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 40 -c 3 -o gpurun_out/r4_prof_traverse $CMD > gpurun_out/r4_ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 40 -c 2 -o gpurun_out/r4_prof_shade $CMD > gpurun_out/r4_ncu_full2.log 2>&1
