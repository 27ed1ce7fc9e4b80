set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest46.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest46.log
for w in mesh c5 c4; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r46_${w}.json 2>> gpurun_out/r46.err
done
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 3 -o gpurun_out/r46_prof_mesh $M > gpurun_out/r46_ncu.log 2>&1
