set -x
for k in 1 2 3 4 5 6 7 8; do
  timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --workload fps$k > gpurun_out/r78_fps$k.json 2>> gpurun_out/r78.err
done
