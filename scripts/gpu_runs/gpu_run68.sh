set -x
for k in 1 2 3 4 5 6 7 8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r68_scene$k.json 2>> gpurun_out/r68.err
done
