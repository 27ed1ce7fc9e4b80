set -x
for b in 8 32 64; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --ppb $b > gpurun_out/r61_c2_ppb$b.json 2>> gpurun_out/r61.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --ppb 32 > gpurun_out/r61_c5_ppb32.json 2>> gpurun_out/r61.err
