set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest62.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest62.log
for w in c2 c1 c5 mesh; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r62_${w}_split.json 2>> gpurun_out/r62.err
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w --split-next 0 > gpurun_out/r62_${w}_nosplit.json 2>> gpurun_out/r62.err
done
