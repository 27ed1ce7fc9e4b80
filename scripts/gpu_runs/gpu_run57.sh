set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest57.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest57.log
for w in c2 c1; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r57_${w}_bins.json 2>> gpurun_out/r57.err
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w --flat-bins 0 > gpurun_out/r57_${w}_nobins.json 2>> gpurun_out/r57.err
done
