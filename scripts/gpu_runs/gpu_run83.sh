set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest83.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest83.log
for w in c2 c5 scene2 scene4 scene8 c1; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r83_${w}.json 2>> gpurun_out/r83.err
done
