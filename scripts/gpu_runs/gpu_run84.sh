set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest84.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest84.log
for w in c5 scene8 scene3 mesh; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r84_${w}.json 2>> gpurun_out/r84.err
done
