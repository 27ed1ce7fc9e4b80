set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest86.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest86.log
for w in c2 c5 scene1; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r86_${w}.json 2>> gpurun_out/r86.err
done
