set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest74.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest74.log
for w in scene8 scene7 scene3 scene5 c5 mesh; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r74_${w}.json 2>> gpurun_out/r74.err
done
