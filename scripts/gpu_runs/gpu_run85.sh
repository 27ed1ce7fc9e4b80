set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest85.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest85.log
timeout 600 python bench.py > gpurun_out/r85_c2_default.json 2> gpurun_out/r85_c2_default.err
for w in c1 c5 mesh; do
  timeout 300 python bench.py --steps 5 --workload $w > gpurun_out/r85_${w}.json 2>> gpurun_out/r85.err
done
for k in 1 2 4 6; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r85_scene$k.json 2>> gpurun_out/r85.err
done
