set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest44.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest44.log
for w in mesh c5; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r44_${w}_wide.json 2>> gpurun_out/r44.err
  for v in narrow wide_s1 wide_s3; do
    MCPT_LIBRARY=$V/$v/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r44_${w}_$v.json 2>> gpurun_out/r44.err
  done
done
