set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest50.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest50.log
for w in c2 c1 c5 mesh; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r50_${w}_rec.json 2>> gpurun_out/r50.err
  MCPT_LIBRARY=$V/soa/libmcpt.so timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r50_${w}_soa.json 2>> gpurun_out/r50.err
done
