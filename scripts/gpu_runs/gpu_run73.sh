set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest73.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest73.log
for w in c5 scene3 scene5 scene7 scene8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r73_${w}.json 2>> gpurun_out/r73.err
done
MCPT_LIBRARY=$V/steps8/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r73_c5_steps8.json 2>> gpurun_out/r73.err
