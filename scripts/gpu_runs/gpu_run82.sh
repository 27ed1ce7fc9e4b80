set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest82.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest82.log
for w in c2 c5 scene2 scene5 scene8 mesh; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r82_${w}_reject.json 2>> gpurun_out/r82.err
  MCPT_LIBRARY=$V/noreject/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r82_${w}_noreject.json 2>> gpurun_out/r82.err
done
