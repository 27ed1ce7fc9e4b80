set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest43.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest43.log
for w in mesh c5; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r43_${w}_stack64.json 2>> gpurun_out/r43.err
  MCPT_LIBRARY=$V/stack32/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r43_${w}_stack32.json 2>> gpurun_out/r43.err
done
timeout 300 python bench.py --steps 10 > gpurun_out/r43_c2.json 2>> gpurun_out/r43.err
