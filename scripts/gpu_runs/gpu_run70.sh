set -x
V=montecarlo_pathtracing_b200/lib/variants
for w in scene8 scene3 c5 scene7; do
  MCPT_LIBRARY=$V/wide_all/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r70_${w}_wide.json 2>> gpurun_out/r70.err
done
