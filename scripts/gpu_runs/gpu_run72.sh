set -x
V=montecarlo_pathtracing_b200/lib/variants
for w in scene8 c5; do
  for r in 8 16 24; do
    timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w --refill $r > gpurun_out/r72_${w}_refill$r.json 2>> gpurun_out/r72.err
  done
  for v in steps2 steps6; do
    MCPT_LIBRARY=$V/$v/libmcpt.so timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w > gpurun_out/r72_${w}_$v.json 2>> gpurun_out/r72.err
  done
done
