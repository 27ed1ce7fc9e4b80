set -x
V=montecarlo_pathtracing_b200/lib/variants
for w in c2 scene1 scene2; do
  timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r90_${w}_base.json 2>> gpurun_out/r90.err
  MCPT_LIBRARY=$V/noround/libmcpt.so timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload $w > gpurun_out/r90_${w}_noround.json 2>> gpurun_out/r90.err
done
