set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r49_c2_base.json 2>> gpurun_out/r49.err
for m in 3 5 6; do
  MCPT_LIBRARY=$V/shade_minb$m/libmcpt.so timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r49_c2_minb$m.json 2>> gpurun_out/r49.err
done
