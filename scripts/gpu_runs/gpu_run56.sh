set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest56.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest56.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r56_c2_pipe.json 2>> gpurun_out/r56.err
for v in pipe_minb3 nopipe; do
  MCPT_LIBRARY=$V/$v/libmcpt.so timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r56_c2_$v.json 2>> gpurun_out/r56.err
done
MCPT_LIBRARY=$V/pipe_minb3/libmcpt.so timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c5 > gpurun_out/r56_c5_pipe_minb3.json 2>> gpurun_out/r56.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c5 > gpurun_out/r56_c5_pipe.json 2>> gpurun_out/r56.err
