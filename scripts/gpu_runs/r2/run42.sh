# default run and the driver's N=1 command with the pooled flat walk in place
set -x
timeout 900 python bench.py --steps 100 > gpurun_out/r2_42_bench_default.json 2> gpurun_out/r2_42_bench_default.err; tail -2 gpurun_out/r2_42_bench_default.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_42_bench_driver.json 2> gpurun_out/r2_42_bench_driver.err; tail -2 gpurun_out/r2_42_bench_driver.err
cat gpurun_out/r2_42_bench_driver.json
