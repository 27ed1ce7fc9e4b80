# final check of the round: whole GPU suite, smoke(), the default run and the driver's N = 1 command, the reference arm
set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_50_pytest.log 2>&1; tail -3 gpurun_out/r2_50_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_50_smoke.log 2>&1; tail -2 gpurun_out/r2_50_smoke.log
timeout 900 python bench.py --steps 100 > gpurun_out/r2_50_bench_default.json 2> gpurun_out/r2_50_bench_default.err; tail -1 gpurun_out/r2_50_bench_default.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_50_bench_driver.json 2> gpurun_out/r2_50_bench_driver.err; tail -1 gpurun_out/r2_50_bench_driver.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r2_50_bench_reference.json 2> gpurun_out/r2_50_bench_reference.err; cat gpurun_out/r2_50_bench_reference.json | cut -c1-400
