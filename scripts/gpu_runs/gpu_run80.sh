set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest80.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest80.log
timeout 600 python bench.py > gpurun_out/r80_c2_default.json 2> gpurun_out/r80_c2_default.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r80_ref_c2.json 2>> gpurun_out/r80.err
timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline --workload fps1 > gpurun_out/r80_fps1.json 2>> gpurun_out/r80.err
python __graft_entry__.py --smoke > gpurun_out/r80_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r80_smoke.log
