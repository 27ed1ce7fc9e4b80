set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest42.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest42.log
timeout 300 python bench.py > gpurun_out/r42_c2.json 2>> gpurun_out/r42.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r42_ref.json 2>> gpurun_out/r42.err
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r42_smoke.log 2>&1
