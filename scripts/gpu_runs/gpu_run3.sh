set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest3.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest3.log
python bench.py --steps 10 > gpurun_out/r3_c2.json 2> gpurun_out/r3_c2.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r3_smoke.log
