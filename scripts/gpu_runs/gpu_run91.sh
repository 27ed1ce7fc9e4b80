set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest91.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest91.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r91_c2.json 2> gpurun_out/r91.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene8 > gpurun_out/r91_scene8.json 2>> gpurun_out/r91.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r91_mesh.json 2>> gpurun_out/r91.err
python __graft_entry__.py --smoke > gpurun_out/r91_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r91_smoke.log
