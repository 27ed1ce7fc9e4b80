set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest29.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest29.log
timeout 600 python bench.py > gpurun_out/r29_c2_default.json 2> gpurun_out/r29_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r29_c1.json 2>> gpurun_out/r29.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r29_c5.json 2>> gpurun_out/r29.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r29_mesh.json 2>> gpurun_out/r29.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r29_ref_c2.json 2>> gpurun_out/r29.err
python __graft_entry__.py --smoke > gpurun_out/r29_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r29_smoke.log
