set -x
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest17.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest17.log
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r17_mesh.json 2>> gpurun_out/r17.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ 2 > gpurun_out/r17_mesh_occ2.json 2>> gpurun_out/r17.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent 2 > gpurun_out/r17_c5_p2.json 2>> gpurun_out/r17.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent 2 --occ 4 > gpurun_out/r17_c5_p2_occ4.json 2>> gpurun_out/r17.err
