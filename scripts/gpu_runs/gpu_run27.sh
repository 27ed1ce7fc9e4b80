set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest27.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest27.log
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r27_mesh.json 2>> gpurun_out/r27.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r27_c5.json 2>> gpurun_out/r27.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --refill 16 > gpurun_out/r27_c5_r16.json 2>> gpurun_out/r27.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --refill 4 > gpurun_out/r27_mesh_r4.json 2>> gpurun_out/r27.err
timeout 600 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -s -k converges > gpurun_out/r27_converge.log 2>&1
