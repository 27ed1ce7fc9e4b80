set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest12.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest12.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r12_c2.json 2>> gpurun_out/r12.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r12_c1.json 2>> gpurun_out/r12.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r12_c5.json 2>> gpurun_out/r12.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r12_mesh.json 2>> gpurun_out/r12.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ 3 > gpurun_out/r12_mesh_occ3.json 2>> gpurun_out/r12.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --occ 3 > gpurun_out/r12_c5_occ3.json 2>> gpurun_out/r12.err
