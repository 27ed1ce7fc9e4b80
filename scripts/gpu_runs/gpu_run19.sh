set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest19.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest19.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r19_c2.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r19_c1.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r19_mesh.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ 2 > gpurun_out/r19_mesh_occ2.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r19_c5.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent 2 > gpurun_out/r19_c5_p2.json 2>> gpurun_out/r19.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent 2 --occ 4 > gpurun_out/r19_c5_p2_occ4.json 2>> gpurun_out/r19.err
