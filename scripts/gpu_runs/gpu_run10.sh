set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest10.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest10.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r10_c2.json 2>> gpurun_out/r10.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r10_c1.json 2>> gpurun_out/r10.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r10_c5.json 2>> gpurun_out/r10.err
timeout 300 python scripts/mesh_probe.py 16 16 > gpurun_out/r10_mesh_probe.txt 2>&1
timeout 300 python scripts/mesh_probe.py 4 12 >> gpurun_out/r10_mesh_probe.txt 2>&1
