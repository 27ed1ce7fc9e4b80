set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest30.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest30.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r30_c2.json 2>> gpurun_out/r30.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r30_c1.json 2>> gpurun_out/r30.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r30_c5.json 2>> gpurun_out/r30.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r30_mesh.json 2>> gpurun_out/r30.err
