set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest69.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest69.log
for k in 3 5 7 8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r69_scene$k.json 2>> gpurun_out/r69.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r69_c5.json 2>> gpurun_out/r69.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r69_mesh.json 2>> gpurun_out/r69.err
