set -x
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest18.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest18.log
for pb in 1 4 8 16; do
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --prim-batch $pb > gpurun_out/r18_mesh_pb$pb.json 2>> gpurun_out/r18.err
done
for pb in 1 6 12; do
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent 2 --occ 4 --prim-batch $pb > gpurun_out/r18_c5_pb$pb.json 2>> gpurun_out/r18.err
done
