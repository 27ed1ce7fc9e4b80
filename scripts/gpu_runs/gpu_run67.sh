set -x
( time timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_render.py tests/test_gpu_traversal.py -m gpu -x -q -k "mesh or persistent or flat or queue_order or additive" ) > gpurun_out/sanitizer67.log 2>&1; echo "sanitizer exit $?" >> gpurun_out/sanitizer67.log
