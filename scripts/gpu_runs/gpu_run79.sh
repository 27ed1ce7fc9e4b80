set -x
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -x -q -k "rotated_elongated" > gpurun_out/pytest79.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest79.log
