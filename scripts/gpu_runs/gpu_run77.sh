set -x
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -x -q -k "top_level_tree" > gpurun_out/pytest77.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest77.log
