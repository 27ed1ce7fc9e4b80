set -x
( time timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q --durations=5 ) > gpurun_out/pytest66.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest66.log
