set -x
timeout 600 python -m pytest tests/test_gpu_unified.py tests/test_gpu_nee.py -m gpu -q -x > gpurun_out/r2_25_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_25_pytest.log
tail -15 gpurun_out/r2_25_pytest.log
true
