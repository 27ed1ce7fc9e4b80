timeout 600 python -m pytest tests/test_gpu_nee.py -m gpu -q -x -s > gpurun_out/r2_18_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_18_pytest.log
grep -E "RMSE|rays|passed|failed|Error|assert|exit" gpurun_out/r2_18_pytest.log | tail -30
