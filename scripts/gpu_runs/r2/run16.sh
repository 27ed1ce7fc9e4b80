timeout 900 python -m pytest tests/test_gpu_golden.py -m gpu -q -x -s -k "converges or sampler or sampling" > gpurun_out/r2_16_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_16_pytest.log
grep -E "RMSE|roughness|passed|failed|Error|assert" gpurun_out/r2_16_pytest.log | tail -30
