set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_06_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_06_pytest.log
tail -30 gpurun_out/r2_06_pytest.log
