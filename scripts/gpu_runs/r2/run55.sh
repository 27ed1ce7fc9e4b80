# last check of the round on the final binary: whole GPU suite and smoke()
set -x
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/r2_55_pytest.log 2>&1; tail -2 gpurun_out/r2_55_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_55_smoke.log 2>&1; tail -1 gpurun_out/r2_55_smoke.log
