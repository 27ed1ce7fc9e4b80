# full ncu capture of the pooled unified kernel (mesh workload, trigger 24), first two traversal launches
set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse_unified -c 2 -o gpurun_out/r2_33_mesh_pooled -f python bench.py --steps 1 --warmup 0 --no-extras --workload mesh --leaf-pool 24 > gpurun_out/r2_33_ncu.log 2>&1
tail -3 gpurun_out/r2_33_ncu.log
timeout 300 python -m pytest tests/test_gpu_unified.py -m gpu -q -x -k pooled > gpurun_out/r2_33_pytest.log 2>&1; tail -3 gpurun_out/r2_33_pytest.log
du -sh gpurun_out
