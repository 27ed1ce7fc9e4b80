MCPT_FLAT_POOL=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse_flat_pooled -c 2 -o gpurun_out/r2_28_pool -f python bench.py --steps 1 --warmup 0 --no-extras --workload c2 > gpurun_out/r2_28_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
