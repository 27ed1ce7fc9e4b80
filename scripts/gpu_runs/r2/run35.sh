# full ncu capture of the pooled flat kernel at C2 (second and third bounce)
set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse_flat_pooled -c 2 -o gpurun_out/r2_35_c2_pooled -f python bench.py --steps 1 --warmup 0 --no-extras --workload c2 > gpurun_out/r2_35_ncu.log 2>&1
tail -3 gpurun_out/r2_35_ncu.log
du -sh gpurun_out
