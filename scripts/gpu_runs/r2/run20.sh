# final evidence of the round: GPU tests, bench lines of every workload, launch lists and ncu captures of the final kernels
set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_20_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_20_pytest.log
tail -3 gpurun_out/r2_20_pytest.log
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
for wl in c2 mesh c5 scene8 c1 c4; do
  timeout 600 ncu --clock-control none --csv --log-file gpurun_out/r2_20_step_$wl.csv --metrics $M python bench.py --steps 1 --warmup 0 --no-extras --workload $wl > gpurun_out/r2_20_ncu_$wl.log 2>&1
done
cp gpurun_out/r2_20_step_mesh.csv gpurun_out/r2_20_step_c3.csv
# full captures: first-bounce and second-bounce launches of the traversal kernels and of the shading kernel
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -c 3 -o gpurun_out/r2_20_c2_traverse -f python bench.py --steps 1 --warmup 0 --no-extras --workload c2 > gpurun_out/r2_20_ncu_full1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade -c 2 -o gpurun_out/r2_20_c2_shade -f python bench.py --steps 1 --warmup 0 --no-extras --workload c2 > gpurun_out/r2_20_ncu_full2.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_20_launches_c2.csv python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2_20_ncu_l1.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_20_launches_mesh.csv python bench.py --steps 1 --warmup 0 --no-extras --workload mesh > gpurun_out/r2_20_ncu_l2.log 2>&1
du -sh gpurun_out; ls -la gpurun_out | tail -30
