# per-launch counters of one step (light ncu pass) for the step-metrics JSON, mesh + scene8
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
for wl in mesh scene8; do
  timeout 600 ncu --clock-control none --csv --log-file gpurun_out/r2_13_step_$wl.csv --metrics $M python bench.py --steps 1 --warmup 0 --no-extras --workload $wl > gpurun_out/r2_13_$wl.log 2>&1
done
ls -la gpurun_out/*.csv
