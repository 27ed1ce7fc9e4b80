# step metrics of the tree / mesh workloads re-taken with the final kernels, then their bench lines
set -x
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
for wl in c5 mesh c4 scene8; do
  timeout 300 ncu --clock-control none --csv --log-file gpurun_out/r2_54_step_$wl.csv --metrics $M python bench.py --steps 1 --warmup 0 --no-extras --workload $wl > gpurun_out/r2_54_ncu_$wl.log 2>&1
done
C=$(cat gpurun_out/r2_54_commit.txt)
for wl in c5 mesh c4 scene8; do python scripts/ncu_step_metrics.py $wl gpurun_out/r2_54_step_$wl.csv $C > /dev/null; done
cp profiles/r02_step_metrics.json gpurun_out/r2_54_step_metrics.json
b() { n=$1; shift; timeout 300 python bench.py "$@" > gpurun_out/r2_54_bench_$n.json 2> gpurun_out/r2_54_bench_$n.err; tail -1 gpurun_out/r2_54_bench_$n.err; }
for wl in c5 mesh c4; do b $wl --workload $wl --steps 40 --no-cpu-baseline; done
b scene8 --workload scene8 --steps 20 --no-cpu-baseline --sustain-s 0
du -sh gpurun_out
