# evidence of the final kernels (fused raygen, async pooled walk): step metrics + full capture for c2 / c1, launch list, bench lines of every workload, default run
set -x
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum
for wl in c2 c1; do
  timeout 600 ncu --clock-control none --csv --log-file gpurun_out/r2_49_step_$wl.csv --metrics $M python bench.py --steps 1 --warmup 0 --no-extras --workload $wl > gpurun_out/r2_49_ncu_$wl.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -c 4 -o gpurun_out/r2_49_c2_traverse -f python bench.py --steps 1 --warmup 0 --no-extras --workload c2 > gpurun_out/r2_49_ncu_full1.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_49_launches_c2.csv python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2_49_ncu_l1.log 2>&1
b() { n=$1; shift; timeout 300 python bench.py "$@" > gpurun_out/r2_49_bench_$n.json 2> gpurun_out/r2_49_bench_$n.err; tail -1 gpurun_out/r2_49_bench_$n.err; }
b c1 --workload c1 --steps 200 --no-cpu-baseline
for k in 1 2 3 4 5 6 7 8; do b scene$k --workload scene$k --steps 20 --no-cpu-baseline --sustain-s 0; done
for k in 1 2 3 4 5 6 7 8; do b fps$k --workload fps$k --steps 200 --warmup 20 --no-cpu-baseline --sustain-s 0; done
du -sh gpurun_out
timeout 900 python bench.py --steps 100 > gpurun_out/r2_49_bench_default.json 2> gpurun_out/r2_49_bench_default.err; tail -2 gpurun_out/r2_49_bench_default.err
for wl in c5 mesh c4; do b $wl --workload $wl --steps 40 --no-cpu-baseline; done
b c3 --workload c3 --steps 6 --warmup 2 --no-cpu-baseline
du -sh gpurun_out
