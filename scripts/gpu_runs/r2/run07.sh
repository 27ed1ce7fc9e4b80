# unified walk: loop variants, then launch list + full ncu capture of the default
set -x
V=montecarlo_pathtracing_b200/lib/variants
b() { # name lib workload extra
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  timeout 90 python bench.py --steps 5 --no-cpu-baseline --workload $3 $4 $5 > gpurun_out/r2_07_$1_$3.json 2>> gpurun_out/r2_07.err
  unset MCPT_LIBRARY
}
b default "" mesh
for v in ul2_m8 ul2_m12 ul2_m16 ul2_m20 un1 un3 up1 up3; do b $v $V/$v/libmcpt.so mesh; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_07_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_07_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_07_mesh_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload mesh > gpurun_out/r2_07_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 2 -c 3 -o gpurun_out/r2_07_mesh_traverse -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload mesh > gpurun_out/r2_07_ncu2.log 2>&1
