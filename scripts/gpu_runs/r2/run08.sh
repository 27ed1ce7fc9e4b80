# unified walk: first-bounce refill threshold, top-of-stack in registers
set -x
V=montecarlo_pathtracing_b200/lib/variants
b() { # name lib workload extra
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  timeout 90 python bench.py --steps 5 --no-cpu-baseline --workload $3 $4 $5 > gpurun_out/r2_08_$1_$3.json 2>> gpurun_out/r2_08.err
  unset MCPT_LIBRARY
}
b default "" mesh
for r in 16 24 32; do MCPT_REFILL_FIRST=$r b rf$r "" mesh; done
for v in topreg topreg_l2 l2; do b $v $V/$v/libmcpt.so mesh; MCPT_REFILL_FIRST=32 b ${v}_rf32 $V/$v/libmcpt.so mesh; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_08_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_08_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
