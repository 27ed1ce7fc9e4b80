V=montecarlo_pathtracing_b200/lib/variants
b() { n=$1; wl=$2; shift; shift; timeout 120 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --sustain-s 0 --workload $wl "$@" > gpurun_out/r2_14_${n}_$wl.json 2>> gpurun_out/r2_14.err; }
for wl in scene8 scene7 scene3; do
  for r in 4 8 16; do b base_r$r $wl --refill $r; done
  for r in 4 8 12 16; do MCPT_UNIFIED_ANALYTIC=1 b ua_r$r $wl --refill $r; done
  for r in 8; do MCPT_LIBRARY=$V/ul1/libmcpt.so MCPT_UNIFIED_ANALYTIC=1 b ua1_r$r $wl --refill $r; done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_14_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_14_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
