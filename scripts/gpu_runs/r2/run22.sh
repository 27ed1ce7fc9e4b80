# unified walk: triangle slots around L1, prefetch of the reached leaf / of the second nearest child
V=montecarlo_pathtracing_b200/lib/variants
b() { # name lib workload
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  timeout 90 python bench.py --steps 5 --no-cpu-baseline --sustain-s 0 --workload $3 > gpurun_out/r2_22_$1_$3.json 2>> gpurun_out/r2_22.err
  unset MCPT_LIBRARY
}
for wl in mesh scene8; do
  b default "" $wl
  for v in trinol1 pf1 pf2 pf3 pf1_nol1; do b $v $V/$v/libmcpt.so $wl; done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_22_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_22_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
