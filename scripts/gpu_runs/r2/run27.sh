# pooled flat walk: rays per lane (pool size per warp) 1 / 2 / 3 / 4
V=montecarlo_pathtracing_b200/lib/variants
b() { # name lib workload
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  MCPT_FLAT_POOL=1 timeout 90 python bench.py --steps 10 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $3 > gpurun_out/r2_27_$1_$3.json 2>> gpurun_out/r2_27.err
  unset MCPT_LIBRARY
}
for wl in c2 c1 scene2 scene4; do
  b r2 "" $wl
  for v in r1 r3 r4; do b $v $V/$v/libmcpt.so $wl; done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_27_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_27_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
