# RNG seeds of k_shade from the per-pixel screen_tc table: parity (RNG bit-exact test, golden images, render tests), bench lines
set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_53_pytest.log 2>&1; tail -3 gpurun_out/r2_53_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_53_${n}_$wl.json 2>> gpurun_out/r2_53.err; }
for wl in c2 c5 scene1 mesh; do b table $wl; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_53_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_53_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
