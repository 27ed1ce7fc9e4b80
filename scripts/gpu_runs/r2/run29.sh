set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_29_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_29_pytest.log
tail -8 gpurun_out/r2_29_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 120 python bench.py --steps 10 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_29_${n}_$wl.json 2>> gpurun_out/r2_29.err; }
for wl in c2 c1 scene1 scene2 scene4 scene6 fps1 fps6; do b pool $wl; done
b base c2 --flat-pool 0
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_29_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_29_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
