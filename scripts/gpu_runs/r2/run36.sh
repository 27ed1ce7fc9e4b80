# pooled flat walk with the next tile's rays copied in by cp.async (MCPT_OPT_FLAT_POOL = 2): parity, A/B
set -x
timeout 300 python -m pytest tests/test_gpu_render.py -m gpu -q -x -k pooled > gpurun_out/r2_36_pytest.log 2>&1; tail -3 gpurun_out/r2_36_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_36_${n}_$wl.json 2>> gpurun_out/r2_36.err; }
for wl in c2 c1 scene1 scene2 scene4 scene6; do b p1 $wl --flat-pool 1; b p2 $wl --flat-pool 2; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_36_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_36_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
tail -3 gpurun_out/r2_36.err
