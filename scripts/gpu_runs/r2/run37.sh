# k_shade with the next path's hit / radiance sectors copied in by cp.async (MCPT_SHADE_ASYNC): A/B against the direct loads, parity tests
set -x
V=montecarlo_pathtracing_b200/lib/variants/shade_sync/libmcpt.so
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --flat-pool 2 --workload $wl "$@" > gpurun_out/r2_37_${n}_$wl.json 2>> gpurun_out/r2_37.err; }
for wl in c2 c1 c5 mesh scene8; do b async $wl; MCPT_LIBRARY=$V b sync $wl; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_37_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_37_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
tail -3 gpurun_out/r2_37.err
timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py -m gpu -q -x > gpurun_out/r2_37_pytest.log 2>&1; tail -3 gpurun_out/r2_37_pytest.log
