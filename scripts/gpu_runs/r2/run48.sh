# fused raygen: no primary ray is stored, k_accumulate takes the direction of never-shaded paths from the camera table: parity tests, A/B
set -x
timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py tests/test_gpu_nee.py -m gpu -q -x > gpurun_out/r2_48_pytest.log 2>&1; tail -3 gpurun_out/r2_48_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_48_${n}_$wl.json 2>> gpurun_out/r2_48.err; }
for wl in c2 c5 scene6; do b fused $wl --fused-raygen 1; b plain $wl --fused-raygen 0; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_48_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_48_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
