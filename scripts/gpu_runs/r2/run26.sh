# flat scenes after the first bounce: exact tests pooled per warp and primitive type (MCPT_FLAT_POOL=1) -- parity, then A/B
set -x
MCPT_FLAT_POOL=1 timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2_26_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_26_pytest.log
tail -12 gpurun_out/r2_26_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 120 python bench.py --steps 10 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_26_${n}_$wl.json 2>> gpurun_out/r2_26.err; }
for wl in c2 c1 scene1 scene2 scene4; do
  b base $wl
  MCPT_FLAT_POOL=1 b pool $wl
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_26_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_26_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
