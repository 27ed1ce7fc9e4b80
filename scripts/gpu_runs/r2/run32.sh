# pooled leaf tests of the unified walk (MCPT_OPT_LEAF_POOL): parity tests, then the trigger sweep on the mesh workload, C3 and scene 8
set -x
timeout 300 python -m pytest tests/test_gpu_unified.py -m gpu -q -x > gpurun_out/r2_32_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_32_pytest.log
tail -15 gpurun_out/r2_32_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 10 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_32_${n}_$wl.json 2>> gpurun_out/r2_32.err; }
for t in 0 4 8 12 16 24 32; do b pool$t mesh --leaf-pool $t; b pool$t scene8 --leaf-pool $t; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_32_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_32_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
tail -5 gpurun_out/r2_32.err
