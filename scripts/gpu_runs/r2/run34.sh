# pooled unified walk with pop passes in front of every node step: value = trigger + 64 * passes
set -x
timeout 300 python -m pytest tests/test_gpu_unified.py -m gpu -q -x -k pooled > gpurun_out/r2_34_pytest.log 2>&1; tail -3 gpurun_out/r2_34_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 10 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_34_${n}_$wl.json 2>> gpurun_out/r2_34.err; }
for v in 88 152 216 280 144 208 160 224; do b v$v mesh --leaf-pool $v; b v$v scene8 --leaf-pool $v; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_34_*.json')):
    try:
        d=json.load(open(f)); w=d['roofline']['fp32']['work_per_step']; print(f.split('r2_34_')[1], round(d['value'],1), 'boxes/ray', round(w['boxes']/w['rays'],2), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
tail -3 gpurun_out/r2_34.err
