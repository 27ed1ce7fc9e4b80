# the unified walker (four-children nodes, lean state) on trees of analytic primitives
b() { n=$1; wl=$2; shift; shift; timeout 120 python bench.py --steps 5 --no-cpu-baseline --sustain-s 0 --workload $wl "$@" > gpurun_out/r2_12_${n}_$wl.json 2>> gpurun_out/r2_12.err; }
for wl in scene8 scene3 scene5 scene7 c5; do
  b base $wl
  MCPT_UNIFIED_ANALYTIC=1 b ua $wl
  MCPT_UNIFIED_ANALYTIC=1 b ua_r8 $wl --refill 8
done
MCPT_UNIFIED_ANALYTIC=1 timeout 300 python -m pytest tests/test_gpu_golden.py tests/test_gpu_traversal.py -m gpu -q -x 2>&1 | tail -3
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_12_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_12_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05}, d['roofline']['fp32']['work_per_step']['boxes'])
    except Exception as e: print(f, 'ERR', e)
P
tail -3 gpurun_out/r2_12.err
