# unified world-space tree: GPU tests, then A/B against the two-level walk
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_05_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_05_pytest.log
tail -15 gpurun_out/r2_05_pytest.log
b() { # name workload extra...
  n=$1; wl=$2; shift; shift
  timeout 120 python bench.py --steps 5 --no-cpu-baseline --workload $wl "$@" > gpurun_out/r2_05_${n}_$wl.json 2>> gpurun_out/r2_05.err
}
b u0 mesh --unified 0
b u1 mesh
b u1o5 mesh --occ 5
b u1o6 mesh --occ 6
b u1o3 mesh --occ 3
b u1r12 mesh --refill 12
b u1r16 mesh --refill 16
b u1r4 mesh --refill 4
b u1 c4
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_05_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_05_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05}, d['roofline']['fp32']['work_per_step'])
    except Exception as e: print(f, 'ERR', e)
P
tail -5 gpurun_out/r2_05.err
