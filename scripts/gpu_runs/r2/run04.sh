# A/B: while-while loop variants of the persistent traversal (MCPT_FLOW=2)
set -x
V=montecarlo_pathtracing_b200/lib/variants
b() { # name lib workload extra
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  timeout 90 python bench.py --steps 5 --no-cpu-baseline --workload $3 $4 $5 > gpurun_out/r2_04_$1_$3.json 2>> gpurun_out/r2_04.err
  unset MCPT_LIBRARY
}
MCPT_LIBRARY=$V/ww/libmcpt.so timeout 200 python -m pytest tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r2_04_pytest_ww.log 2>&1; echo "exit $?" >> gpurun_out/r2_04_pytest_ww.log
tail -3 gpurun_out/r2_04_pytest_ww.log
for wl in mesh c5 scene8 scene3; do
  b default "" $wl
  for v in ww ww_m4 ww_m8 ww_m12 ww_s2; do b $v $V/$v/libmcpt.so $wl; done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_04_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_04_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
