# A/B of the persistent loop variants (flow pipeline, batched deliver, node steps, pops). Every command has a short timeout:
# a variant that hangs costs 90 s, not the call.
set -x
V=montecarlo_pathtracing_b200/lib/variants
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r2_02_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_02_pytest.log
tail -3 gpurun_out/r2_02_pytest.log
# the no-alloc entry point's test alone, in a fresh process
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_02_pytest_zz.log 2>&1; echo "exit $?" >> gpurun_out/r2_02_pytest_zz.log
tail -3 gpurun_out/r2_02_pytest_zz.log
b() { # name lib workload extra
  if [ -n "$2" ]; then export MCPT_LIBRARY=$2; else unset MCPT_LIBRARY; fi
  timeout 90 python bench.py --steps 5 --no-cpu-baseline --workload $3 $4 $5 > gpurun_out/r2_02_$1_$3.json 2>> gpurun_out/r2_02.err
  unset MCPT_LIBRARY
}
# parity of the batched-deliver variant first (short): hit records of every walker
MCPT_LIBRARY=$V/batch/libmcpt.so timeout 150 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -k "traversal" > gpurun_out/r2_02_pytest_batch.log 2>&1; echo "exit $?" >> gpurun_out/r2_02_pytest_batch.log
for wl in mesh c5 scene8 scene3; do
  b default "" $wl
  for v in base batch flow_n3 flow_n4 flow_p1 flow_p3; do b $v $V/$v/libmcpt.so $wl; done
done
b default_r12 "" mesh --refill 12
b default_r16 "" mesh --refill 16
b default_r4 "" mesh --refill 4
b default_r16 "" c5 --refill 16
b default_r24 "" c5 --refill 24
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_02_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_02_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
