# pooled flat walk: candidates ranked through match.any + REDUX instead of 64-bit atomicMin, bit-sliced ballot scan instead of the shuffle scan
set -x
timeout 300 python -m pytest tests/test_gpu_render.py tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/r2_41_pytest.log 2>&1; tail -3 gpurun_out/r2_41_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_41_${n}_$wl.json 2>> gpurun_out/r2_41.err; }
for wl in c2 c1 scene2 scene4; do
  b m1s1 $wl
  for v in m0s0 m1s0 m0s1; do MCPT_LIBRARY=montecarlo_pathtracing_b200/lib/variants/pool_$v/libmcpt.so b $v $wl; done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_41_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_41_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
