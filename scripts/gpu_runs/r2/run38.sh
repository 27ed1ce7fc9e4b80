# async ray copies of the pooled flat walk: .cg against .ca, then the whole GPU suite with the new default
set -x
V=montecarlo_pathtracing_b200/lib/variants/cp_ca/libmcpt.so
b() { n=$1; wl=$2; shift; shift; timeout 200 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 --workload $wl "$@" > gpurun_out/r2_38_${n}_$wl.json 2>> gpurun_out/r2_38.err; }
for wl in c2 scene1 scene6; do b cg $wl; MCPT_LIBRARY=$V b ca $wl; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_38_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_38_')[1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_38_pytest.log 2>&1; tail -3 gpurun_out/r2_38_pytest.log
