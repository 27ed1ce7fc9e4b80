timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_15_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_15_pytest.log
tail -5 gpurun_out/r2_15_pytest.log
b() { n=$1; wl=$2; shift; shift; timeout 120 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --sustain-s 0 --workload $wl "$@" > gpurun_out/r2_15_${n}_$wl.json 2>> gpurun_out/r2_15.err; }
for wl in scene8 scene7 scene3 mesh; do
  for k in 0 64 256; do MCPT_U_SMEM_NODES=$k b sm$k $wl; done
done
b dflt scene5; b dflt c5; b dflt c2
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_15_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_15_')[1], round(d['value'],1), {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items() if v>0.05})
    except Exception as e: print(f, 'ERR', e)
P
