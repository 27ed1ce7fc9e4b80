set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_09_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_09_pytest.log
tail -30 gpurun_out/r2_09_pytest.log
timeout 300 python bench.py --steps 10 --workload mesh > gpurun_out/r2_09_mesh.json 2> gpurun_out/r2_09_mesh.err; tail -3 gpurun_out/r2_09_mesh.err
timeout 300 python bench.py --steps 10 --workload mesh --host-tree 1 --no-cpu-baseline > gpurun_out/r2_09_mesh_ht.json 2> gpurun_out/r2_09_mesh_ht.err; tail -3 gpurun_out/r2_09_mesh_ht.err
timeout 300 python bench.py --steps 20 > gpurun_out/r2_09_c2.json 2> gpurun_out/r2_09_c2.err; tail -3 gpurun_out/r2_09_c2.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_09_ref.json 2> gpurun_out/r2_09_ref.err; tail -3 gpurun_out/r2_09_ref.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_09_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_09_')[1], round(d['value'],1), 'e2e', d['e2e'], d.get('sustained'), d.get('roofline',{}).get('stage_ms_per_step'))
    except Exception as e: print(f, 'ERR', e)
P
