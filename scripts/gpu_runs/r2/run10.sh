set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_10_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_10_pytest.log
tail -5 gpurun_out/r2_10_pytest.log
timeout 300 python bench.py --steps 10 --workload mesh --no-cpu-baseline > gpurun_out/r2_10_mesh.json 2> gpurun_out/r2_10_mesh.err; tail -3 gpurun_out/r2_10_mesh.err
timeout 300 python bench.py --steps 10 --workload mesh --host-tree 1 --no-cpu-baseline > gpurun_out/r2_10_mesh_ht.json 2> gpurun_out/r2_10_mesh_ht.err
timeout 300 python bench.py --steps 10 --workload c4 --no-cpu-baseline > gpurun_out/r2_10_c4.json 2> gpurun_out/r2_10_c4.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_10_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_10_')[1], round(d['value'],1), 'e2e', d['e2e'])
    except Exception as e: print(f, 'ERR', e)
P
