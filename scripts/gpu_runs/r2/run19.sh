# the driver's own commands, as it runs them at round end
set -x
time (python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_19_ref.json 2> gpurun_out/r2_19_ref.err)
time (python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_19_c2.json 2> gpurun_out/r2_19_c2.err)
tail -3 gpurun_out/r2_19_c2.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_19_smoke.log 2>&1; tail -2 gpurun_out/r2_19_smoke.log
python - <<'P'
import json
d=json.load(open('gpurun_out/r2_19_c2.json')); r=json.load(open('gpurun_out/r2_19_ref.json'))
print('value', d['value'], 'e2e', d['e2e'], 'ref', r['value'], 'ratio', d['value']/r['value'], 'e2e ratio', d['e2e']['value']/r['value'])
print('same config', d['config']==r['config'], {k:(d['config'].get(k), r['config'].get(k)) for k in set(d['config'])|set(r['config']) if d['config'].get(k)!=r['config'].get(k)})
print('sustained', d['sustained'])
print('other', json.dumps(d['other_configs'], indent=1)[:3000])
print('roofline', {k:v for k,v in d['roofline'].items() if k not in ('fp32','stage_ms_per_step')})
print('cpu', d['cpu_baseline'])
P
