# the driver's scaling command at N = 2 (default workload, other_configs.c4_strong under torchrun), both arms
set -x
time (python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_51_ref_n2.json 2> gpurun_out/r2_51_ref_n2.err)
time (python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_51_c2_n2.json 2> gpurun_out/r2_51_c2_n2.err)
tail -5 gpurun_out/r2_51_c2_n2.err
python - <<'P'
import json
d=json.load(open('gpurun_out/r2_51_c2_n2.json')); r=json.load(open('gpurun_out/r2_51_ref_n2.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'reduce', d['reduce'], 'same config', d['config']==r['config'])
print('other', {k:(round(v['value'],1), round(v['ms_per_step'],2), v['n_gpus'], v['scaling']) for k,v in (d['other_configs'] or {}).items() if v})
print('ref', r['value'], r['n_gpus'])
P
