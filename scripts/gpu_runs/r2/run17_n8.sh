# 8-GPU box: strong scaling of BASELINE config 4 (3840x2160, 1024 spp split over N, one reduce, resolve on rank 0) at N = 1, 2, 4, 8,
# weak scaling of C2 at N = 8 through both reduce paths, and the C-ABI reduce check
set -x
nvidia-smi --query-gpu=index,name --format=csv | head -10
run() { # n outfile args...
  n=$1; out=$2; shift; shift
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n "$@" > gpurun_out/$out.json 2> gpurun_out/$out.err
  fi
  tail -2 gpurun_out/$out.err
}
for n in 8 4 2 1; do run $n r2_17_c4_strong_n$n --workload c4 --scaling strong --steps 3 --warmup 1 --no-cpu-baseline --sustain-s 0 --reduce capi; done
run 8 r2_17_c4_strong_torch_n8 --workload c4 --scaling strong --steps 3 --warmup 1 --no-cpu-baseline --sustain-s 0 --reduce torch
run 8 r2_17_c2_weak_n8 --steps 20 --warmup 5 --no-cpu-baseline
run 8 r2_17_c2_weak_capi_n8 --steps 20 --warmup 5 --no-cpu-baseline --reduce capi
run 8 r2_17_c2_strong_n8 --steps 20 --warmup 5 --no-cpu-baseline --scaling strong
run 1 r2_17_c2_strong_n1 --steps 5 --warmup 2 --no-cpu-baseline --scaling strong --sustain-s 0
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tests/nccl_reduce_check.py > gpurun_out/r2_17_nccl_check_n8.log 2>&1; tail -2 gpurun_out/r2_17_nccl_check_n8.log
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_17_*.json')):
    try:
        d=json.load(open(f)); print(f.split('r2_17_')[1], 'N', d['n_gpus'], d['scaling'], round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'reduce', d.get('reduce'), 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
P
