set -x
nvidia-smi -L | head -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r89_c2_n8.json 2> gpurun_out/r89_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --workload c4 > gpurun_out/r89_c4_n8.json 2>> gpurun_out/r89_n8.err
