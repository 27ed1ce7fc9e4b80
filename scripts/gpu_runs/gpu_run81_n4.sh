set -x
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r81_c2_n4.json 2> gpurun_out/r81_n4.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu-baseline --workload c4 > gpurun_out/r81_c4_n4.json 2>> gpurun_out/r81_n4.err
