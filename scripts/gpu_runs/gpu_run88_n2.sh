set -x
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r88_c2_n2.json 2> gpurun_out/r88_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/nccl_reduce_check.py > gpurun_out/r88_nccl_check.log 2>&1; echo "exit $?" >> gpurun_out/r88_nccl_check.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --workload c4 > gpurun_out/r88_c4_n2.json 2>> gpurun_out/r88_n2.err
