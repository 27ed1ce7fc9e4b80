set -x
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/nccl_reduce_check.py > gpurun_out/r21_nccl_check.log 2>&1; echo "exit $?" >> gpurun_out/r21_nccl_check.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r21_c2_n2.json 2> gpurun_out/r21_n2.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r21_c2_n1.json 2>> gpurun_out/r21_n1.err
