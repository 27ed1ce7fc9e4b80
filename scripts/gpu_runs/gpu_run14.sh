set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest14.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest14.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r14_c2.json 2>> gpurun_out/r14.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r14_c1.json 2>> gpurun_out/r14.err
for pers in 1 0; do
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --persistent $pers > gpurun_out/r14_c5_p$pers.json 2>> gpurun_out/r14.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --persistent $pers > gpurun_out/r14_mesh_p$pers.json 2>> gpurun_out/r14.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ 3 > gpurun_out/r14_mesh_p1_occ3.json 2>> gpurun_out/r14.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 --occ 3 > gpurun_out/r14_c5_p1_occ3.json 2>> gpurun_out/r14.err
