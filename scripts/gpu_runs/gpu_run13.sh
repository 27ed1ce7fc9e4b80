set -x
timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest13.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest13.log
for occ in 4 3 2; do
timeout 300 python bench.py --steps 10 --no-cpu-baseline --occ $occ > gpurun_out/r13_c2_occ$occ.json 2>> gpurun_out/r13.err
done
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 --occ 3 > gpurun_out/r13_c1_occ3.json 2>> gpurun_out/r13.err
