set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest24.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest24.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r24_c2.json 2>> gpurun_out/r24.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r24_c1.json 2>> gpurun_out/r24.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r24_c5.json 2>> gpurun_out/r24.err
for r in 0 4 12 16; do
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --refill $r > gpurun_out/r24_mesh_r$r.json 2>> gpurun_out/r24.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ 3 > gpurun_out/r24_mesh_occ3.json 2>> gpurun_out/r24.err
