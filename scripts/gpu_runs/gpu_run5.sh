set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest5.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest5.log
for ppb in 0 8 16 32; do
python bench.py --steps 10 --no-cpu-baseline --ppb $ppb > gpurun_out/r5_c2_ppb$ppb.json 2>> gpurun_out/r5.err
done
python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r5_c5.json 2>> gpurun_out/r5.err
python bench.py --steps 10 --no-cpu-baseline --workload c1 > gpurun_out/r5_c1.json 2>> gpurun_out/r5.err
python bench.py --steps 3 --no-cpu-baseline --workload mesh > gpurun_out/r5_mesh.json 2>> gpurun_out/r5.err
