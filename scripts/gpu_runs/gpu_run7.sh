set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest7.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest7.log
for sort in 0 1; do
python bench.py --steps 10 --no-cpu-baseline --sort $sort > gpurun_out/r7_c2_sort$sort.json 2>> gpurun_out/r7.err
python bench.py --steps 10 --no-cpu-baseline --sort $sort --workload c1 > gpurun_out/r7_c1_sort$sort.json 2>> gpurun_out/r7.err
done
python bench.py --steps 10 --no-cpu-baseline --fast-math 0 > gpurun_out/r7_c2_fm0.json 2>> gpurun_out/r7.err
