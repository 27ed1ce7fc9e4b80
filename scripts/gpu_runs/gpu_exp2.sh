set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest2.log
for occ in 2 3 4; do for fm in 0 1; do
python bench.py --steps 10 --no-cpu-baseline --occ $occ --fast-math $fm > gpurun_out/exp2_c2_occ${occ}_fm${fm}.json 2>> gpurun_out/exp2.err
done; done
python bench.py --steps 5 --no-cpu-baseline --workload c5 --occ 3 --fast-math 1 > gpurun_out/exp2_c5_occ3_fm1.json 2>> gpurun_out/exp2.err
python bench.py --steps 5 --no-cpu-baseline --workload c5 --occ 4 --fast-math 1 > gpurun_out/exp2_c5_occ4_fm1.json 2>> gpurun_out/exp2.err
python bench.py --steps 3 --no-cpu-baseline --workload mesh --occ 3 > gpurun_out/exp2_mesh_occ3.json 2>> gpurun_out/exp2.err
python bench.py --steps 10 --no-cpu-baseline --workload c1 --occ 3 --fast-math 1 > gpurun_out/exp2_c1_occ3_fm1.json 2>> gpurun_out/exp2.err
