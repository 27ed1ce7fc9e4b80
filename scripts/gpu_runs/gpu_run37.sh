set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest37.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest37.log
timeout 600 python bench.py > gpurun_out/r37_c2_default.json 2> gpurun_out/r37_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r37_c1.json 2>> gpurun_out/r37.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r37_c5.json 2>> gpurun_out/r37.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r37_mesh.json 2>> gpurun_out/r37.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r37_ref_c2.json 2>> gpurun_out/r37.err
python __graft_entry__.py --smoke > gpurun_out/r37_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r37_smoke.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r37_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r37_launches.csv $CMD > gpurun_out/r37_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 6 -o gpurun_out/r37_prof_traverse $CMD > gpurun_out/r37_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 4 -o gpurun_out/r37_prof_shade $CMD > gpurun_out/r37_ncu3.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
$M > gpurun_out/r37_mesh_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r37_prof_mesh $M > gpurun_out/r37_ncu4.log 2>&1
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
$C5 > gpurun_out/r37_c5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r37_prof_c5 $C5 > gpurun_out/r37_ncu5.log 2>&1
timeout 600 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -s -k converges > gpurun_out/r37_converge.log 2>&1
