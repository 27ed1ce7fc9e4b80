set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest87.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest87.log
timeout 600 python bench.py > gpurun_out/r87_c2_default.json 2> gpurun_out/r87_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r87_c1.json 2>> gpurun_out/r87.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r87_c5.json 2>> gpurun_out/r87.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r87_mesh.json 2>> gpurun_out/r87.err
timeout 300 python bench.py --steps 5 --workload c4 --no-cpu-baseline > gpurun_out/r87_c4.json 2>> gpurun_out/r87.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r87_ref_c2.json 2>> gpurun_out/r87.err
for k in 1 2 3 4 5 6 7 8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r87_scene$k.json 2>> gpurun_out/r87.err
  timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --workload fps$k > gpurun_out/r87_fps$k.json 2>> gpurun_out/r87.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene8 --top-sah 0 > gpurun_out/r87_scene8_nosah.json 2>> gpurun_out/r87.err
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 330 --csv --log-file gpurun_out/r87_launches.csv $CMD > gpurun_out/r87_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r87_prof_traverse $CMD > gpurun_out/r87_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 3 -o gpurun_out/r87_prof_shade $CMD > gpurun_out/r87_ncu3.log 2>&1
python __graft_entry__.py --smoke > gpurun_out/r87_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r87_smoke.log
