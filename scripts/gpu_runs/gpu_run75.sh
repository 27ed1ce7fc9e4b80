set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest75.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest75.log
timeout 600 python bench.py > gpurun_out/r75_c2_default.json 2> gpurun_out/r75_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r75_c1.json 2>> gpurun_out/r75.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r75_c5.json 2>> gpurun_out/r75.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r75_mesh.json 2>> gpurun_out/r75.err
timeout 300 python bench.py --steps 5 --workload c4 --no-cpu-baseline > gpurun_out/r75_c4.json 2>> gpurun_out/r75.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r75_ref_c2.json 2>> gpurun_out/r75.err
for k in 1 2 3 4 5 6 7 8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r75_scene$k.json 2>> gpurun_out/r75.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene8 --top-sah 0 > gpurun_out/r75_scene8_nosah.json 2>> gpurun_out/r75.err
python __graft_entry__.py --smoke > gpurun_out/r75_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r75_smoke.log
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 2 -o gpurun_out/r75_prof_c5 $C5 > gpurun_out/r75_ncu5.log 2>&1
