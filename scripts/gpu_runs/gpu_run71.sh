set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest71.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest71.log
timeout 600 python bench.py > gpurun_out/r71_c2_default.json 2> gpurun_out/r71_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r71_c1.json 2>> gpurun_out/r71.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r71_c5.json 2>> gpurun_out/r71.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r71_mesh.json 2>> gpurun_out/r71.err
timeout 300 python bench.py --steps 5 --workload c4 --no-cpu-baseline > gpurun_out/r71_c4.json 2>> gpurun_out/r71.err
for k in 1 2 3 4 5 6 7 8; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene$k > gpurun_out/r71_scene$k.json 2>> gpurun_out/r71.err
done
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload scene8 --top-sah 0 > gpurun_out/r71_scene8_nosah.json 2>> gpurun_out/r71.err
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 2 -o gpurun_out/r71_prof_c5 $C5 > gpurun_out/r71_ncu5.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 3 -o gpurun_out/r71_prof_mesh $M > gpurun_out/r71_ncu4.log 2>&1
