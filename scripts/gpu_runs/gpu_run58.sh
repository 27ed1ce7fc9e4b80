set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest58.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest58.log
timeout 600 python bench.py > gpurun_out/r58_c2_default.json 2> gpurun_out/r58_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r58_c1.json 2>> gpurun_out/r58.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r58_c5.json 2>> gpurun_out/r58.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r58_mesh.json 2>> gpurun_out/r58.err
timeout 300 python bench.py --steps 5 --workload c4 --no-cpu-baseline > gpurun_out/r58_c4.json 2>> gpurun_out/r58.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r58_ref_c2.json 2>> gpurun_out/r58.err
python __graft_entry__.py --smoke > gpurun_out/r58_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r58_smoke.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r58_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r58_launches.csv $CMD > gpurun_out/r58_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 4 -o gpurun_out/r58_prof_traverse $CMD > gpurun_out/r58_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 0 -c 3 -o gpurun_out/r58_prof_shade $CMD > gpurun_out/r58_ncu3.log 2>&1
M="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 3 -o gpurun_out/r58_prof_mesh $M > gpurun_out/r58_ncu4.log 2>&1
C5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5"
ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 0 -c 2 -o gpurun_out/r58_prof_c5 $C5 > gpurun_out/r58_ncu5.log 2>&1
timeout 600 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -s -k converges > gpurun_out/r58_converge.log 2>&1
timeout 300 python -m montecarlo_pathtracing_b200.render --scene 6 --width 1920 --height 1080 --spp 64 --bounces 8 --ior 1.5 --roughness 0.5 --out gpurun_out/r58_scene6_c2.png > gpurun_out/r58_render.json 2>> gpurun_out/r58.err
ls -la gpurun_out | tail -30
