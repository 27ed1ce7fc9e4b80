set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest63.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest63.log
timeout 600 python bench.py > gpurun_out/r63_c2_default.json 2> gpurun_out/r63_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r63_c1.json 2>> gpurun_out/r63.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r63_c5.json 2>> gpurun_out/r63.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r63_mesh.json 2>> gpurun_out/r63.err
timeout 300 python bench.py --steps 5 --workload c4 --no-cpu-baseline > gpurun_out/r63_c4.json 2>> gpurun_out/r63.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r63_ref_c2.json 2>> gpurun_out/r63.err
python __graft_entry__.py --smoke > gpurun_out/r63_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r63_smoke.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r63_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 330 --csv --log-file gpurun_out/r63_launches.csv $CMD > gpurun_out/r63_ncu1.log 2>&1
timeout 600 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -s -k converges > gpurun_out/r63_converge.log 2>&1
timeout 300 python -m montecarlo_pathtracing_b200.render --scene 6 --width 1920 --height 1080 --spp 256 --bounces 8 --ior 1.5 --roughness 0.5 --out gpurun_out/r63_scene6_c2.png > gpurun_out/r63_render.json 2>> gpurun_out/r63.err
