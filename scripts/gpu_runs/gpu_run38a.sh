set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest38.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest38.log
timeout 600 python bench.py > gpurun_out/r38_c2_default.json 2> gpurun_out/r38_c2_default.err
timeout 300 python bench.py --steps 10 --workload c1 > gpurun_out/r38_c1.json 2>> gpurun_out/r38.err
timeout 300 python bench.py --steps 5 --workload c5 > gpurun_out/r38_c5.json 2>> gpurun_out/r38.err
timeout 300 python bench.py --steps 5 --workload mesh > gpurun_out/r38_mesh.json 2>> gpurun_out/r38.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r38_ref_c2.json 2>> gpurun_out/r38.err
python __graft_entry__.py --smoke > gpurun_out/r38_smoke.log 2>&1; echo "exit $?" >> gpurun_out/r38_smoke.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r38_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r38_launches.csv $CMD > gpurun_out/r38_ncu1.log 2>&1
timeout 600 python -m pytest tests/test_gpu_golden.py -m gpu -x -q -s -k converges > gpurun_out/r38_converge.log 2>&1
