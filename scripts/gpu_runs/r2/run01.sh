set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_01_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_01_pytest.log
tail -5 gpurun_out/r2_01_pytest.log
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r2_01_mesh.json 2> gpurun_out/r2_01_mesh.err
MCPT_DUMP_SPANS=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mesh > /dev/null 2> gpurun_out/r2_01_mesh_spans.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload scene8 > gpurun_out/r2_01_scene8.json 2> gpurun_out/r2_01_scene8.err
timeout 300 python bench.py --steps 20 --no-cpu-baseline > gpurun_out/r2_01_c2.json 2> gpurun_out/r2_01_c2.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_01_smoke.log 2>&1
