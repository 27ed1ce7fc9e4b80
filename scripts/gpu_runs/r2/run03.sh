# round-2 baseline of the restored checkpoint: GPU tests, bench lines, launch list and one full ncu capture of the mesh traversal
set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_03_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_03_pytest.log
tail -3 gpurun_out/r2_03_pytest.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload mesh > gpurun_out/r2_03_mesh.json 2> gpurun_out/r2_03_mesh.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload scene8 > gpurun_out/r2_03_scene8.json 2> gpurun_out/r2_03_scene8.err
timeout 300 python bench.py --steps 20 --no-cpu-baseline > gpurun_out/r2_03_c2.json 2> gpurun_out/r2_03_c2.err
timeout 300 python bench.py --steps 10 --no-cpu-baseline --workload c5 > gpurun_out/r2_03_c5.json 2> gpurun_out/r2_03_c5.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_03_mesh_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload mesh > gpurun_out/r2_03_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -s 2 -c 3 -o gpurun_out/r2_03_mesh_traverse -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload mesh > gpurun_out/r2_03_ncu2.log 2>&1
ls -la gpurun_out
