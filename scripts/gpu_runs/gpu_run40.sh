set -x
V=gpurun_variants
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest40.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest40.log
timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r40_c2_pf.json 2>> gpurun_out/r40.err
MCPT_LIBRARY=$PWD/$V/libmcpt_nopf.so timeout 300 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/r40_c2_nopf.json 2>> gpurun_out/r40.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload c5 > gpurun_out/r40_c5_pf.json 2>> gpurun_out/r40.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh > gpurun_out/r40_mesh_pf.json 2>> gpurun_out/r40.err
