# final evidence, part 2: full ncu captures of the unified traversal kernel (1 M triangles, repo scene 8)
set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -c 2 -o gpurun_out/r2_21_mesh_traverse -f python bench.py --steps 1 --warmup 0 --no-extras --workload mesh > gpurun_out/r2_21_ncu_full3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_traverse -c 2 -o gpurun_out/r2_21_scene8_traverse -f python bench.py --steps 1 --warmup 0 --no-extras --workload scene8 > gpurun_out/r2_21_ncu_full4.log 2>&1
du -sh gpurun_out; ls -la gpurun_out
# ... and the bench lines of every workload with the final kernels (profiles/r02_bench_*.json)
b() { n=$1; shift; timeout 300 python bench.py "$@" > gpurun_out/r2_21_bench_$n.json 2> gpurun_out/r2_21_bench_$n.err; tail -1 gpurun_out/r2_21_bench_$n.err; }
b c2 --steps 100 --warmup 5
b c1 --workload c1 --steps 200 --no-cpu-baseline
b c5 --workload c5 --steps 30
b mesh --workload mesh --steps 30
b c3 --workload c3 --steps 4 --warmup 1 --no-cpu-baseline
b c4 --workload c4 --steps 30 --no-cpu-baseline
b mesh_hosttree --workload mesh --steps 20 --no-cpu-baseline --host-tree 1
b mesh_twolevel --workload mesh --steps 20 --no-cpu-baseline --unified 0
b c2_nee_off_ref --impl reference --steps 20 --warmup 5
for k in 1 2 3 4 5 6 7 8; do b scene$k --workload scene$k --steps 20 --no-cpu-baseline --sustain-s 0; done
for k in 1 2 3 4 5 6 7 8; do b fps$k --workload fps$k --steps 200 --warmup 20 --no-cpu-baseline --sustain-s 0; done
du -sh gpurun_out
