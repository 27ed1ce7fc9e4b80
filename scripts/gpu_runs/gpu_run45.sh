set -x
for o in 2 3; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --occ $o > gpurun_out/r45_mesh_occ$o.json 2>> gpurun_out/r45.err
done
for r in 4 12 16 24; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload mesh --refill $r > gpurun_out/r45_mesh_refill$r.json 2>> gpurun_out/r45.err
done
