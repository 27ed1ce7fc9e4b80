set -x
for w in scene1 scene2 scene4 c2 c1; do
  timeout 300 python bench.py --steps 5 --no-cpu-baseline --workload $w --flat-walk 0 > gpurun_out/r76_${w}_tree.json 2>> gpurun_out/r76.err
done
