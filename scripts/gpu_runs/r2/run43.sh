# experiment: time of the production traversal kernel on one batch of secondary rays in different orders
set -x
for wl in mesh scene8; do timeout 300 python scripts/exp_ray_order.py --workload $wl > gpurun_out/r2_43_order_$wl.json 2> gpurun_out/r2_43_order_$wl.err; tail -2 gpurun_out/r2_43_order_$wl.err; cat gpurun_out/r2_43_order_$wl.json; done
