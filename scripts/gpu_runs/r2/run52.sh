# after the per-device shared-memory attribute helper: whole GPU suite, smoke(), a short default-workload line
set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2_52_pytest.log 2>&1; tail -3 gpurun_out/r2_52_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_52_smoke.log 2>&1; tail -1 gpurun_out/r2_52_smoke.log
timeout 300 python bench.py --steps 20 --no-cpu-baseline --sustain-s 0 --other-configs 0 > gpurun_out/r2_52_bench_c2.json 2> gpurun_out/r2_52_bench_c2.err; python -c "
import json; d=json.load(open('gpurun_out/r2_52_bench_c2.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'])"
