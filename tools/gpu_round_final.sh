#!/bin/bash
# last call of the round: the whole GPU suite and the default bench line on the final tree (no profiler)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/f_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/f_pytest_gpu.log
tail -3 gpurun_out/f_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/f_bench.json'))
print('value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['parity_on_bench_scene']['disagreements'], d['roofline']['kernel_ms'], d['steps'], d['warmup'])
PY
