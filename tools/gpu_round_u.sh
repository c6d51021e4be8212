#!/bin/bash
# call U (1 GPU): k_edges walk stage in ray-major turns with early exit; parity tests first, then timings with both ray orders
mkdir -p gpurun_out
python -m pytest tests/test_gpu_connect.py tests/test_gpu_bench_scene_parity.py tests/test_gpu_host.py -m gpu -x -q > gpurun_out/u_pytest.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/u_pytest.log
python tools/profile_target.py 2>&1 | grep -E "^connect"
SSPP_RAY_STRIDE=1 python tools/profile_target.py 2>&1 | grep -E "^connect" | sed 's/^/stride1 /'
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
python tools/config2_full.py 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print(\"config2\", d[\"connect_no_map\"][\"device_ms\"], d[\"connect_with_map\"][\"device_ms\"])"
python bench.py --steps 10 --warmup 3 --no-visibility > gpurun_out/u_bench.json 2> gpurun_out/u_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/u_bench.json'))
print('value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['parity_on_bench_scene']['disagreements'], d['roofline']['kernel_ms'])
PY
