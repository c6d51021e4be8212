#!/bin/bash
# call W (1 GPU): lock-step walk stage with two/three turns in a row for many open pairs; S4a, S4b, S5, config2 with both ray orders
mkdir -p gpurun_out
python -m pytest tests/test_gpu_connect.py tests/test_gpu_bench_scene_parity.py tests/test_gpu_config2_full.py -m gpu -x -q > gpurun_out/w_pytest.log 2>&1; echo "pytest rc $?"
tail -2 gpurun_out/w_pytest.log
python tools/profile_target.py 2>&1 | grep -E "^connect"
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
for s in 0 1; do
SSPP_RAY_STRIDE=$s python tools/config2_full.py 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print(\"config2 stride env $s\", d[\"connect_no_map\"][\"device_ms\"], d[\"connect_with_map\"][\"device_ms\"])"
done
