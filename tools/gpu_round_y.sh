#!/bin/bash
# call Y (1 GPU): k_count leaves out cells that are out of reach; variant a = 4 CTAs per SM, b = 5 (48 registers)
mkdir -p gpurun_out
for v in a b; do
cp build_variants/lib_$v.so sspp_b200/csrc/libsspp_gpu.so
echo "variant $v"
python tools/profile_target.py 2>&1 | grep -E "^connect"
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
done
cp build_variants/lib_a.so sspp_b200/csrc/libsspp_gpu.so
python -m pytest tests/test_gpu_connect.py tests/test_gpu_bench_scene_parity.py tests/test_gpu_config2_full.py tests/test_gpu_host.py tests/test_gpu_peer_assembly.py -m gpu -x -q > gpurun_out/y_pytest.log 2>&1; echo "pytest rc $?"
tail -2 gpurun_out/y_pytest.log
