#!/bin/bash
# timing only: CTAs per SM of the other edge-kernel instantiations (S5, config2) and of k_count
mkdir -p gpurun_out
cp sspp_b200/csrc/libsspp_gpu.so /tmp/lib_keep.so
for v in base all5 all6; do
cp build_variants/lib_$v.so sspp_b200/csrc/libsspp_gpu.so
echo "variant $v"
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1 | cut -c1-20
python tools/config2_full.py 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('config2', d['connect_no_map']['device_ms'], d['connect_with_map']['device_ms'])"
done
for v in c6 c8; do
cp build_variants/lib_$v.so sspp_b200/csrc/libsspp_gpu.so
echo "variant $v"
python tools/profile_target.py 2>&1 | grep -E "^connect"
done
cp /tmp/lib_keep.so sspp_b200/csrc/libsspp_gpu.so
