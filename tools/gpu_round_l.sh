#!/bin/bash
# call L (1 GPU): start-voxel stage for lattices up to 128 rays: connect tests + parity on the big scenes + probes
python -m pytest tests/test_gpu_connect.py tests/test_gpu_bench_scene_parity.py tests/test_gpu_config2_full.py tests/test_gpu_queries.py -x -q 2>&1 | tail -3
python tools/profile_target.py connect 2>&1 | tail -1
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
python tools/config2_full.py 2>/dev/null | cut -c1-700
