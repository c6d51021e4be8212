#!/bin/bash
python -m pytest tests/test_gpu_connect.py tests/test_gpu_dedup.py tests/test_gpu_peer_assembly.py tests/test_gpu_bench_scene_parity.py tests/test_gpu_host.py -x -q 2>&1 | tail -2
python tools/profile_target.py connect 2>&1 | tail -1
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
