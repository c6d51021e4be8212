#!/bin/bash
python -m pytest tests/test_gpu_connect.py tests/test_gpu_queries.py tests/test_gpu_host.py tests/test_gpu_bench_scene_parity.py -x -q 2>&1 | tail -1
for i in 1 2; do python tools/profile_target.py connect 2>&1 | tail -1; done
