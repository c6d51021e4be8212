#!/bin/bash
python -m pytest tests/test_gpu_visibility.py tests/test_gpu_bench_scene_parity.py::test_bench_visibility_scene_against_oracle_sample tests/test_gpu_config2_full.py tests/test_gpu_peer_assembly.py tests/test_gpu_host.py -x -q 2>&1 | tail -2
python tools/profile_target.py 2>&1 | tail -2
