#!/bin/bash
# call I (1 GPU): viewpoint processing order (tests + timing), k_edges tuning switches
mkdir -p gpurun_out
python -m pytest tests/test_gpu_visibility.py tests/test_gpu_bench_scene_parity.py::test_bench_visibility_scene_against_oracle_sample tests/test_gpu_config2_full.py tests/test_gpu_peer_assembly.py -x -q 2>&1 | tail -3
python tools/profile_target.py 2>&1 | tail -2
for v in "-DSSPP_RAY_BOX=0" "-DSSPP_PAIR_COARSE=0" "-DSSPP_RAY_BOX=0 -DSSPP_PAIR_COARSE=0"; do
  SSPP_NVCC_EXTRA="$v" python -c "from sspp_b200 import build; build.build_gpu(force=True)" && echo "variant $v" && python tools/profile_target.py connect 2>&1 | tail -1
done
