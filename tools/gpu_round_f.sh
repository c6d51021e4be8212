#!/bin/bash
# call F: gather variants (visibility only)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_visibility.py tests/test_gpu_bench_scene_parity.py::test_bench_visibility_scene_against_oracle_sample -x -q 2>&1 | tail -3
python tools/profile_target.py vis 2>&1 | tail -1
for v in 6 8; do
  SSPP_NVCC_EXTRA="-DGATHER_MIN_BLOCKS=$v" python -c "from sspp_b200 import build; build.build_gpu(force=True)" && echo "GATHER_MIN_BLOCKS=$v" && python tools/profile_target.py vis 2>&1 | tail -1
done
