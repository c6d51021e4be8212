#!/bin/bash
# round-2 GPU call B: k_edges v2 (batch key box scan + projections): tests, bench, full capture, occupancy variant
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/b_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/b_pytest_gpu.log
tail -4 gpurun_out/b_pytest_gpu.log
python bench.py --steps 10 --warmup 3 --no-visibility > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc $?"
python tools/profile_target.py connect > gpurun_out/b_profile_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --profile-from-start off \
    --metrics lts__t_bytes.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__t_bytes.sum \
    -f -o gpurun_out/r02b_prof python tools/profile_target.py connect > gpurun_out/b_ncu_full.log 2>&1; echo "ncu full rc $?"
cat gpurun_out/b_profile_plain.log
for v in 3 5; do
  SSPP_NVCC_EXTRA="-DEDGE_MIN_BLOCKS=$v" python -c "from sspp_b200 import build; build.build_gpu(force=True)" && \
  python tools/profile_target.py connect > gpurun_out/b_profile_minblocks$v.log 2>&1; cat gpurun_out/b_profile_minblocks$v.log
done
