#!/bin/bash
# round-2 GPU call A: tests, bench, launch list, one full capture of every kernel of the step
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
python -m pytest tests -m gpu -x -q -s > gpurun_out/a_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/a_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err; echo "ref rc $?"
python tools/profile_target.py > gpurun_out/a_profile_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --profile-from-start off \
    --metrics lts__t_bytes.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__t_bytes.sum \
    -f -o gpurun_out/r02a_prof python tools/profile_target.py > gpurun_out/a_ncu_full.log 2>&1; echo "ncu full rc $?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-visibility > gpurun_out/a_bench_short.json 2> gpurun_out/a_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02a_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-visibility > gpurun_out/a_ncu_launches.log 2>&1; echo "ncu launches rc $?"
tail -5 gpurun_out/a_pytest_gpu.log
