#!/bin/bash
# last call of the round: the whole GPU suite, the default bench line, configs[1] at full size and the ncu captures on the final tree
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/g_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/g_pytest_gpu.log
tail -3 gpurun_out/g_pytest_gpu.log
python bench.py > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; echo "bench rc $?"
python tools/config2_full.py > gpurun_out/g_config2.json 2>/dev/null; echo "config2 rc $?"
python tools/s4b_probe.py 2>&1 | tail -1
python tools/s5_probe.py 2>&1 | grep -E "connect" | tail -1
python tools/profile_target.py > gpurun_out/g_profile_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --profile-from-start off \
    --metrics lts__t_bytes.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__t_bytes.sum \
    -f -o gpurun_out/r02g_prof python tools/profile_target.py > gpurun_out/g_ncu_full.log 2>&1; echo "ncu full rc $?"
cat gpurun_out/g_profile_plain.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-visibility > gpurun_out/g_bench_short.json 2> gpurun_out/g_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02g_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-visibility > gpurun_out/g_ncu_launches.log 2>&1; echo "ncu launches rc $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/g_bench.json'))
print('value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['parity_on_bench_scene']['disagreements'], d['roofline']['kernel_ms'])
v=d['visibility']; print('vis %.3e rays/s %.3f ms e2e %.3f ms' % (v['value'], v['ms_per_pass'], v['e2e']['ms_per_pass']), v.get('parity_on_bench_scene'))
PY
