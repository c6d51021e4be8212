#!/bin/bash
# round-2 GPU call E (1 GPU): visibility fast path + frustum row culling; full tests, bench, config 2 at full size, full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/e_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/e_pytest_gpu.log
tail -4 gpurun_out/e_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench rc $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/e_bench_ref.json 2> gpurun_out/e_bench_ref.err; echo "ref rc $?"
python tools/config2_full.py > gpurun_out/e_config2.json 2> gpurun_out/e_config2.err; echo "config2 rc $?"
python tools/profile_target.py > gpurun_out/e_profile_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --profile-from-start off \
    --metrics lts__t_bytes.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__t_bytes.sum \
    -f -o gpurun_out/r02e_prof python tools/profile_target.py > gpurun_out/e_ncu_full.log 2>&1; echo "ncu full rc $?"
cat gpurun_out/e_profile_plain.log
python - <<'PY'
import json
d=json.load(open('gpurun_out/e_bench.json'))
print('value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['parity_on_bench_scene'])
v=d['visibility']; print('vis %.3e rays/s %.3f ms e2e %.3f ms' % (v['value'], v['ms_per_pass'], v['e2e']['ms_per_pass']), v['phase_ms'], v.get('parity_on_bench_scene'))
print(open('gpurun_out/e_config2.json').read()[:2500])
PY
