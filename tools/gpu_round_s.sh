#!/bin/bash
# call S (8 GPUs): fused exchange at 8 GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29677"
$TR bench.py --gpus 8 --steps 10 --warmup 3 --assembler fused --no-s5 --no-visibility > gpurun_out/s_n8_fused.json 2> gpurun_out/s_n8_fused.err; echo "n8 fused rc $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/s_n8_fused.json')); print('value %.3e ms %.3f' % (d['value'], d['ms_per_step']), d['config'].get('csr_assembly'), d['multi_gpu']['connect_device_ms_per_rank'], d['multi_gpu']['connect_host_ms_per_rank'])
PY
