#!/bin/bash
# call R (2 GPUs): fused exchange: tests + N=2 bench, fused against peer
mkdir -p gpurun_out
python -m pytest tests/test_gpu_peer_assembly.py tests/test_gpu_connect.py -x -q 2>&1 | tail -3
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29666"
for a in fused peer; do
  $TR bench.py --gpus 2 --steps 10 --warmup 3 --assembler $a --no-s5 --no-visibility > gpurun_out/r_n2_$a.json 2> gpurun_out/r_n2_$a.err; echo "n2 $a rc $?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r_n2_*.json')):
    try:
        d=json.load(open(f)); print(f, 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), d['config'].get('csr_assembly'), d['multi_gpu']['connect_device_ms_per_rank'], d['multi_gpu']['connect_host_ms_per_rank'])
    except Exception as e: print(f, 'ERR', e); print(open(f.replace('.json','.err')).read()[-1500:])
PY
