#!/bin/bash
# scale line for one world size after the lock-step walk stage (no reference arm)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29655 --nproc-per-node $N"
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/z_scale_n$N.json 2> gpurun_out/z_scale_n$N.err; echo "n$N rc $?"
python - <<PY
import json
d=json.load(open('gpurun_out/z_scale_n$N.json')); print('value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['config'].get('csr_assembly'), json.dumps(d.get('multi_gpu'))[:300])
print(json.dumps(d.get('s5'))[:2600])
PY
