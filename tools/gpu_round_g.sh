#!/bin/bash
# call G (8 GPUs): the default N=8 line with the S5 object, the overlapped assembler for comparison, N=4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29633"
$TR --nproc-per-node 8 bench.py --gpus 8 --steps 10 --warmup 3 --assembler streamed --no-s5 --no-visibility > gpurun_out/g_n8_streamed.json 2> gpurun_out/g_n8_streamed.err; echo "n8 streamed rc $?"
$TR --nproc-per-node 8 bench.py --gpus 8 --steps 10 --warmup 3 --assembler peer --no-s5 --no-visibility > gpurun_out/g_n8_peer.json 2> gpurun_out/g_n8_peer.err; echo "n8 peer rc $?"
$TR --nproc-per-node 4 bench.py --gpus 4 --steps 10 --warmup 3 --no-s5 --no-visibility > gpurun_out/g_n4.json 2> gpurun_out/g_n4.err; echo "n4 rc $?"
$TR --nproc-per-node 8 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/g_n8_default.json 2> gpurun_out/g_n8_default.err; echo "n8 default rc $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/g_n*.json')):
    try:
        d=json.load(open(f)); print(f, 'value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['config'].get('csr_assembly'), json.dumps(d.get('multi_gpu'))[:420])
        if d.get('s5'): print(json.dumps(d['s5'])[:3500])
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/g_n8_default.err
