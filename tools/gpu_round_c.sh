#!/bin/bash
# round-2 GPU call C (2 GPUs): tests incl. the two-process peer tests, N=2 bench with both assemblers, S5 leg
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/c_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/c_pytest_gpu.log
tail -4 gpurun_out/c_pytest_gpu.log
python bench.py --steps 10 --warmup 3 --no-visibility --no-cpu-baseline > gpurun_out/c_n1.json 2> gpurun_out/c_n1.err; echo "n1 rc $?"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
$TR bench.py --gpus 2 --steps 10 --warmup 3 --assembler peer --no-s5 --no-visibility > gpurun_out/c_n2_peer.json 2> gpurun_out/c_n2_peer.err; echo "n2 peer rc $?"
$TR bench.py --gpus 2 --steps 10 --warmup 3 --assembler streamed --no-s5 --no-visibility > gpurun_out/c_n2_streamed.json 2> gpurun_out/c_n2_streamed.err; echo "n2 streamed rc $?"
SSPP_PUSH_ROOM=0 $TR bench.py --gpus 2 --steps 10 --warmup 3 --assembler streamed --no-s5 --no-visibility > gpurun_out/c_n2_streamed_room0.json 2> gpurun_out/c_n2_streamed_room0.err; echo "n2 streamed room0 rc $?"
$TR bench.py --gpus 2 --steps 5 --warmup 3 --no-visibility > gpurun_out/c_n2_s5.json 2> gpurun_out/c_n2_s5.err; echo "n2 s5 rc $?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/c_n*.json')):
    try:
        d=json.load(open(f)); print(f, 'value %.3e ms %.3f e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['config'].get('csr_assembly'), json.dumps(d.get('multi_gpu'))[:300])
        if d.get('s5'): print(json.dumps(d['s5'])[:3000])
    except Exception as e: print(f, 'ERR', e)
PY
