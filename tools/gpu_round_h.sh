#!/bin/bash
# call H (8 GPUs): chunk-count / push-grid variants of the overlapped assembler
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29644 --nproc-per-node 8"
run() { name=$1; shift; env "$@" $TR bench.py --gpus 8 --steps 10 --warmup 3 --assembler streamed --no-s5 --no-visibility > gpurun_out/h_$name.json 2> gpurun_out/h_$name.err; echo "$name rc $?"; }
run k2 SSPP_HOST_CHUNKS=2
run k3 SSPP_HOST_CHUNKS=3
run k2_room0 SSPP_HOST_CHUNKS=2 SSPP_PUSH_ROOM=0
run k2_pb32 SSPP_HOST_CHUNKS=2 SSPP_PUSH_BLOCKS=32 SSPP_PUSH_ROOM=32
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_*.json')):
    try:
        d=json.load(open(f)); print(f, 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), d['multi_gpu']['connect_device_ms_per_rank'][:3], d['multi_gpu']['connect_host_ms_per_rank'][:2])
    except Exception as e: print(f, 'ERR', e)
PY
