#!/bin/bash
# tools/bench_ref_modes.sh N : bench.py --impl reference (the unmodified reference code) in the three sync modes on N GPUs.
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
for m in all_reduce gather ddp; do
  timeout 300 $T --master-port 29503 bench.py --gpus $N --steps 30 --warmup 5 --mode $m --impl reference 2>$OUT/bench_ref_${N}_$m.err | tee $OUT/bench_ref_${N}_$m.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ref N=$N $m:', d.get('ms_per_step'), 'ms/step', d.get('value'), d.get('unavailable'))
"
done
