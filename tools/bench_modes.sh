#!/bin/bash
# tools/bench_modes.sh N : bench.py (our arm) in the three gradient-sync modes on N GPUs; JSON lines land in gpurun_out/.
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
for m in all_reduce gather ddp; do
  timeout 300 $T --master-port 29502 bench.py --gpus $N --steps 30 --warmup 5 --mode $m 2>$OUT/bench_ours_${N}_$m.err | tee $OUT/bench_ours_${N}_$m.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('N=$N $m:', round(d['ms_per_step'],4), 'ms/step', round(d['value']), 'img/s, e2e', round(d['e2e']['value']), 'clocks', d.get('clocks'))
"
done
