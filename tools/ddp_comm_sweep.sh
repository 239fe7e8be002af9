#!/bin/bash
# tools/ddp_comm_sweep.sh N : ddp-mode bench over gradient bucket size x grid cap of the collective kernels
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
for cfg in "25 16" "25 32" "10 16" "10 32"; do
  set -- $cfg
  DPB_COMM_BLOCKS=$2 timeout 200 $T --master-port 29502 bench.py --gpus $N --steps 30 --warmup 5 --mode ddp --bucket-cap-mb $1 2>$OUT/ddp_sweep.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ddp cap=$1 MB comm blocks=$2:', round(d['ms_per_step'],4), 'ms/step', round(d['value']), 'img/s')
"
done
