#!/bin/bash
# tools/ddp_bucket_sweep.sh N : correctness of every sync mode on N GPUs, then bench.py in ddp mode over a few gradient
# bucket sizes (smaller buckets = less of the last all-reduce exposed after backward) and the other two modes.
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version'; }
echo "== multi_gpu_check N=$N"
timeout 600 $T --master-port 29500 tools/multi_gpu_check.py 2>&1 | filter | tail -3 | cut -c1-400 | tee $OUT/multi_check_$N.log
show() { python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', d.get('ms_per_step'), 'ms/step', d.get('value'), d.get('unit'), 'e2e', d.get('e2e',{}).get('value'))
"; }
for cap in 25 8 2 0.5; do
  timeout 300 $T --master-port 29502 bench.py --gpus $N --steps 40 --warmup 5 --mode ddp --bucket-cap-mb $cap 2>$OUT/bench_ddp_${N}_cap$cap.err | tee $OUT/bench_ddp_${N}_cap$cap.json | show "ddp cap=$cap"
done
for m in all_reduce gather; do
  timeout 300 $T --master-port 29502 bench.py --gpus $N --steps 40 --warmup 5 --mode $m 2>$OUT/bench_ours_${N}_$m.err | tee $OUT/bench_ours_${N}_$m.json | show "$m"
done
