#!/bin/bash
# Multi-GPU measurement pass on one box:  tools/run_multi.sh N [quick]
#   1. correctness of every sync mode on N real GPUs (tools/multi_gpu_check.py)
#   2. gradient-sync sweep vs NCCL (benchmarks/sync_sweep.py)
#   3. bench.py, ours and the unmodified reference, for the all_reduce / gather / ddp modes
# Everything lands in gpurun_out/ as JSON.
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version'; }

echo "== multi_gpu_check N=$N"
timeout 600 $T --master-port 29500 tools/multi_gpu_check.py 2>&1 | filter | tail -3 | cut -c1-400 | tee $OUT/multi_check_$N.log

echo "== sync sweep N=$N"
timeout 600 $T --master-port 29501 benchmarks/sync_sweep.py --out $OUT/sync_sweep_$N.json 2>&1 | filter | tail -16
if [ "$2" != "quick" ]; then
  for b in 32 128; do
    echo "-- whole arena with comm grid cap $b"
    DPB_COMM_BLOCKS=$b timeout 300 $T --master-port 29504 benchmarks/sync_sweep.py --out $OUT/sync_sweep_${N}_blocks$b.json 2>&1 | filter | grep -E "whole_arena|64MB|16MB"
  done
fi

echo "== bench N=$N"
for m in all_reduce gather ddp; do
  timeout 300 $T --master-port 29502 bench.py --gpus $N --steps 30 --warmup 5 --mode $m 2>$OUT/bench_ours_${N}_$m.err | tee $OUT/bench_ours_${N}_$m.json | cut -c1-230
  timeout 400 $T --master-port 29503 bench.py --gpus $N --steps 30 --warmup 5 --mode $m --impl reference 2>$OUT/bench_ref_${N}_$m.err | tee $OUT/bench_ref_${N}_$m.json | cut -c1-230
done
