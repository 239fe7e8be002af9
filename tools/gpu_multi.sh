#!/bin/bash
# Multi-GPU measurement pass on one box:  tools/gpu_multi.sh N [full|quick]
#   1. correctness of every sync mode against the NCCL + torch.optim oracle (tools/multi_gpu_check.py)
#   2. gradient-sync sweep vs NCCL incl. the NVLS arm (benchmarks/sync_sweep.py)
#   3. bench.py in every mode / variant, ours and the unmodified reference
#   4. timeline of one graph replay (tools/overlap_timeline.py)
#   5. (full) the flag-launched entry scripts through tests/test_gpu_cli.py
# Everything lands in gpurun_out/.
N=${1:-2}
WHAT=${2:-full}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version\|^W[0-9]'; }
PORT=29500
next() { PORT=$((PORT+1)); }

echo "== multi_gpu_check N=$N"
CHECK_OUT=$OUT/multi_gpu_check_${N}gpu.json timeout 900 $T --master-port $PORT tools/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
echo "rc=$?"; grep MULTI_GPU_CHECK $OUT/multi_check_$N.log | cut -c1-300; grep -v MULTI_GPU_CHECK $OUT/multi_check_$N.log | filter | tail -15
python - <<PY
import json
try:
    r=json.load(open("$OUT/multi_gpu_check_${N}gpu.json"))
    for k,v in r.items():
        if isinstance(v,dict):
            print(f"  {k:24s} ok={v['ok']} nvls={v['nvls']} grad_err={v.get('grad_vs_oracle_rel_err')} param_err_max={max(v['param_vs_oracle_rel_err']):.2e} mom_err={v.get('momentum_vs_oracle_rel_err')}")
except Exception as e: print("no report", e)
PY

next; echo "== sync sweep N=$N"
timeout 600 $T --master-port $PORT benchmarks/sync_sweep.py --out $OUT/sync_sweep_$N.json 2>&1 | filter | tail -18

bench() { # name, extra args...
  name=$1; shift; next
  timeout 400 $T --master-port $PORT bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>$OUT/bench_${N}_$name.err > $OUT/bench_${N}_$name.json
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_${N}_$name.json") if l.startswith("{")][-1]
    print(f"N=$N {'$name':28s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms {d['e2e']['value']:9.0f} img/s | launches/step {d.get('gpu_launches',0)/d['steps']:.0f} notes {d.get('impl_notes',{}).get('nvls')} clocks {d['clocks'].get('sm_mhz')}/{d['clocks'].get('reasons')}")
except Exception as e:
    print("N=$N $name FAILED", e); print(open("$OUT/bench_${N}_$name.err").read()[-1200:])
PY
}
echo "== bench N=$N"
bench all_reduce --mode all_reduce
bench all_reduce_p2p --mode all_reduce --no-nvls
bench all_reduce_overlap --mode all_reduce --overlap on
bench all_reduce_overlap_p2p --mode all_reduce --overlap on --no-nvls
bench ddp --mode ddp
bench ddp_p2p --mode ddp --no-nvls
bench ddp_unfused --mode ddp --no-fuse-optimizer
bench gather --mode gather
if [ "$WHAT" = "full" ]; then
  bench all_reduce_b256 --mode all_reduce --per-rank-batch 256
  bench ddp_b256 --mode ddp --per-rank-batch 256
  bench gather_b256 --mode gather --per-rank-batch 256
  bench ref_all_reduce --mode all_reduce --impl reference
  bench ref_ddp --mode ddp --impl reference
  bench ref_gather --mode gather --impl reference
  bench ref_all_reduce_b256 --mode all_reduce --impl reference --per-rank-batch 256
  bench ref_ddp_b256 --mode ddp --impl reference --per-rank-batch 256
  bench ref_gather_b256 --mode gather --impl reference --per-rank-batch 256
fi

echo "== timelines N=$N"
for m in ddp all_reduce_overlap all_reduce; do
  next
  timeout 400 $T --master-port $PORT tools/overlap_timeline.py --mode $m --out $OUT/overlap_${m}_$N.txt > $OUT/overlap_${m}_$N.log 2>&1
  echo "timeline $m rc=$?"; grep "^# " $OUT/overlap_${m}_$N.txt 2>/dev/null | tail -6 | cut -c1-260
done
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)" > /dev/null 2>&1

if [ "$WHAT" = "full" ]; then
  echo "== entry scripts N=$N"
  timeout 1500 python -m pytest tests/test_gpu_cli.py -q -p no:cacheprovider -k "flag_launched or start_ddp" 2>&1 | tail -8
fi
