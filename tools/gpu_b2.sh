#!/bin/bash
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version\|^W[0-9]'; }
echo "== multi_gpu_check N=$N"
CHECK_OUT=$OUT/multi_gpu_check_${N}gpu.json timeout 900 $T --master-port 29500 tools/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
echo "rc=$?"; grep -v MULTI_GPU_CHECK $OUT/multi_check_$N.log | filter | grep -v "^\[rank" | tail -5
python - <<PY
import json
try:
    r=json.load(open("$OUT/multi_gpu_check_${N}gpu.json"))
    print("overall ok:", r["ok"])
    for k,v in r.items():
        if isinstance(v,dict):
            print(f"  {k:26s} ok={v['ok']} nvls={v['nvls']} grad_err={v.get('grad_vs_oracle_rel_err')} param_err_max={max(v['param_vs_oracle_rel_err']):.2e} mom_err={v.get('momentum_vs_oracle_rel_err')} ident={all(x for kk,x in v.items() if kk.startswith('identical'))}")
except Exception as e: print("no report", e)
PY
echo "== sync sweep N=$N"
timeout 600 $T --master-port 29501 benchmarks/sync_sweep.py --out $OUT/sync_sweep_$N.json 2>&1 | filter | grep -E "whole|world|ddp_bucket|64MB|16MB| 4KB"
PORT=29510
bench() { name=$1; shift; PORT=$((PORT+1))
  timeout 400 env $ENVV $T --master-port $PORT bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>$OUT/bench_${N}_$name.err > $OUT/bench_${N}_$name.json
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_${N}_$name.json") if l.startswith("{")][-1]
    print(f"N=$N {'$name':30s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms | launches/step {d.get('gpu_launches',0)/d['steps']:.0f} nvls {d.get('impl_notes',{}).get('nvls')}")
except Exception as e:
    print("N=$N $name FAILED", e); print(open("$OUT/bench_${N}_$name.err").read()[-1500:])
PY
}
ENVV="A=1"
bench all_reduce --mode all_reduce
bench all_reduce_p2p --mode all_reduce --no-nvls
bench all_reduce_overlap --mode all_reduce --overlap on
bench all_reduce_overlap_p2p --mode all_reduce --overlap on --no-nvls
bench ddp --mode ddp
bench ddp_p2p --mode ddp --no-nvls
bench ddp_cap25 --mode ddp --bucket-cap-mb 25
bench ddp_cap5 --mode ddp --bucket-cap-mb 5
ENVV="DPB_DDP_BLOCKS=16"; bench ddp_blocks16 --mode ddp
ENVV="DPB_DDP_BLOCKS=64"; bench ddp_blocks64 --mode ddp
ENVV="DPB_DDP_BLOCKS=64"; bench ddp_p2p_blocks64 --mode ddp --no-nvls
ENVV="A=1"
bench gather --mode gather
echo "== timelines"
for m in ddp all_reduce_overlap; do
  PORT=$((PORT+1))
  timeout 400 $T --master-port $PORT tools/overlap_timeline.py --mode $m --out $OUT/overlap_${m}_$N.txt > $OUT/overlap_${m}_$N.log 2>&1
  echo "timeline $m rc=$?"; grep "^# " $OUT/overlap_${m}_$N.txt 2>/dev/null | tail -8 | cut -c1-230
done
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)" > /dev/null 2>&1
if [ "$2" = "cli" ]; then
  timeout 900 python -m pytest tests/test_gpu_cli.py -q -p no:cacheprovider -k "overlap or start_ddp" 2>&1 | tail -5
fi
