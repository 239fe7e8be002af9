#!/bin/bash
# 4- / 8-GPU pass: oracle check, sync sweep, in-job variant A/B, official bench lines, weak-scaling rows, timeline.
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version\|^W[0-9]'; }
echo "== multi_gpu_check N=$N"
CHECK_P2P_TOO=0 CHECK_OUT=$OUT/multi_gpu_check_${N}gpu.json timeout 600 $T --master-port 29500 tools/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
echo "rc=$?"; grep -v MULTI_GPU_CHECK $OUT/multi_check_$N.log | filter | grep -v "^\[rank" | tail -3
python - <<PY
import json
try:
    r=json.load(open("$OUT/multi_gpu_check_${N}gpu.json"))
    print("overall ok:", r["ok"])
    for k,v in r.items():
        if isinstance(v,dict):
            print(f"  {k:26s} ok={v['ok']} nvls={v['nvls']} grad_err={v.get('grad_vs_oracle_rel_err')} param_err_max={max(v['param_vs_oracle_rel_err']):.2e} loss_diff_max={max(v['loss_vs_oracle']):.1e} mom_err={v.get('momentum_vs_oracle_rel_err')} ident={all(x for kk,x in v.items() if kk.startswith('identical'))}")
except Exception as e: print("no report", e)
PY
echo "== mode sweep N=$N"
timeout 600 $T --master-port 29502 tools/mode_sweep.py --out $OUT/mode_sweep_$N.json 2>&1 | filter | grep "^N="
echo "== sync sweep N=$N"
timeout 600 $T --master-port 29501 benchmarks/sync_sweep.py --out $OUT/sync_sweep_$N.json 2>&1 | filter | grep -v "^\[rank" | tail -24
PORT=29510
bench() { name=$1; shift; PORT=$((PORT+1))
  timeout 400 $T --master-port $PORT bench.py --gpus $N --steps 40 --warmup 8 "$@" 2>$OUT/bench_${N}_$name.err > $OUT/bench_${N}_$name.json
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_${N}_$name.json") if l.startswith("{")][-1]
    print(f"N=$N {'$name':26s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms {d['e2e']['value']:9.0f} img/s | launches/step {d.get('gpu_launches',0)/d['steps']:.0f} nvls {d.get('impl_notes',{}).get('nvls')} overlap {d.get('impl_notes',{}).get('overlap')} clocks {d['clocks'].get('sm_mhz')}/{d['clocks'].get('reasons')}")
except Exception as e:
    print("N=$N $name FAILED", e); print(open("$OUT/bench_${N}_$name.err").read()[-1200:])
PY
}
echo "== bench N=$N"
bench all_reduce --mode all_reduce
bench ddp --mode ddp
bench gather --mode gather
bench all_reduce_b256 --mode all_reduce --per-rank-batch 256
bench ddp_b256 --mode ddp --per-rank-batch 256
bench gather_b256 --mode gather --per-rank-batch 256
bench ref_all_reduce_b256 --mode all_reduce --impl reference --per-rank-batch 256
if [ "$2" = "ref" ]; then
  bench ref_all_reduce --mode all_reduce --impl reference
  bench ref_ddp --mode ddp --impl reference
  bench ref_gather --mode gather --impl reference
fi
echo "== timeline N=$N"
timeout 300 $T --master-port 29530 tools/overlap_timeline.py --mode ddp --out $OUT/overlap_ddp_$N.txt > $OUT/overlap_ddp_$N.log 2>&1
echo "timeline ddp rc=$?"; grep "^# " $OUT/overlap_ddp_$N.txt 2>/dev/null | tail -9 | cut -c1-230
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)" > /dev/null 2>&1
