#!/bin/bash
# final multi-GPU pass: official bench lines (with the settle phase) + in-job sweep with the shipped defaults + timeline
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version\|^W[0-9]'; }
PORT=29510
bench() { name=$1; shift; PORT=$((PORT+1))
  timeout 400 $T --master-port $PORT bench.py --gpus $N --steps 40 --warmup 8 "$@" 2>$OUT/bench_${N}_$name.err > $OUT/bench_${N}_$name.json
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_${N}_$name.json") if l.startswith("{")][-1]
    print(f"N=$N {'$name':26s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms {d['e2e']['value']:9.0f} img/s | launches/step {d.get('gpu_launches',0)/d['steps']:.0f} nvls {d.get('impl_notes',{}).get('nvls')} settle {d['clocks'].get('settle_steps')} clocks {d['clocks'].get('sm_mhz')}/{d['clocks'].get('reasons')}")
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
[ "$2" = "ref" ] && bench ref_all_reduce --mode all_reduce --impl reference
echo "== mode sweep N=$N"
timeout 600 $T --master-port 29502 tools/mode_sweep.py --out $OUT/mode_sweep_final_$N.json 2>&1 | filter | grep "^N="
echo "== timeline N=$N"
for m in ddp all_reduce_overlap; do
timeout 300 $T --master-port 29530 tools/overlap_timeline.py --mode $m --out $OUT/overlap_${m}_$N.txt > $OUT/overlap_${m}_$N.log 2>&1
echo "timeline $m rc=$?"; grep "^# " $OUT/overlap_${m}_$N.txt 2>/dev/null | tail -7 | cut -c1-200
done
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)" > /dev/null 2>&1
