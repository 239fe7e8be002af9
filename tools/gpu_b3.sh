#!/bin/bash
N=${1:-2}
OUT=gpurun_out
mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
filter() { grep -v '^\*\|OMP_NUM_THREADS\|^$\|NCCL version\|^W[0-9]'; }
echo "== multi_gpu_check N=$N"
CHECK_OUT=$OUT/multi_gpu_check_${N}gpu.json timeout 900 $T --master-port 29500 tools/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
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
PORT=29510
bench() { name=$1; shift; PORT=$((PORT+1))
  timeout 400 $T --master-port $PORT bench.py --gpus $N --steps 40 --warmup 8 "$@" 2>$OUT/bench_${N}_$name.err > $OUT/bench_${N}_$name.json
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_${N}_$name.json") if l.startswith("{")][-1]
    print(f"N=$N {'$name':30s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms | launches/step {d.get('gpu_launches',0)/d['steps']:.0f} nvls {d.get('impl_notes',{}).get('nvls')} overlap {d.get('impl_notes',{}).get('overlap')}")
except Exception as e:
    print("N=$N $name FAILED", e); print(open("$OUT/bench_${N}_$name.err").read()[-1500:])
PY
}
bench all_reduce --mode all_reduce
bench all_reduce_nooverlap --mode all_reduce --overlap off
bench ddp --mode ddp
bench gather --mode gather
echo "== cli"
timeout 1200 python -m pytest tests/test_gpu_cli.py tests/test_gpu_multi.py -q -p no:cacheprovider -k "not single_process" 2>&1 | tail -5
