#!/bin/bash
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest"
timeout 900 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider -k "not cli" > $OUT/a3_pytest.log 2>&1
echo "pytest rc=$?"; tail -12 $OUT/a3_pytest.log
b() { name=$1; shift
  timeout 300 env "$@" > $OUT/a3_bench_$name.json 2> $OUT/a3_bench_$name.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/a3_bench_$name.json") if l.startswith("{")][-1]
    print(f"{'$name':34s} {d['ms_per_step']:.4f} ms/step e2e {d['e2e']['ms_per_step']:.4f} launches/step {d['gpu_launches']/d['steps']:.0f} loss {d['final_loss']:.3f}")
except Exception as e:
    print("$name FAILED", e); print(open("$OUT/a3_bench_$name.err").read()[-1200:])
PY
}
for bs in 256 32; do
  B="python bench.py --steps 30 --warmup 5 --mode single --global-batch $bs"
  b b${bs} A=1 $B
  b b${bs}_nocluster DPB_CONV_CLUSTER_MAX=1 $B
  b b${bs}_cluster8 DPB_CONV_CLUSTER_MAX=8 $B
  b b${bs}_cluster4 DPB_CONV_CLUSTER_MAX=4 $B
  b b${bs}_nowgradovl DPB_WGRAD_OVERLAP=0 $B
  b b${bs}_bwd222 DPB_BWD_TARGET_CTAS=222 $B
  b b${bs}_bwd296 DPB_BWD_TARGET_CTAS=296 $B
  b b${bs}_cluster16 DPB_CONV_CLUSTER_MAX=16 $B
  b b${bs}_mink4 DPB_CONV_MIN_K_ITERS=4 $B
  b b${bs}_mink16 DPB_CONV_MIN_K_ITERS=16 $B
done
python tools/conv_timing.py 32 > $OUT/conv_timing_b32.txt 2>&1; tail -25 $OUT/conv_timing_b32.txt
for bs in 32 256; do
  python tools/step_timeline.py $bs $OUT/timeline_b$bs.txt > $OUT/a3_timeline_$bs.log 2>&1; echo "timeline b$bs rc=$?"; tail -14 $OUT/timeline_b$bs.txt
done
python -c "from distributed_pytorch_b200 import _native as N; N.build(force=True)"
