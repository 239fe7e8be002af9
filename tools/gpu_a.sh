#!/bin/bash
# round-2 call A (1 GPU): full GPU test-suite, bench at batch 256 / 32 with variants, launch lists
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.used --format=csv > $OUT/a_smi.txt 2>&1
echo "== pytest" 
timeout 1100 python -m pytest tests -m gpu -q --maxfail=12 --junitxml=$OUT/gpu_tests.xml -p no:cacheprovider > $OUT/a_pytest.log 2>&1
echo "pytest rc=$?"; tail -25 $OUT/a_pytest.log
b() { # name, env..., args
  name=$1; shift
  echo "== bench $name"
  timeout 300 env "$@" > $OUT/a_bench_$name.json 2> $OUT/a_bench_$name.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/a_bench_$name.json") if l.startswith("{")][-1]
    print("$name", round(d["ms_per_step"],4), "ms/step", round(d["value"]), "img/s e2e", round(d["e2e"]["ms_per_step"],4), "launches/step", d["gpu_launches"]/d["steps"], "loss", round(d["final_loss"],3), d["clocks"])
except Exception as e:
    print("$name FAILED", e); print(open("$OUT/a_bench_$name.err").read()[-1500:])
PY
}
b b256 A=1 python bench.py --steps 30 --warmup 5 --mode single
b b32 A=1 python bench.py --steps 30 --warmup 5 --mode single --global-batch 32
b b256_nocluster DPB_CONV_CLUSTER_MAX=1 python bench.py --steps 30 --warmup 5 --mode single
b b32_nocluster DPB_CONV_CLUSTER_MAX=1 python bench.py --steps 30 --warmup 5 --mode single --global-batch 32
b b256_cluster8 DPB_CONV_CLUSTER_MAX=8 python bench.py --steps 30 --warmup 5 --mode single
b b32_cluster8 DPB_CONV_CLUSTER_MAX=8 python bench.py --steps 30 --warmup 5 --mode single --global-batch 32
b b256_nofold DPB_BN_FOLD=0 python bench.py --steps 30 --warmup 5 --mode single
b b32_nofold DPB_BN_FOLD=0 python bench.py --steps 30 --warmup 5 --mode single --global-batch 32
b b256_allreduce1 A=1 python bench.py --steps 30 --warmup 5
b b256_ddp1 A=1 python bench.py --steps 30 --warmup 5 --mode ddp
b ref256 A=1 python bench.py --steps 20 --warmup 5 --impl reference
echo "== launch lists"
for bs in 32 256; do
  python tools/step_profile.py $bs 3 > $OUT/a_step_$bs.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:dpb -s 120 -c 80 --csv --log-file $OUT/launches_b$bs.csv python tools/step_profile.py $bs 4 > $OUT/a_ncu_$bs.log 2>&1
  echo "ncu b$bs rc=$?"; tail -3 $OUT/a_step_$bs.log
done
