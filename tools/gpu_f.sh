#!/bin/bash
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_cli.py -q -p no:cacheprovider 2>&1 | tail -3
for m in all_reduce ddp gather single; do
timeout 300 python bench.py --steps 40 --warmup 8 --mode $m > $OUT/bench_1_$m.json 2> $OUT/bench_1_$m.err
python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/bench_1_$m.json") if l.startswith("{")][-1]
    print(f"N=1 {'$m':12s} {d['ms_per_step']:.4f} ms/step {d['value']:9.0f} img/s | e2e {d['e2e']['ms_per_step']:.4f} ms | launches/step {d['gpu_launches']/d['steps']:.0f} settle {d['clocks'].get('settle_steps')}")
except Exception as e:
    print("$m FAILED", e); print(open("$OUT/bench_1_$m.err").read()[-1500:])
PY
done
