"""Run bench.py under a few conv tuning settings (env DPB_CONV_NT / DPB_CONV_TARGET_CTAS) and print ms/step."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
configs = [(128, 296), (256, 296), (128, 148), (256, 148), (128, 444)]
for batch in (256, 32):
    for nt, ctas in configs:
        env = dict(os.environ, DPB_CONV_NT=str(nt), DPB_CONV_TARGET_CTAS=str(ctas))
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "30", "--warmup", "5", "--per-rank-batch", str(batch)]
        p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
        try:
            d = json.loads(p.stdout.strip().splitlines()[-1])
            print(f"batch {batch:3d} NT {nt:3d} target_ctas {ctas:3d}: {d['ms_per_step']:.4f} ms/step device, "
                  f"{d['e2e']['ms_per_step']:.4f} ms e2e, {d['gpu_launches']} launches", flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"batch {batch} NT {nt} ctas {ctas}: FAILED {e} {p.stderr[-300:]}", flush=True)
