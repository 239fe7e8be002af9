"""Run bench.py under a few tuning settings (environment knobs read by _native.load) and print ms/step.
usage: python tools/tune_sweep.py "DPB_PDL=0" "DPB_PDL=1" "DPB_PDL=1 DPB_CONV_MIN_K_ITERS=4" ..."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
configs = sys.argv[1:] or ["DPB_PDL=1", "DPB_PDL=0"]
for batch in (256, 32):
    for cfg in configs:
        env = dict(os.environ)
        for kv in cfg.split():
            k, v = kv.split("=")
            env[k] = v
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "40", "--warmup", "5", "--per-rank-batch", str(batch)]
        p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
        try:
            d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
            print(f"batch {batch:3d} [{cfg:45s}] {d['ms_per_step']:.4f} ms/step device, {d['e2e']['ms_per_step']:.4f} ms e2e, "
                  f"{d['gpu_launches'] // d['steps']} launches/step, loss {d['final_loss']:.3f}", flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"batch {batch} [{cfg}]: FAILED {e} {p.stderr[-400:]}", flush=True)
