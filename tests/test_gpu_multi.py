"""Real multi-GPU run (one process per GPU over NVLink peer memory) of all sync modes; skipped on a 1-GPU box.
The checks themselves live in tools/multi_gpu_check.py."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(900)
def test_all_sync_modes_on_real_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tools", "multi_gpu_check.py")]
    env = dict(os.environ)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    env["CHECK_OUT"] = os.path.join(ROOT, "gpurun_out", f"multi_gpu_check_{world}gpu.json")
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=850, env=env)
    lines = [l for l in proc.stdout.splitlines() if l.startswith("MULTI_GPU_CHECK ")]
    assert proc.returncode == 0 and lines, (proc.stdout[-2000:], proc.stderr[-3000:])
    rep = json.loads(lines[-1][len("MULTI_GPU_CHECK "):])
    assert rep["ok"], rep
