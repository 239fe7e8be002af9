"""The five entry scripts on a real GPU, started the way the reference's README and start_ddp.sh start them
(reference README.md:4, start_ddp.sh:1, main_gather.py:95-146): one process per GPU, ``tcp://127.0.0.1:<port>``
rendezvous from ``--master-ip/--num-nodes/--rank`` (torchrun environment for main_ddp.py), 41+ iterations so that both
"Average Loss" lines and the "Avg Time for iteration 2-40" line print, a ragged last batch (eager fallback next to the
captured graph), the full evaluation pass and ``--save``.  Each run must exit 0, print the reference's line formats, show
a falling loss, and leave a checkpoint that loads into the reference's own ``model.VGG11()``."""
import os
import re
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PER_RANK = 41 * 256 + 80  # 42 iterations per rank, the last one with 80 images


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _world():
    n = torch.cuda.device_count()
    return 8 if n >= 8 else (4 if n >= 4 else (2 if n >= 2 else 1))


def _check_output(out: str, what: str):
    # (under torchrun the ranks share one stdout, so a window's line appears once per rank)
    first = [float(x) for x in re.findall(r"Epoch: 1, Iteration: 1-20, Average Loss: (\d+\.\d{3})", out)]
    second = [float(x) for x in re.findall(r"Epoch: 1, Iteration: 21-40, Average Loss: (\d+\.\d{3})", out)]
    assert first and len(first) == len(second), (what, out[-1500:])
    losses = [sum(first) / len(first), sum(second) / len(second)]
    m = re.search(r"Avg Time for iteration 2-40: ([\d.e-]+) seconds\.", out)
    assert m, (what, out[-1500:])
    assert re.search(r"Test set: Average loss: \d+\.\d{4}, Accuracy: \d+/\d+ \(\d+%\)", out), (what, out[-1500:])
    assert all(l == l and l < 50 for l in first + second), (what, first, second)
    assert losses[1] < losses[0], (what, "loss did not fall", losses)
    return float(m.group(1))


def _check_checkpoint(path: str, ref_model_module, ddp_prefix: bool):
    sd = torch.load(path, map_location="cpu")
    if ddp_prefix:
        assert all(k.startswith("module.") for k in sd), list(sd)[:3]
        sd = {k[len("module."):]: v for k, v in sd.items()}
    ref = ref_model_module.VGG11()
    ref.load_state_dict(sd)  # strict: all 58 keys, reference shapes
    opt = torch.load(path + ".optim", map_location="cpu")
    assert len(opt["state"]) == 34 and all("momentum_buffer" in st for st in opt["state"].values())
    assert all(torch.isfinite(st["momentum_buffer"]).all() for st in opt["state"].values())


def _run(cmds, what, timeout=600):
    procs = [subprocess.Popen(c, cwd=ROOT, env=e, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
             for c, e in cmds]
    outs = []
    for p in procs:
        try:
            o, err = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, o, err))
    for rc, o, err in outs:
        assert rc == 0, (what, rc, o[-1500:], err[-3000:])
    return outs


def _env(local_rank=None):
    e = dict(os.environ)
    e["PYTHONPATH"] = ROOT + os.pathsep + e.get("PYTHONPATH", "")
    if local_rank is not None:
        e["LOCAL_RANK"] = str(local_rank)
    return e


@pytest.mark.timeout(900)
def test_main_single_process(tmp_path, ref_model_module):
    save = str(tmp_path / "single.pt")
    outs = _run([([sys.executable, "main.py", "--synthetic", "--train-size", str(PER_RANK), "--test-size", "1000",
                   "--save", save], _env())], "main.py")
    t = _check_output(outs[0][1], "main.py")
    assert t < 0.1  # seconds per iteration: the loop is GPU-bound, not loader-bound
    _check_checkpoint(save, ref_model_module, ddp_prefix=False)


@pytest.mark.timeout(1200)
@pytest.mark.parametrize("script,extra", [("main_gather.py", []), ("main_all_reduce.py", []),
                                          ("main_all_reduce.py", ["--no-fuse-optimizer"]),
                                          ("main_all_reduce.py", ["--no-overlap"]), ("main_part3.py", []),
                                          ("main_part3.py", ["--no-fuse-optimizer"])])
def test_flag_launched_modes(tmp_path, ref_model_module, script, extra):
    world = _world()
    port = _free_port()
    save = str(tmp_path / "model.pt")
    cmds = []
    for r in range(world):
        cmds.append(([sys.executable, script, "--master-ip", "127.0.0.1", "--num-nodes", str(world), "--rank", str(r),
                      "--master-port", str(port), "--synthetic", "--train-size", str(PER_RANK * world), "--test-size",
                      "1000", "--save", save, *extra], _env(local_rank=r)))
    outs = _run(cmds, script)
    for rc, o, _ in outs:  # the reference prints on every rank
        _check_output(o, script)
    if script == "main_part3.py":
        assert "test" in outs[0][1].splitlines()[0:3]
    _check_checkpoint(save, ref_model_module, ddp_prefix=(script == "main_part3.py"))


@pytest.mark.timeout(1200)
def test_start_ddp_sh(tmp_path, ref_model_module):
    world = _world()
    save = str(tmp_path / "ddp.pt")
    e = _env()
    e.update(NPROC=str(world), MASTER_PORT=str(_free_port()))
    outs = _run([(["sh", "start_ddp.sh", "--synthetic", "--train-size", str(PER_RANK * world), "--test-size", "1000",
                   "--save", save], e)], "start_ddp.sh")
    out = outs[0][1]
    assert re.search(r"\[\d+\] Initializing process group with: \{'MASTER_ADDR': '127.0.0.1'", out), out[:800]
    _check_output(out, "start_ddp.sh")
    _check_checkpoint(save, ref_model_module, ddp_prefix=True)
