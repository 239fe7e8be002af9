#!/usr/bin/env python
"""Install the UNMODIFIED reference (BrianZCS/distributed_pytorch) under baseline/_ref for `bench.py --impl reference`.

1. The prescribed way: `pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target
   baseline/_ref <copy of /root/reference>`.  The reference is a flat directory of scripts with no setup.py /
   pyproject.toml, so pip cannot build it (outcome recorded in DESIGN.md).
2. Fallback: byte-for-byte copy of its files into baseline/_ref (git-ignored; travels to the GPU box with gpurun).

Nothing in baseline/_ref is edited: bench.py imports the reference's own `model.py` and `main_*.py` from there and calls
their `VGG11()` / `train_model()` as they are.
"""
import filecmp
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SRC = os.environ.get("REFERENCE_SRC", "/root/reference")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"reference source {SRC} not found", file=sys.stderr)
        return 1
    os.makedirs(DEST, exist_ok=True)
    tmp = tempfile.mkdtemp(prefix="ref_src_")
    work = os.path.join(tmp, "reference")
    shutil.copytree(SRC, work)
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
           "/opt/wheelhouse", "--target", DEST, work]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(HERE, "install_reference.log")
    with open(log, "w") as fh:
        fh.write("$ " + " ".join(cmd) + f"\nexit code {proc.returncode}\n" + proc.stdout[-2000:] + proc.stderr[-2000:])
    if proc.returncode == 0 and os.path.exists(os.path.join(DEST, "model.py")):
        print("pip install succeeded")
    else:
        print(f"pip install failed (exit {proc.returncode}; see {log}): copying the sources verbatim")
        for name in sorted(os.listdir(SRC)):
            s = os.path.join(SRC, name)
            if os.path.isfile(s):
                shutil.copy2(s, os.path.join(DEST, name))
    shutil.rmtree(tmp, ignore_errors=True)
    same = all(filecmp.cmp(os.path.join(SRC, n), os.path.join(DEST, n), shallow=False)
               for n in os.listdir(SRC) if os.path.isfile(os.path.join(SRC, n)))
    print(f"baseline/_ref populated; identical to {SRC}: {same}")
    return 0 if same else 2


if __name__ == "__main__":
    raise SystemExit(main())
