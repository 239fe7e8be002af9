"""Turn the raw measurement files that gpurun brought back (gpurun_out/, scratch) into the tracked summaries under
profiles/: bench lines (ours vs the unmodified reference), the gradient-sync sweep vs NCCL with roofline fractions, and
the per-kernel launch lists from ncu."""
import csv
import collections
import glob
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")


def load_json_line(path):
    try:
        with open(path) as fh:
            lines = [l for l in fh.read().splitlines() if l.startswith("{")]
        return json.loads(lines[-1]) if lines else None
    except (OSError, ValueError):
        return None


def bench_table():
    rows = []
    for path in sorted(glob.glob(os.path.join(OUT, "bench_ours_*_*.json"))):
        m = re.match(r"bench_ours_(\d+)_(\w+)\.json", os.path.basename(path))
        n, mode = int(m.group(1)), m.group(2)
        ours = load_json_line(path)
        ref = load_json_line(os.path.join(OUT, f"bench_ref_{n}_{mode}.json"))
        if ours:
            rows.append((n, mode, ours, ref))
    if not rows:
        return
    rows.sort(key=lambda r: (r[1], r[0]))
    with open(os.path.join(PROF, "bench_multi_gpu.md"), "w") as fh:
        fh.write("# bench.py, ours vs the unmodified reference (`--impl reference`), global batch 256\n\n"
                 "Our rows are the latest run of each configuration; the reference rows were taken earlier in the round on a box of\n"
                 "the same type (8 x B200, same image) -- the reference code did not change.  1-GPU: see the README / DESIGN table.\n"
                 "Device-timed (CUDA events, max over ranks), 30 steps after 5 warm-ups; `e2e` adds the pinned-host -> device\n"
                 "copy of every batch and a `loss.item()` per step.  Raw JSON lines: `profiles/raw/`.\n\n"
                 "| mode | GPUs | ours ms/step | ours img/s | ours e2e img/s | ref ms/step | ref img/s | ref e2e img/s | speed-up (e2e) | clocks (MHz, reasons) |\n"
                 "|---|---|---|---|---|---|---|---|---|---|\n")
        for n, mode, o, r in rows:
            ck = o.get("clocks", {})
            cks = f"{ck.get('sm_mhz')} / {ck.get('sm_max_mhz')} {ck.get('reasons')}"
            if r and "value" in r:
                fh.write(f"| {mode} | {n} | {o['ms_per_step']:.3f} | {o['value']:.0f} | {o['e2e']['value']:.0f} | "
                         f"{r['ms_per_step']:.3f} | {r['value']:.0f} | {r['e2e']['value']:.0f} | "
                         f"{o['e2e']['value'] / r['e2e']['value']:.2f}x | {cks} |\n")
            else:
                fh.write(f"| {mode} | {n} | {o['ms_per_step']:.3f} | {o['value']:.0f} | {o['e2e']['value']:.0f} | - | - | - | - | {cks} |\n")
    os.makedirs(os.path.join(PROF, "raw"), exist_ok=True)
    for n, mode, o, r in rows:
        with open(os.path.join(PROF, "raw", f"bench_{n}gpu_{mode}.json"), "w") as fh:
            json.dump({"ours": o, "reference": r}, fh, indent=1)


def sweep_tables():
    for path in sorted(glob.glob(os.path.join(OUT, "sync_sweep_*.json"))):
        name = os.path.basename(path)[:-5]
        with open(path) as fh:
            d = json.load(fh)
        with open(os.path.join(PROF, name + ".md"), "w") as fh:
            fh.write(f"# Gradient-sync sweep on {d['world']} B200 (benchmarks/sync_sweep.py), ours vs NCCL\n\n"
                     f"Eager launches, CUDA-event time, max over ranks.  Roofline: all-reduce 2(N-1)/N*S, gather->scatter (N-1)*S\n"
                     f"(rank-0 ingress) over the measured {d['link_gbs']:.0f} GB/s per direction per GPU.\n\n"
                     "| message | bytes | all-reduce ours us | NCCL all_reduce+div us | speed-up | bus GB/s | of roofline | "
                     "gather-scatter ours us | NCCL gather+mean+scatter us | speed-up | of roofline |\n|---|---|---|---|---|---|---|---|---|---|---|\n")
            for r in d["rows"]:
                fh.write(f"| {r['label']} | {r['bytes']} | {r['ours_allreduce_us']:.1f} | {r['nccl_allreduce_us']:.1f} | "
                         f"{r['nccl_allreduce_us'] / r['ours_allreduce_us']:.2f}x | {r['allreduce_busbw_gbs']:.0f} | "
                         f"{r['allreduce_roofline_frac']:.2f} | {r['ours_gather_scatter_us']:.1f} | "
                         f"{r['nccl_gather_scatter_us']:.1f} | {r['nccl_gather_scatter_us'] / r['ours_gather_scatter_us']:.2f}x | "
                         f"{r['gather_roofline_frac']:.2f} |\n")


def launch_tables():
    for path in sorted(glob.glob(os.path.join(OUT, "launches_b*.csv"))):
        name = os.path.basename(path)[:-4]
        with open(path, newline="") as fh:
            lines = [l for l in fh if not l.startswith("==")]
        rows = []
        for r in csv.DictReader(lines):
            if r.get("Metric Name") == "gpu__time_duration.sum":
                rows.append((re.sub(r"\(.*$", "", re.sub(r"^void ", "", r["Kernel Name"])), r.get("Grid Size", ""),
                             float(r["Metric Value"].replace(",", "")) / 1e3))
        if not rows:
            continue
        tot = sum(v for _, _, v in rows)
        agg = collections.OrderedDict()
        for k, _, v in rows:
            a = agg.setdefault(k, [0, 0.0])
            a[0] += 1
            a[1] += v
        with open(os.path.join(PROF, name + ".md"), "w") as fh:
            fh.write(f"# Kernel launch list of one eager training step ({name}), ncu --metrics gpu__time_duration.sum\n\n"
                     f"{len(rows)} launches, {tot:.1f} us summed (serialised, cold caches: compare shares, not absolutes).\n\n"
                     "| kernel | calls | us | share |\n|---|---|---|---|\n")
            for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                fh.write(f"| `{k}` | {n} | {v:.1f} | {100 * v / tot:.1f}% |\n")
            fh.write("\nPer launch, in order:\n\n```\n")
            for k, g, v in rows:
                fh.write(f"{v:9.1f} us  grid {g:16s} {k}\n")
            fh.write("```\n")


if __name__ == "__main__":
    os.makedirs(PROF, exist_ok=True)
    bench_table()
    sweep_tables()
    launch_tables()
    print("profiles written:", sorted(os.listdir(PROF)))
