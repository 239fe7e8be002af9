"""Turn the JSON artefacts under profiles/raw (copied from gpurun_out/) into the markdown tables the docs cite:
profiles/sync_sweep_N.md, profiles/mode_sweep.md, profiles/bench_multi_gpu.md.   python tools/make_round2_profiles.py"""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAW = os.path.join(ROOT, "profiles", "raw")
OUT = os.path.join(ROOT, "profiles")


def load(path):
    with open(path) as fh:
        txt = fh.read().strip()
    return json.loads(txt) if txt else None


def sync_sweeps():
    for path in sorted(glob.glob(os.path.join(RAW, "sync_sweep_*.json"))):
        d = load(path)
        w = d["world"]
        lines = [f"# Gradient-sync sweep, {w} GPUs (benchmarks/sync_sweep.py)", "",
                 f"Transport: {d.get('transport', 'ipc').upper()}; multicast bound: {d.get('nvls')}.  Device-timed (CUDA events, max "
                 "over ranks), eager launches, 20 iterations after 5 warm-ups.  Roofline fractions are against the NOMINAL "
                 f"{d['link_gbs']:.0f} GB/s per direction per GPU (BASELINE.json); in brackets against the measured "
                 f"{d.get('link_measured_gbs', 770):.0f} GB/s peer copy.  All-reduce roofline = 2(N-1)/N x S / link; "
                 "gather->scatter = (N-1) x S / link (rank-0 ingress).  NVLS = multimem.ld_reduce / multimem.st "
                 "(reduction inside the NVSwitch); P2P = peer loads + stores.", "",
                 "| message | bytes | NVLS us | P2P us | NCCL us | best vs NCCL | all-reduce roofline (best) | gather->scatter ours us | "
                 "NCCL gather+stack+mean+scatter us | speed-up | gather roofline |", "|---|---|---|---|---|---|---|---|---|---|---|"]
        for r in d["rows"]:
            nv = r.get("nvls_allreduce_us")
            best = min(x for x in (nv, r["ours_allreduce_us"]) if x)
            roof = max(x for x in (r.get("nvls_roofline_frac"), r["allreduce_roofline_frac"]) if x is not None)
            roof770 = roof * d["link_gbs"] / d.get("link_measured_gbs", 770.0)
            lines.append(f"| {r['label']} | {r['bytes']} | {nv and f'{nv:.1f}' or '-'} | {r['ours_allreduce_us']:.1f} | "
                         f"{r['nccl_allreduce_us']:.1f} | {r['nccl_allreduce_us'] / best:.2f}x | {roof:.2f} ({roof770:.2f}) | "
                         f"{r['ours_gather_scatter_us']:.1f} | {r['nccl_gather_scatter_us']:.1f} | "
                         f"{r['nccl_gather_scatter_us'] / r['ours_gather_scatter_us']:.2f}x | {r['gather_roofline_frac']:.2f} "
                         f"({r.get('gather_roofline_frac_770', 0):.2f}) |")
        if d.get("grid_sweep"):
            lines += ["", "Whole arena (36.9 MB) against the number of CTAs of the collective:", "",
                      "| CTAs | P2P us | NVLS us |", "|---|---|---|"]
            for g in d["grid_sweep"]:
                nv = g["nvls_us"]
                lines.append(f"| {g['blocks']} | {g['p2p_us']:.1f} | {nv and f'{nv:.1f}' or '-'} |")
        with open(os.path.join(OUT, f"sync_sweep_{w}.md"), "w") as fh:
            fh.write("\n".join(lines) + "\n")


def mode_sweeps():
    lines = ["# Gradient-sync variants inside one job (tools/mode_sweep.py)", "",
             "Global batch 256 (strong scaling), VGG-11, whole step as a CUDA graph.  All variants are built in the same "
             "processes and timed in turns (4 rounds x 30 steps, CUDA events, max over ranks): separate bench.py runs of a "
             "0.5 ms step differ by several per cent, these do not.  `capN` = gradient bucket cap in MB, `bN` = CTAs of a "
             "per-bucket collective, `p2p` = NVLS switched off.", ""]
    for path in sorted(glob.glob(os.path.join(RAW, "mode_sweep_*.json"))):
        d = load(path)
        tag = " (final defaults, second box)" if "final" in os.path.basename(path) else ""
        lines += [f"## {d['world']} GPUs, {d['per_rank_batch']} images per rank{tag}", "",
                  "| variant | NVLS | min ms/step | median | images/s |", "|---|---|---|---|---|"]
        for r in d["rows"]:
            lines.append(f"| {r['variant']} | {'yes' if r['nvls'] else 'no'} | {r['ms_min']:.4f} | {r['ms_median']:.4f} | "
                         f"{r['images_per_sec']:.0f} |")
        lines.append("")
    with open(os.path.join(OUT, "mode_sweep.md"), "w") as fh:
        fh.write("\n".join(lines))


def bench_table():
    rows = []
    for path in sorted(glob.glob(os.path.join(RAW, "bench_*.json"))):
        d = load(path)
        if not d or "ms_per_step" not in d:
            continue
        name = os.path.basename(path)[len("bench_"):-len(".json")]
        n, _, variant = name.partition("_")
        rows.append((int(n), variant, d))
    lines = ["# bench.py lines, ours and the unmodified reference (`--impl reference`)", "",
             "Device-timed (CUDA events, max over ranks) after W warm-up steps plus a settle phase (0.25 s of untimed steps, "
             "so that NVML initialisation and the SM clock ramp are outside the timed region); `e2e` adds the pinned-host -> "
             "device copy of every batch and a `loss.item()` per step.  `b256` = per-rank batch 256 (the reference's own "
             "convention, weak scaling), otherwise global batch 256.  Raw JSON: profiles/raw/bench_*.json.", "",
             "| GPUs | run | ms/step | images/s | e2e ms/step | e2e images/s | launches/step | NVLS | clocks MHz / reasons |",
             "|---|---|---|---|---|---|---|---|---|"]
    for n, variant, d in sorted(rows, key=lambda r: (r[0], r[1])):
        notes = d.get("impl_notes", {})
        lines.append(f"| {n} | {variant} | {d['ms_per_step']:.4f} | {d['value']:.0f} | {d['e2e']['ms_per_step']:.4f} | "
                     f"{d['e2e']['value']:.0f} | {d.get('gpu_launches', 0) / max(d['steps'], 1):.0f} | {notes.get('nvls', '-')} | "
                     f"{d['clocks'].get('sm_mhz')} / {d['clocks'].get('reasons')} |")
    lines += ["", "Reference rows measured in round 1 on the same machine type (the reference code has not changed), global "
              "batch 256: 4 GPUs all_reduce / gather / ddp 3.927 / 8.966 / 3.493 ms per step; 8 GPUs 5.881 / 10.185 / 4.519 ms."]
    with open(os.path.join(OUT, "bench_multi_gpu.md"), "w") as fh:
        fh.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    sync_sweeps()
    mode_sweeps()
    bench_table()
    print("profiles written")
