"""Summarise the trimmed `ncu --set full ... --page raw --csv` exports of tools/gpu_ncu.sh into profiles/*.md:
   python tools/summarize_ncu_csv.py gpurun_out/ncu_conv.csv profiles/ncu_conv_gemm.md "title" """
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
note = sys.argv[4] if len(sys.argv) > 4 else ""
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:  # noqa: BLE001
    pass
hbm = float(peaks.get("hbm_gbs", 6551.0))
rows = list(csv.reader(open(path)))
hdr, data = rows[0], rows[2:]
col = {h: i for i, h in enumerate(hdr)}


def g(r, name, default=0.0):
    i = col.get(name)
    if i is None or i >= len(r) or r[i] == "":
        return default
    try:
        return float(r[i].replace(",", ""))
    except ValueError:
        return default


def unit(name):
    i = col.get(name)
    return rows[1][i] if i is not None else ""


def to_mb(v, u):
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)


lines = [f"# {title}", "", f"`ncu --set full --clock-control none`, one B200, values per launch.  DRAM % is of the measured "
         f"copy bandwidth in MEASURED_PEAKS.json ({hbm:.0f} GB/s); tensor % = `sm__pipe_tensor_cycles_active` of elapsed "
         "cycles.  ncu serialises launches and flushes caches between them, so durations are larger than in the replayed "
         "CUDA graph: compare shares and utilisation, not absolutes." + (" " + note if note else ""), "",
         "| kernel | grid | block | cluster | regs | us | DRAM read MB | DRAM write MB | DRAM GB/s | DRAM % of measured | L2 % | tensor % | warps/SM |",
         "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
tot_t = tot_tensor = 0.0
for r in data:
    name = re.sub(r"\(.*$", "", r[col["Kernel Name"]]).replace("void ", "")
    t = g(r, "gpu__time_duration.sum")
    t_us = t * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit("gpu__time_duration.sum"), 1.0)
    rd = to_mb(g(r, "dram__bytes_read.sum"), unit("dram__bytes_read.sum"))
    wr = to_mb(g(r, "dram__bytes_write.sum"), unit("dram__bytes_write.sum"))
    gbs = (rd + wr) / t_us * 1e3 if t_us else 0.0
    tensor = g(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed")
    tot_t += t_us
    tot_tensor += tensor * t_us
    lines.append(f"| `{name}` | {r[col['Grid Size']]} | {r[col['Block Size']]} | {int(g(r, 'launch__cluster_size'))} | "
                 f"{int(g(r, 'launch__registers_per_thread'))} | {t_us:.1f} | {rd:.1f} | {wr:.1f} | {gbs:.0f} | "
                 f"{100 * gbs / hbm:.0f} | {g(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | {tensor:.0f} | "
                 f"{g(r, 'sm__warps_active.avg.per_cycle_active'):.1f} |")
lines += ["", f"{len(data)} launches, {tot_t:.0f} us in total; time-weighted tensor-pipe activity "
          f"{tot_tensor / max(tot_t, 1e-9):.1f} % of elapsed cycles."]
with open(out, "w") as fh:
    fh.write("\n".join(lines) + "\n")
print(lines[-1])
