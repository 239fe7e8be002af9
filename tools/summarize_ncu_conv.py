"""Summarise an `ncu --set full` capture of the conv GEMM launches of one training step (page raw, csv) into a table:
python tools/summarize_ncu_conv.py gpurun_out/prof_conv.ncu-rep profiles/ncu_conv_gemm.md"""
import csv
import io
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]


def col(name):
    return hdr.index(name) if name in hdr else None


def find(sub):
    for i, h in enumerate(hdr):
        if sub in h:
            return i
    return None


C = {
    "name": col("Kernel Name"), "grid": col("Grid Size"),
    "time": col("gpu__time_duration.sum"),
    "tensor": col("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
    "tensor_max": col("sm__pipe_tensor_cycles_active.max.pct_of_peak_sustained_elapsed"),
    "clk": col("sm__cycles_elapsed.max.per_second"),
    "tc_inst": col("sm__inst_executed_pipe_tc.sum.pct_of_peak_sustained_active"),
    "dram_r": col("dram__bytes_read.sum"), "dram_w": col("dram__bytes_write.sum"),
    "dram_pct": col("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    "l2_pct": col("lts__t_sectors.avg.pct_of_peak_sustained_elapsed"),
    "warps": col("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "regs": col("launch__registers_per_thread"), "smem": col("launch__shared_mem_per_block_dynamic"),
    "occ_smem": col("launch__occupancy_limit_shared_mem"), "waves": col("launch__waves_per_multiprocessor"),
    "sm_clk": col("sm__cycles_elapsed.max"), "bank": col("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
}


def val(r, key, scale=1.0):
    i = C[key]
    if i is None:
        return float("nan")
    try:
        v = float(r[i].replace(",", ""))
    except ValueError:
        return float("nan")
    u = units[i]
    if key in ("dram_r", "dram_w"):
        v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
    if key == "time":
        v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(u, 1.0)
    return v * scale


MODE = {"0": "fprop", "1": "dgrad", "2": "wgrad"}
with open(out, "w") as fh:
    fh.write("# tcgen05 conv GEMM kernels of one VGG-11 training step (batch 256), `ncu --set full --clock-control none`\n\n"
             "One row per launch, in launch order (7 fprop, then per layer wgrad + dgrad going backward).  `tensor pipe` is\n"
             "`sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed`: the share of the kernel's elapsed cycles the\n"
             "5th-gen tensor cores were busy, averaged over the 148 SMs (`max` = the busiest SM).  Times under ncu are serialised single launches with cold caches.\n\n"
             "| # | kernel | grid | us | tensor pipe % (avg / max SM) | SM GHz | DRAM read MB | DRAM write MB | DRAM % | L2 % | warps active % | regs | dyn smem KB | CTAs/SM (smem) | waves |\n"
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
    tot_t = tot_tt = 0.0
    for k, r in enumerate(data):
        if C["name"] is None or len(r) <= C["name"]:
            continue
        name = r[C["name"]]
        tpl = name[name.find("<") + 1:name.find(">")].split(",") if "<" in name else []
        label = f"{MODE.get(tpl[0].strip(), '?')} NT={tpl[1].strip()} stages={tpl[2].strip()}" if len(tpl) >= 3 else name[:40]
        t = val(r, "time")
        tp = val(r, "tensor")
        tot_t += t
        tot_tt += t * tp / 100.0
        fh.write(f"| {k} | {label} | {r[C['grid']] if C['grid'] is not None else ''} | {t:.1f} | {tp:.1f} / {val(r, 'tensor_max'):.1f} | {val(r, 'clk'):.2f} | {val(r, 'dram_r'):.1f} | "
                 f"{val(r, 'dram_w'):.1f} | {val(r, 'dram_pct'):.1f} | {val(r, 'l2_pct'):.1f} | {val(r, 'warps'):.1f} | {val(r, 'regs'):.0f} | "
                 f"{val(r, 'smem'):.0f} | {val(r, 'occ_smem'):.0f} | {val(r, 'waves'):.2f} |\n")
    fh.write(f"\nAll {len(data)} launches: {tot_t:.0f} us, time-weighted tensor pipe activity {100 * tot_tt / max(tot_t, 1e-9):.1f} %.\n")
print(open(out).read())
