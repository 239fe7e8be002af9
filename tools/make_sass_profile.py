"""SASS evidence for profiles/sass/: per-kernel mnemonic summary (tcgen05 = UTC*MMA, TMA = UTMALDG, tcgen05.ld = LDTM,
tcgen05.commit = UTCBAR, mbarrier = SYNCS, system-scope peer loads/stores) and full listings of the flagship kernels.
Runs on the CPU build box: cuobjdump -sass on the in-tree libdpb200.so."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_pytorch_b200", "_lib", "libdpb200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
FULL = {"conv_gemm_kernel<0, 128, 3, true>": "conv_gemm_fprop_nt128_cluster.sass",
        "conv_gemm_kernel<2, 128, 3, false>": "conv_gemm_wgrad_nt128.sass",
        "allreduce_nvls_kernel<16>": "allreduce_nvls.sass",
        "first_conv_fprop_tc_kernel": "first_conv_fprop_tc.sass",
        "first_conv_wgrad_tc_kernel": "first_conv_wgrad_tc.sass",
        "allreduce_twoshot_kernel<false, 8, 2>": "allreduce_twoshot_8ranks.sass"}


def demangle(name: str) -> str:
    try:
        return subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    except OSError:
        return name


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs = {}
    for part in re.split(r"\n\s*Function : ", txt)[1:]:
        funcs[part.split("\n", 1)[0].strip()] = part
    key = re.compile(r"\b(UTC[A-Z0-9]*MMA[A-Z0-9_.]*|UTMALDG[A-Z0-9_.]*|UTMAPF[A-Z0-9_.]*|LDTM[A-Z0-9_.]*|UTCBAR[A-Z0-9_.]*|"
                     r"SYNCS[A-Z0-9_.]*|HMMA[A-Z0-9_.]*|ELECT[A-Z0-9_.]*|UCGABAR[A-Z0-9_.]*|LDGMC[A-Z0-9_.]*|STGMC[A-Z0-9_.]*|REDGMC[A-Z0-9_.]*|"
                     r"LDS[A-Z0-9_.]*CLUSTER[A-Z0-9_.]*|MAPA[A-Z0-9_.]*|ATOMG[A-Z0-9_.]*|RED[A-Z0-9_.]*64[A-Z0-9_.]*)\b")
    sysop = re.compile(r"\b((?:LDG|STG|LD|ST)\.E[A-Z0-9_.]*SYS[A-Z0-9_.]*|MEMBAR\.[A-Z.]*SYS[A-Z.]*)")
    rows = []
    for name, body in funcs.items():
        d = re.sub(r"\(.*", "", demangle(name)).replace("void dpb::", "").replace("dpb::", "")
        ins = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        tc = collections.Counter(m.group(1) for l in ins for m in [key.search(l)] if m)
        so = collections.Counter(m.group(1) for l in ins for m in [sysop.search(l)] if m)
        rows.append((d, len(ins), tc, so))
        for want, fname in FULL.items():
            if want in demangle(name):
                with open(os.path.join(OUT, fname), "w") as fh:
                    fh.write("Function : " + body)
    rows.sort()
    with open(os.path.join(OUT, "mnemonic_summary.md"), "w") as fh:
        fh.write("# SASS evidence (cuobjdump -sass libdpb200.so, sm_100a)\n\n"
                 "Per kernel: instruction count, the Blackwell tensor/TMA/mbarrier mnemonics (`UTC*MMA` = tcgen05.mma, `UTMALDG` = "
                 "cp.async.bulk.tensor, `LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier, `UCGABAR` = cluster "
                 "barrier, `LDGMC.E.ADD.F32` = multimem.ld_reduce, the in-switch NVLS load-reduce) and system-scope memory operations (NVLink peer loads/stores and flags).  No `HMMA` (legacy mma.sync) "
                 "anywhere.\n\n| kernel | instrs | tcgen05 / TMA / mbarrier | system-scope (peer) ops |\n|---|---|---|---|\n")
        for d, n, tc, so in rows:
            fh.write(f"| `{d}` | {n} | {', '.join(f'{k} x{v}' for k, v in sorted(tc.items())) or '-'} | "
                     f"{', '.join(f'{k} x{v}' for k, v in sorted(so.items())) or '-'} |\n")
    print(f"{len(rows)} kernels summarised in {OUT}")


if __name__ == "__main__":
    main()
