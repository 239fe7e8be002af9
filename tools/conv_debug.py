"""GPU bring-up aid for the tcgen05 conv: runs each (mode, nt, splits) case in its own process (a trapped kernel
poisons the CUDA context) and prints error structure, not just pass/fail.
usage: python tools/conv_debug.py            # all cases, each in a subprocess
       python tools/conv_debug.py one MODE NT SPLITS N H CIN COUT"""
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(mode, nt, splits, N, H, Cin, Cout):
    import torch
    import torch.nn.functional as F
    from distributed_pytorch_b200 import _native
    from distributed_pytorch_b200.ops import kernels as K

    def tf32(t):
        return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)

    def nhwc(t):
        return t.contiguous(memory_format=torch.channels_last)

    torch.manual_seed(0)
    x = tf32(torch.randn(N, Cin, H, H))
    w = tf32(torch.randn(Cout, Cin, 3, 3) * 0.1).requires_grad_(True)
    b = torch.randn(Cout)
    xr = x.clone().requires_grad_(True)
    y = F.conv2d(xr, w, b, padding=1)
    dy = tf32(torch.randn_like(y))
    y.backward(dy)
    xd, wd, bd, dyd = nhwc(x.cuda()), nhwc(w.detach().cuda()), b.cuda(), nhwc(dy.cuda())
    res = {"mode": mode, "nt": nt, "splits": splits, "shape": [N, H, Cin, Cout]}
    dump = torch.zeros(48 * 1024 // 4, device="cuda")
    _native.require().dpb_debug_set_conv_dump(dump.data_ptr())
    try:
        if mode == 0:
            out = K.nhwc_empty(N, Cout, H, H, "cuda"); out.fill_(float("nan"))
            K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, out, bd, nt=nt, splits=splits)
            ref = y.detach()
        elif mode == 1:
            out = K.nhwc_empty(N, Cin, H, H, "cuda"); out.fill_(float("nan"))
            K.conv3x3_gemm(1, N, H, H, Cin, Cout, dyd, None, wd, out, nt=nt, splits=splits)
            ref = xr.grad
        else:
            out = torch.empty_like(wd); out.fill_(float("nan"))
            K.conv3x3_gemm(2, N, H, H, Cin, Cout, dyd, xd, wd, out, nt=nt, splits=splits)
            ref = w.grad
        torch.cuda.synchronize()
        o = out.cpu()
        diff = (o - ref).abs()
        bad = ~(diff <= 2e-4 * ref.abs().max() + 1e-6)
        res.update(ok=bool(not bad.any()), max_err=float(torch.nan_to_num(diff, nan=1e30).max()),
                   scale=float(ref.abs().max()), frac_bad=float(bad.float().mean()), nan=int(torch.isnan(o).sum()))
        d = dump.cpu()
        a_part, b_part = d[:4096], d[4096:]
        res["dump_nonzero"] = {"A": int((a_part != 0).sum()), "B": int((b_part != 0).sum())}

        def unswizzle(atom):  # atom: [rows][32] floats written by TMA with SWIZZLE_128B_ATOM_32B (MN-major operands)
            rows = atom.shape[0]
            out_ = torch.empty_like(atom)
            for r_ in range(rows):
                for j in range(4):
                    out_[r_, 8 * j:8 * j + 8] = atom[r_, 8 * (j ^ (r_ % 4)):8 * (j ^ (r_ % 4)) + 8]
            return out_
        if mode == 1:  # first k-iter: tap'=0 -> weight tap 8 (r=2,s=2), co 0..31, ci 0..31
            exp = w.detach()[0:32, 0:32, 2, 2]            # [co][ci]
            got = unswizzle(b_part[:1024].view(32, 32))
            res["B_atom0_matches"] = bool(torch.equal(got, exp))
            res["B_atom0_linear_matches"] = bool(torch.equal(b_part[:1024].view(32, 32), exp))
        if mode == 2:  # tap 0 (r=0,s=0), pixels 0..31, co 0..31 ; B: shifted pixels x ci 0..31
            dyp = dy.permute(0, 2, 3, 1).reshape(-1, Cout)
            got = unswizzle(a_part[:1024].view(32, 32))
            res["A_atom0_matches"] = bool(torch.equal(got, dyp[0:32, 0:32]))
        if bad.any():
            # structure: which output channels / pixels are wrong (logical NCHW or OIHW index)
            idx = bad.nonzero()
            res["bad_dim_hist"] = {f"dim{d}": sorted(set(idx[:, d].tolist()))[:16] for d in range(idx.shape[1])}
            # ratio out/ref on bad entries to spot scaling / partial sums
            r = (o[bad] / ref[bad])
            r = r[torch.isfinite(r)]
            if r.numel():
                res["ratio_median"] = float(r.median())
            res["sample"] = [[float(o.flatten()[i]), float(ref.flatten()[i])] for i in range(4)]
    except Exception as e:  # noqa: BLE001
        res.update(ok=False, error=str(e)[:400])
        try:
            import ctypes
            res["device_error"] = _native.device_error_message()
        except Exception as e2:  # noqa: BLE001
            res["device_error"] = "unreadable: " + str(e2)[:100]
    print("RESULT " + json.dumps(res))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        one(*[int(a) for a in sys.argv[2:9]])
        return
    cases = []
    for mode in (0, 1, 2):
        cases += [(mode, 64, 1, 2, 8, 64, 64), (mode, 128, 1, 2, 8, 128, 128), (mode, 256, 1, 2, 8, 256, 256),
                  (mode, 128, 3, 4, 8, 128, 256), (mode, 0, 0, 2, 16, 64, 128), (mode, 0, 0, 32, 2, 512, 512),
                  (mode, 0, 0, 3, 4, 256, 512)]
    out = []
    for c in cases:
        try:
            p = subprocess.run([sys.executable, __file__, "one", *map(str, c)], capture_output=True, text=True, timeout=120)
            line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            rec = json.loads(line[-1][7:]) if line else {"case": c, "ok": False, "rc": p.returncode, "stderr": p.stderr[-600:]}
        except subprocess.TimeoutExpired:
            rec = {"case": c, "ok": False, "error": "timeout (hang)"}
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/conv_debug.json", "w") as fh:
        json.dump(out, fh, indent=1)
    print(f"{sum(1 for r in out if r.get('ok'))}/{len(out)} conv cases ok")


if __name__ == "__main__":
    main()
