"""GPU bring-up / diagnostics for the tcgen05 kernels (developer tool, not a test).

Runs each case in its own subprocess with a timeout so that a device trap in one kernel cannot
take the others down:   python tools/gpu_bringup.py            (all cases)
                        python tools/gpu_bringup.py --case pq  (one case, in-process)
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _time(fn, iters=20, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(iters):
        fn()
    en.record()
    torch.cuda.synchronize()
    return st.elapsed_time(en) / iters * 1e3   # us


def case_pq():
    import torch
    from probtalk_b200 import ops
    torch.manual_seed(0)
    dev = "cuda"
    for scale, cbk in [(1.0, "xavier"), (0.1, "xavier"), (1.0, "normal")]:
        B, W, G, K, D = 256, 22, 4, 128, 128
        z = (torch.randn(B, G * D, W, device=dev) * scale)
        if cbk == "xavier":
            cb = (torch.rand(G, K, D, device=dev) * 2 - 1) * (6.0 / 256) ** 0.5
        else:
            cb = torch.randn(G, K, D, device=dev)
        idx, gap, dump, nres = ops.pq_search(z, cb, want_gap=True, score_dump=True, count_rescored=True)
        torch.cuda.synchronize()
        flat = z.permute(0, 2, 1).reshape(-1, G, D).double()
        mism_tot, maxerr, bound_violation = 0, 0.0, 0
        for g in range(G):
            x, e = flat[:, g], cb[g].double()
            sc = (e ** 2).sum(1)[None] - 2 * x @ e.t()
            err = (dump.view(-1, G, K)[:, g].double() - sc).abs().max(1).values
            bound = 2.5e-4 * x.norm(dim=1) * e.norm(dim=1).max()
            bound_violation += int((err > bound).sum())
            maxerr = max(maxerr, float((err / bound).max()))
            d = (x ** 2).sum(1, keepdim=True) + sc
            t2 = torch.topk(d, 2, dim=1, largest=False)
            ref = t2.indices[:, 0]
            mine = idx.view(-1, G)[:, g]
            bad = mine != ref
            mism_tot += int(bad.sum())
            if bad.any():
                g2 = (t2.values[:, 1] - t2.values[:, 0])[bad]
                print("   group %d: %d mismatches, fp64 gaps of mismatches: max %.3e" % (g, int(bad.sum()), float(g2.max())))
        print("pq scale=%s cb=%s: mismatches vs fp64 argmin = %d / %d, rescored rows = %d, "
              "max |tc score err| / bound = %.3f, bound violations = %d"
              % (scale, cbk, mism_tot, B * W * G, int(nres), maxerr, bound_violation))
        zc = z.permute(0, 2, 1).contiguous().permute(0, 2, 1)       # channels-last storage, same logical tensor
        idx2 = ops.pq_search(zc, cb)
        print("   channels-last path equal to channel-first path:", bool((idx2 == idx).all()))
    z = torch.randn(256, 512, 22, device=dev)
    cb = (torch.rand(4, 128, 128, device=dev) * 2 - 1) * (6.0 / 256) ** 0.5
    us = _time(lambda: ops.pq_search(z, cb), iters=50)
    print("pq_search 256x22x4: %.2f us  -> %.1f GB/s algorithmic (11.97 MB)" % (us, 11.97e6 / us / 1e3))
    z = torch.randn(4096, 512, 22, device=dev)
    us = _time(lambda: ops.pq_search(z, cb), iters=20)
    print("pq_search 4096x22x4: %.2f us -> %.1f GB/s" % (us, (z.numel() * 4 + 4096 * 22 * 4 * 8) / us / 1e3))


def _gemm_ref(a, w, bias):
    return a.float() @ w.float().t() + (bias if bias is not None else 0)


def case_gemm_plain():
    import torch
    from probtalk_b200 import ops
    torch.manual_seed(1)
    dev = "cuda"
    for (M, N, K) in [(128, 128, 64), (128, 128, 256), (1000, 256, 512), (5000, 1536, 512), (333, 384, 1024)]:
        a = torch.randn(M, K, device=dev).bfloat16()
        w = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
        bias = torch.randn(N, device=dev)
        for out_dtype in (torch.float32, torch.bfloat16):
            out = torch.full((M, N), float("nan"), device=dev, dtype=out_dtype)
            ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N)
            torch.cuda.synchronize()
            ref = _gemm_ref(a, w, bias)
            err = (out.float() - ref).abs().max().item()
            print("gemm %s M=%d N=%d K=%d: max abs err %.4g (ref max %.3g) nan=%d"
                  % (str(out_dtype)[6:], M, N, K, err, ref.abs().max().item(), int(torch.isnan(out.float()).sum())))
            if not (err < 0.05):
                bad = (out.float() - ref).abs() > 0.05
                rows = bad.any(1).nonzero().flatten()[:8].tolist()
                cols = bad.any(0).nonzero().flatten()[:8].tolist()
                print("   first bad rows", rows, "cols", cols, "frac bad %.4f" % bad.float().mean().item())
                print("   out[0,:8]", out[0, :8].float().tolist())
                print("   ref[0,:8]", ref[0, :8].tolist())


def case_gemm_conv():
    import torch
    import torch.nn.functional as F
    from probtalk_b200 import ops
    torch.manual_seed(2)
    dev = "cuda"
    B, T, Cin, Cout = 5, 45, 128, 256
    H = 1
    # ---- k3 s1 p1
    Tp = T + 2 * H
    x = torch.randn(B, Cin, T, device=dev)
    xb = torch.zeros(B, Tp, Cin, device=dev, dtype=torch.bfloat16)
    xb[:, H:H + T] = x.transpose(1, 2).bfloat16()
    wt = torch.randn(Cout, Cin, 3, device=dev) / (3 * Cin) ** 0.5
    wk = wt.permute(0, 2, 1).reshape(Cout, 3 * Cin).contiguous().bfloat16()       # [N, tap*C + c]
    bias = torch.randn(Cout, device=dev)
    out = torch.zeros(B, Tp, Cout, device=dev, dtype=torch.bfloat16)
    ops.gemm(xb.view(-1), wk, out.view(-1), M=B * Tp, C_=Cin, taps=3, s=1, base_row=H - 1, rows_per_sample=Tp,
             t_valid=T, oss=Tp * Cout, ors=Cout, out_offset=H * Cout, bias=bias, act=ops.ACT_LEAKY, slope=0.2)
    torch.cuda.synchronize()
    ref = F.leaky_relu(F.conv1d(xb[:, H:H + T].float().transpose(1, 2), wk.float().view(Cout, 3, Cin).permute(0, 2, 1), bias, padding=1), 0.2)
    err = (out[:, H:H + T].float().transpose(1, 2) - ref).abs().max().item()
    halo = out[:, :H].abs().max().item() + out[:, H + T:].abs().max().item()
    print("conv k3s1: max err %.4g (ref max %.3g), halo rows untouched: %s" % (err, ref.abs().max().item(), halo == 0))
    # ---- k4 s2 p1  (Tp even)
    Tp2 = T + 2 * H + ((T + 2 * H) % 2)
    xb2 = torch.zeros(B, Tp2, Cin, device=dev, dtype=torch.bfloat16)
    xb2[:, H:H + T] = x.transpose(1, 2).bfloat16()
    w4 = torch.randn(Cout, Cin, 4, device=dev) / (4 * Cin) ** 0.5
    wk4 = w4.permute(0, 2, 1).reshape(Cout, 4 * Cin).contiguous().bfloat16()
    To = (T - 2) // 2 + 1
    Tpo = To + 2
    out2 = torch.zeros(B, Tpo, Cout, device=dev, dtype=torch.float32)
    ops.gemm(xb2.view(-1), wk4, out2.view(-1), M=B * (Tp2 // 2), C_=Cin, taps=4, s=2, base_row=0,
             rows_per_sample=Tp2 // 2, t_valid=To, oss=Tpo * Cout, ors=Cout, out_offset=Cout, bias=bias)
    torch.cuda.synchronize()
    ref2 = F.conv1d(xb2[:, H:H + T].float().transpose(1, 2), wk4.float().view(Cout, 4, Cin).permute(0, 2, 1), bias, stride=2, padding=1)
    err2 = (out2[:, 1:1 + To].transpose(1, 2) - ref2).abs().max().item()
    print("conv k4s2: max err %.4g (ref max %.3g) To=%d" % (err2, ref2.abs().max().item(), To))
    # ---- transposed k4 s2 p1: two phase launches, channel-first fp32 output
    wtT = torch.randn(Cin, Cout, 4, device=dev) / (2 * Cin) ** 0.5
    wtb = wtT.bfloat16().float()
    outT = torch.zeros(B, Cout, 2 * T, device=dev, dtype=torch.float32)
    for ph in (0, 1):
        taps = (3, 1) if ph == 0 else (2, 0)
        wk2 = torch.cat([wtb[:, :, taps[0]].t(), wtb[:, :, taps[1]].t()], dim=1).contiguous().bfloat16()   # [Cout, 2*Cin]
        ops.gemm(xb.view(-1), wk2, outT.view(-1), M=B * Tp, C_=Cin, taps=2, s=1, base_row=H - 1 + ph,
                 rows_per_sample=Tp, t_valid=T, oss=Cout * 2 * T, ors=2, ocs=2 * T, out_offset=ph, bias=bias)
    torch.cuda.synchronize()
    refT = F.conv_transpose1d(xb[:, H:H + T].float().transpose(1, 2), wtb, bias, stride=2, padding=1)
    errT = (outT - refT).abs().max().item()
    print("convT k4s2 (channel-first fp32 out): max err %.4g (ref max %.3g)" % (errT, refT.abs().max().item()))
    # ---- gated epilogue + residual
    M, Cc = 700, 256
    a = torch.randn(M, Cc, device=dev).bfloat16()
    w = (torch.randn(2 * Cc, Cc, device=dev) / Cc ** 0.5)
    wi = torch.stack([w[:Cc], w[Cc:]], dim=1).reshape(2 * Cc, Cc).contiguous().bfloat16()     # interleave a/b rows
    b2 = torch.randn(2 * Cc, device=dev)
    bi = torch.stack([b2[:Cc], b2[Cc:]], dim=1).reshape(-1).contiguous()
    og = torch.zeros(M, Cc, device=dev, dtype=torch.bfloat16)
    ops.gemm(a, wi, og, M=M, C_=Cc, bias=bi, act=ops.ACT_GATE, ors=Cc)
    h = a.float() @ w.bfloat16().float().t() + b2
    refg = torch.tanh(h[:, :Cc]) * torch.sigmoid(h[:, Cc:])
    torch.cuda.synchronize()
    print("gated epilogue: max err %.4g" % (og.float() - refg).abs().max().item())
    res = torch.randn(M, Cc, device=dev).bfloat16()
    w1 = (torch.randn(Cc, Cc, device=dev) / Cc ** 0.5).bfloat16()
    orr = torch.zeros(M, Cc, device=dev, dtype=torch.bfloat16)
    ops.gemm(a, w1, orr, M=M, C_=Cc, residual=res, rss=0, rrs=Cc, act=ops.ACT_RELU, ors=Cc)
    torch.cuda.synchronize()
    refr = torch.relu(a.float() @ w1.float().t() + res.float())
    print("residual+relu epilogue: max err %.4g" % (orr.float() - refr).abs().max().item())


def case_gemm_perf():
    import torch
    from probtalk_b200 import ops
    dev = "cuda"
    for (M, N, K) in [(90112, 512, 512), (90112, 1536, 512), (90112, 2048, 512), (90112, 512, 2048),
                      (186368, 256, 768), (23552, 1024, 3072), (8192, 8192, 8192)]:
        a = torch.randn(M, K, device=dev).bfloat16()
        w = torch.randn(N, K, device=dev).bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        us = _time(lambda: ops.gemm(a, w, out, M=M, C_=K, ors=N), iters=10)
        us_t = _time(lambda: torch.matmul(a, w.t()), iters=10)
        print("gemm M=%d N=%d K=%d: %.1f us = %.1f TFLOP/s   (torch.matmul %.1f us = %.1f TFLOP/s)"
              % (M, N, K, us, 2.0 * M * N * K / us / 1e6, us_t, 2.0 * M * N * K / us_t / 1e6))


CASES = {"pq": case_pq, "gemm_plain": case_gemm_plain, "gemm_conv": case_gemm_conv, "gemm_perf": case_gemm_perf}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default=None)
    ap.add_argument("--timeout", type=int, default=240)
    args = ap.parse_args()
    if args.case:
        CASES[args.case]()
        sys.exit(0)
    rc = 0
    for name in CASES:
        t0 = time.time()
        print("=== case %s" % name, flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", name], timeout=args.timeout,
                               capture_output=True, text=True)
            print(r.stdout[-6000:])
            if r.returncode != 0:
                rc = 1
                print("!!! case %s exited %d\n%s" % (name, r.returncode, r.stderr[-3000:]))
        except subprocess.TimeoutExpired as e:
            rc = 1
            print("!!! case %s timed out\n%s" % (name, (e.stdout or b"")[-3000:]))
        print("=== case %s done in %.1fs" % (name, time.time() - t0), flush=True)
    sys.exit(rc)
