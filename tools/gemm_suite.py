"""The transformer-stack GEMM shapes of one bench step, two launches each (for one ncu --set full capture):
   python tools/gemm_suite.py [M]      launch order: rn2 K512 | rn1 K512 | plain N512 K512 | relu N2048 K512 | rn1 K2048 | N1536 K512"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 90112
dev = "cuda"


def mk(N, K):
    return (torch.randn(M, K, device=dev).bfloat16(), (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16(),
            torch.randn(N, device=dev), torch.randn(M, N, device=dev).bfloat16())


na = torch.rand(4, 512, device=dev) + 0.5
nb = torch.randn(4, 512, device=dev)
labels = (torch.arange((M + 87) // 88, device=dev) % 4).int()
cases = []
a, w, b, o = mk(512, 512)
cases.append(("rn2 N512 K512", 512, 512, lambda a=a, w=w, b=b, o=o: ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512, residual=o, rrs=512, row_norm=2, norm_a=na, norm_b=nb)))
cases.append(("rn1 N512 K512", 512, 512, lambda a=a, w=w, b=b, o=o: ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512, row_norm=1, norm_a=na, norm_b=nb, labels=labels, label_rows=88)))
cases.append(("plain N512 K512", 512, 512, lambda a=a, w=w, b=b, o=o: ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512)))
a2, w2, b2, o2 = mk(2048, 512)
cases.append(("relu N2048 K512", 2048, 512, lambda: ops.gemm(a2, w2, o2, M=M, C_=512, bias=b2, ors=2048, act=ops.ACT_RELU)))
a3, w3, b3, o3 = mk(512, 2048)
cases.append(("rn1 N512 K2048", 512, 2048, lambda: ops.gemm(a3, w3, o3, M=M, C_=2048, bias=b3, ors=512, row_norm=1, norm_a=na, norm_b=nb, labels=labels, label_rows=88)))
a4, w4, b4, o4 = mk(1536, 512)
cases.append(("plain N1536 K512", 1536, 512, lambda: ops.gemm(a4, w4, o4, M=M, C_=512, bias=b4, ors=1536)))
iters = int(os.environ.get("ITERS", "2"))
for name, N, K, fn in cases:
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters - 1):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / max(iters - 1, 1) * 1e3
    print("%-18s %8.1f us  %7.1f TFLOP/s" % (name, us, 2.0 * M * N * K / us / 1e6), flush=True)
