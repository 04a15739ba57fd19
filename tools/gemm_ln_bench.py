"""Row-norm GEMM microbench (for ncu):  python tools/gemm_ln_bench.py M K mode [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

M, K, mode = (int(x) for x in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 10
N = 512
a = torch.randn(M, K, device="cuda").bfloat16()
w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
bias = torch.randn(N, device="cuda")
na = torch.rand(4, N, device="cuda") + 0.5
nb = torch.randn(4, N, device="cuda")
labels = (torch.arange((M + 87) // 88, device="cuda") % 4).int()
out = torch.randn(M, N, device="cuda").bfloat16()


def run():
    if mode == 1:
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N, row_norm=1, norm_a=na, norm_b=nb, labels=labels, label_rows=88)
    elif mode == 2:
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N, residual=out, rrs=N, row_norm=2, norm_a=na, norm_b=nb)
    else:
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    run()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / iters * 1e3
print("gemm_ln M=%d K=%d mode=%d: %.1f us = %.1f TFLOP/s" % (M, K, mode, us, 2.0 * M * N * K / us / 1e6))
