"""Single-shape GEMM microbench (for ncu captures):  python tools/gemm_bench.py M N K [iters] [out=bf16|f32] [res=0|1]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

M, N, K = (int(x) for x in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 10
odt = torch.float32 if (len(sys.argv) > 5 and sys.argv[5] == "f32") else torch.bfloat16
use_res = len(sys.argv) > 6 and sys.argv[6] == "1"
a = torch.randn(M, K, device="cuda").bfloat16()
w = torch.randn(N, K, device="cuda").bfloat16()
bias = torch.randn(N, device="cuda")
out = torch.empty(M, N, device="cuda", dtype=odt)
res = torch.randn(M, N, device="cuda") if use_res else None


def run():
    ops.gemm(a, w, out, M=M, C_=K, ors=N, bias=bias, residual=res, rrs=N)


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
print("gemm M=%d N=%d K=%d out=%s res=%d: %.1f us = %.1f TFLOP/s" % (M, N, K, str(odt)[6:], use_res, us, 2.0 * M * N * K / us / 1e6))
