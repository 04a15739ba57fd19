"""clock64 stamps of the GEMM kernel's CTA 0 (developer hook PT_GEMM_TRACE): python tools/gemm_trace.py [rn2|rn1|plain] [M]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "rn2"
M = int(sys.argv[2]) if len(sys.argv) > 2 else 90112
a = torch.randn(M, 512, device="cuda").bfloat16()
w = (torch.randn(512, 512, device="cuda") / 512 ** 0.5).bfloat16()
b = torch.randn(512, device="cuda")
o = torch.randn(M, 512, device="cuda").bfloat16()
na, nb = torch.rand(4, 512, device="cuda") + 0.5, torch.randn(4, 512, device="cuda")
labels = (torch.arange((M + 87) // 88, device="cuda") % 4).int()
trace = torch.zeros(3 * 16 * 8, dtype=torch.int64, device="cuda")
os.environ["PT_GEMM_TRACE"] = hex(trace.data_ptr())


def run():
    if mode == "rn2":
        ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512, residual=o, rrs=512, row_norm=2, norm_a=na, norm_b=nb)
    elif mode == "rn1":
        ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512, row_norm=1, norm_a=na, norm_b=nb, labels=labels, label_rows=88)
    else:
        ops.gemm(a, w, o, M=M, C_=512, bias=b, ors=512)


for _ in range(3):
    trace.zero_()
    run()
    torch.cuda.synchronize()
t = trace.cpu().view(3, 16, 8)
t0 = int(t[0, 0, 0])
f = lambda x: int(x) - t0 if int(x) else None
print(mode, "M =", M)
print("per tile of CTA 0: producer [first load | last slot wait | loads issued]   issuer [wait acc | acc free | first operands | last operands | issued]"
      "   epilogue warp 0 [wait acc | acc ready | acc read | stats sent | stats here | stored | chunk 1 ready | chunk 1 tile free]")
for u in range(10):
    print(u, [f(x) for x in t[0, u, :3]], [f(x) for x in t[1, u, :5]], [f(x) for x in t[2, u, :8]])
