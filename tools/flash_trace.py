"""clock64 stamps of the streamed-key attention kernel's CTA 0 (developer hook PT_FLASH_TRACE): python tools/flash_trace.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

B, H, N = 8, 12, 2999
qkv = torch.randn(B * N, 3 * H * 64, device="cuda").bfloat16()
q, k, v = qkv[:, :H * 64], qkv[:, H * 64:2 * H * 64], qkv[:, 2 * H * 64:]
trace = torch.zeros(3 * 16 * 8, dtype=torch.int64, device="cuda")
os.environ["PT_FLASH_TRACE"] = hex(trace.data_ptr())
for _ in range(3):
    trace.zero_()
    ops.attention(q, k, v, B, N, N, heads=H)
    torch.cuda.synchronize()
t = trace.cpu().view(3, 16, 8)
t0 = int(t[0, 0, 0])
f = lambda x: int(x) - t0 if int(x) else None
print("softmax rows (tile A | tile B), per block: wait S | S ready | S read | exps done | P tile free | P published | PV(j-1) added")
for j in range(12):
    print(j, [f(x) for x in t[0, j, :7]], "|", [f(x) for x in t[1, j, :7]])
print("issuer, per block: S_A issued | S_B issued | PV_A issued | PV_B issued")
for j in range(12):
    print(j, [f(x) for x in t[2, j, :7]], int(t[2, j, 7]))
