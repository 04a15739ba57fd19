"""clock64 stamps of the tcgen05 attention kernel's CTA 0 (developer hook PT_ATTN_TRACE): python tools/attn_trace.py [B Nq Nk]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
Nq = int(sys.argv[2]) if len(sys.argv) > 2 else 88
Nk = int(sys.argv[3]) if len(sys.argv) > 3 else 88
H = 8
q = torch.randn(B * Nq, H * 64, device="cuda").bfloat16()
kv = torch.randn(B * Nk, 2 * H * 64, device="cuda").bfloat16()
trace = torch.zeros(5 * 16 * 4, dtype=torch.int64, device="cuda")
os.environ["PT_ATTN_TRACE"] = hex(trace.data_ptr())
for _ in range(3):
    trace.zero_()
    ops.attention(q, kv[:, :H * 64], kv[:, H * 64:], B, Nq, Nk, heads=H)
    torch.cuda.synchronize()
t = trace.cpu().view(5, 16, 4)
t0 = int(t[0, 0, 0])
f = lambda x: int(x) - t0 if int(x) else None
print("per unit of CTA 0:  producer [wait slot | slot free | loads issued]   S issuer [Q,K landed | issue | issued]   "
      "PV issuer [V landed | issue | issued]   softmax [wait S | S ready | P published]   output [wait O | O ready | stored]")
for u in range(12):
    print(u, " ".join(str([f(x) for x in t[r, u, :3]]) for r in range(5)))
