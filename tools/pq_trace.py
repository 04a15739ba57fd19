"""Per-phase clock stamps of the PQ search kernel's CTA 0 (developer hook PT_PQ_TRACE):  python tools/pq_trace.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import _lib   # noqa: E402

L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
zs = [torch.randn(256, 22, 512, device="cuda") for _ in range(16)]
cb = ((torch.rand(4, 128, 128, device="cuda") * 2 - 1) * (6.0 / 256) ** 0.5).contiguous()
idx = torch.empty(256, 22, 4, dtype=torch.int64, device="cuda")
trace = torch.zeros(96, dtype=torch.int64, device="cuda")
os.environ["PT_PQ_TRACE"] = hex(trace.data_ptr())
for i in range(8):
    trace.zero_()
    _lib.check(L.pt_pq_search(zs[i].data_ptr(), 22 * 512, 1, 512, 256, 22, 4, 128, 128, cb.data_ptr(), idx.data_ptr(), None, None,
                              None, st))
    torch.cuda.synchronize()
t = trace.cpu().view(3, 32)
t0 = int(t[0, 0])
names = ["loader: start | set-up done | (rows landed, operands free, converted) per tile",
         "mma: set-up done | (ready, issued) per tile", "scan: set-up done | (scores ready, scanned, written) per tile"]
for r in range(3):
    print(names[r])
    print("   ", [int(x) - t0 for x in t[r] if int(x) != 0])
