"""Launches for the round-2 ncu captures (one process, few launches):  python tools/prof_suite.py
   PQ search config 2 (channels-last rows, rotating z buffers), the three bench attention shapes, J1 search (K=2048, D=64)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import _lib   # noqa: E402

L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
B = 1024
# ---- PQ search, BASELINE config 2: 256 clips x 22 steps x 4 groups
zs = [torch.randn(256, 22, 512, device="cuda") for _ in range(16)]
cb = ((torch.rand(4, 128, 128, device="cuda") * 2 - 1) * (6.0 / 256) ** 0.5).contiguous()
idx = torch.empty(256, 22, 4, dtype=torch.int64, device="cuda")
for i in range(6):
    _lib.check(L.pt_pq_search(zs[i].data_ptr(), 22 * 512, 1, 512, 256, 22, 4, 128, 128, cb.data_ptr(), idx.data_ptr(), None, None,
                              None, st))
# ---- J1: one codebook of 2048 x 64 (VectorQuantizerEMA), 256 x 22 queries
z1 = torch.randn(256, 22, 64, device="cuda")
cb1 = (torch.randn(1, 2048, 64, device="cuda") * 0.3).contiguous()
idx1 = torch.empty(256, 22, 1, dtype=torch.int64, device="cuda")
for i in range(2):
    _lib.check(L.pt_pq_search(z1.data_ptr(), 22 * 64, 1, 64, 256, 22, 1, 2048, 64, cb1.data_ptr(), idx1.data_ptr(), None, None,
                              None, st))
# ---- attention: self 88 x 88, cross 88 x 22 (predictor), self 180 x 180 (refiner), 1024 clips x 8 heads
for Nq, Nk, mode in ((88, 88, "self"), (88, 22, "cross"), (180, 180, "self")):
    out = torch.zeros(B * Nq, 512, device="cuda", dtype=torch.bfloat16)
    if mode == "self":
        qkv = torch.randn(B * Nq, 1536, device="cuda").bfloat16()
        args = (qkv.data_ptr(), 1536, qkv.data_ptr() + 1024, 1536, qkv.data_ptr() + 2048, 1536)
    else:
        q = torch.randn(B * Nq, 512, device="cuda").bfloat16()
        kv = torch.randn(B * Nk, 6144, device="cuda").bfloat16()
        args = (q.data_ptr(), 512, kv.data_ptr(), 6144, kv.data_ptr() + 1024, 6144)
    for _ in range(2):
        _lib.check(L.pt_attention(*args, out.data_ptr(), 512, B, 8, 64, Nq, Nk, 0.125, st))
torch.cuda.synchronize()
print("prof_suite ok")
