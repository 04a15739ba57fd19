"""Attention microbench:  python tools/attn_bench.py B Nq Nk [self|cross] [iters]"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import _lib   # noqa: E402

B, Nq, Nk = (int(x) for x in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "self"
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 10
L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
out = torch.zeros(B * Nq, 512, device="cuda", dtype=torch.bfloat16)
if mode == "self":
    qkv = torch.randn(B * Nq, 1536, device="cuda").bfloat16()
    args = (qkv.data_ptr(), 1536, qkv.data_ptr() + 1024, 1536, qkv.data_ptr() + 2048, 1536)
    byts = qkv.numel() * 2 + out.numel() * 2
else:
    q = torch.randn(B * Nq, 512, device="cuda").bfloat16()
    kv = torch.randn(B * Nk, 6144, device="cuda").bfloat16()
    args = (q.data_ptr(), 512, kv.data_ptr(), 6144, kv.data_ptr() + 1024, 6144)
    byts = q.numel() * 2 + B * Nk * 1024 * 2 + out.numel() * 2


def run():
    _lib.check(L.pt_attention(*args, out.data_ptr(), 512, B, 8, 64, Nq, Nk, 0.125, torch.cuda.current_stream().cuda_stream))


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
fl = 4.0 * B * 8 * Nq * Nk * 64
print("attention B=%d Nq=%d Nk=%d %s: %.1f us  %.1f TFLOP/s  %.0f GB/s" % (B, Nq, Nk, mode, us, fl / us / 1e6, byts / us / 1e3))

# the same launch replayed from a CUDA graph (no host cost per launch)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(iters):
        run()
g.replay(); torch.cuda.synchronize()
e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
print("   graph replay: %.1f us per launch" % (e0.elapsed_time(e1) / iters * 1e3))
