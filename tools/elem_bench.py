"""Microbench of the layout kernels at bench shapes: python tools/elem_bench.py [B]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import _lib   # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
C_, T = 376, 180
L = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
src = torch.randn(B, C_, T, device="cuda"); pred = torch.randn(B, C_, T, device="cuda")
mask = (torch.rand(B, 1, T, device="cuda") < 0.5).float()
dst = torch.zeros(B, T + 8, 384, device="cuda", dtype=torch.bfloat16)
out = torch.zeros(B, C_, T, device="cuda")


def timeit(name, fn, byts):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    print("%-28s %8.1f us  %7.0f GB/s" % (name, us, byts / us / 1e3), flush=True)


n = B * C_ * T
for mode in (0, 1, 2):
    timeit("pack_rows mode %d" % mode, lambda: _lib.check(L.pt_pack_rows(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), T, dst.data_ptr(), None, B, C_, T, T + 8, 4, 384, 0, mode, st)),
           n * 4 * (2 if mode == 2 else 1) + n * 2)
timeit("interp_blend pow 1", lambda: _lib.check(L.pt_interp_blend(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), T, out.data_ptr(), B, C_, T, 1, st)), n * 12)
timeit("interp_blend pow 2 (Tpred 176)", lambda: _lib.check(L.pt_interp_blend(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), 176, out.data_ptr(), B, C_, T, 2, st)), n * 12)
