"""PQ-search microbench driver (BASELINE config 2) used for ncu captures and quick timings."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import _lib   # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    layout = sys.argv[3] if len(sys.argv) > 3 else "cf"
    W, G = 22, 4
    torch.manual_seed(0)
    # 16 distinct z buffers (16 x 11.5 MB = 184 MB > the 126 MB L2) rotate, so that every launch reads its z from HBM
    zs = [torch.randn(B, 512, W, device="cuda") for _ in range(16)]
    if layout == "cl":
        zs = [z.permute(0, 2, 1).contiguous().permute(0, 2, 1) for z in zs]
    z = zs[0]
    n = [0]
    cb = ((torch.rand(G, 128, 128, device="cuda") * 2 - 1) * (6.0 / 256) ** 0.5).contiguous()
    idx = torch.empty(B, W, G, dtype=torch.int64, device="cuda")
    L = _lib.lib()
    sb, sc, sw = z.stride()
    st = torch.cuda.current_stream().cuda_stream

    def run():
        zz = zs[n[0] & 15]
        n[0] += 1
        rc = L.pt_pq_search(zz.data_ptr(), sb, sc, sw, B, W, G, 128, 128, cb.data_ptr(), idx.data_ptr(), None, None, None, st)
        assert rc == 0

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
    byts = z.numel() * 4 + cb.numel() * 4 + idx.numel() * 8
    print("pq_search B=%d layout=%s: %.2f us/launch, %.1f GB/s algorithmic (%.2f MB)" % (B, layout, us, byts / us / 1e3, byts / 1e6))
    # device time without host launch cost: replay a CUDA graph of `iters` launches
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st2 = torch.cuda.current_stream().cuda_stream
        for i in range(iters):
            assert L.pt_pq_search(zs[i & 15].data_ptr(), sb, sc, sw, B, W, G, 128, 128, cb.data_ptr(), idx.data_ptr(), None, None, None, st2) == 0
    g.replay(); torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    print("   graph replay: %.2f us/launch, %.1f GB/s algorithmic" % (us, byts / us / 1e3))


if __name__ == "__main__":
    main()
