"""pt_attention microbench (long sequences): python tools/attn_bench.py [B heads N iters]; prints us per launch and TF/s."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import ops   # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    N = int(sys.argv[3]) if len(sys.argv) > 3 else 2999
    iters = int(sys.argv[4]) if len(sys.argv) > 4 else 10
    g = torch.Generator(device="cuda").manual_seed(0)
    qkv = torch.randn(B * N, 3 * H * 64, device="cuda", generator=g).bfloat16()
    q, k, v = qkv[:, :H * 64], qkv[:, H * 64:2 * H * 64], qkv[:, 2 * H * 64:]
    for _ in range(3):
        ops.attention(q, k, v, B, N, N, heads=H)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.attention(q, k, v, B, N, N, heads=H)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    print("pt_attention B=%d heads=%d N=%d: %.1f us  %.0f TF/s" % (B, H, N, us, 4.0 * B * H * N * N * 64 / us / 1e6))


if __name__ == "__main__":
    main()
