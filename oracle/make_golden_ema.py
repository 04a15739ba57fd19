"""Generate tests/golden/ema_vectors.pt: three consecutive TRAINING-mode forwards of the UNMODIFIED reference's
ProductQuantization (nets/spg/vqvae_modules.py:437-545) on seeded inputs -- codebooks, EMA state and loss after every step
(TEST INFRASTRUCTURE; authoring container only: python -m oracle.make_golden_ema)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH   # noqa: E402

CASES = [dict(seed=61, B=8, W=22, G=4, K=128, D=128, steps=3, scale=0.3),
         dict(seed=62, B=96, W=22, G=4, K=128, D=128, steps=2, scale=1.0),
         dict(seed=63, B=40, W=9, G=2, K=64, D=64, steps=3, scale=0.5)]


def make_batches(c):
    g = torch.Generator().manual_seed(c["seed"])
    cb = [(torch.rand(c["K"], c["D"], generator=g) * 2 - 1) * (6.0 / (c["K"] + c["D"])) ** 0.5 for _ in range(c["G"])]
    zs = [torch.randn(c["B"], c["G"] * c["D"], c["W"], generator=g) * c["scale"] for _ in range(c["steps"])]
    return cb, zs


def main():
    RH.load_reference()
    from nets.spg.vqvae_modules import ProductQuantization
    out = {"cases": []}
    for c in CASES:
        cb, zs = make_batches(c)
        pq = ProductQuantization(c["G"] * c["D"], c["K"], 0.25, 0.99, num_chunks=c["G"])
        for g in range(c["G"]):
            getattr(pq, "embeddings_%d" % g).copy_(cb[g])
        pq.train()
        steps = []
        for z in zs:
            with torch.no_grad():
                idx = torch.stack([pq.get_code_indices(z.permute(0, 2, 1).reshape(-1, c["G"], c["D"])[:, g], g)
                                   for g in range(c["G"])], dim=1).view(c["B"], c["W"], c["G"])
            qo = pq(z)                                                  # training forward: updates the buffers
            steps.append(dict(idx=idx.to(torch.int16), loss=float(qo.loss),
                              codebooks=[getattr(pq, "embeddings_%d" % g).clone() for g in range(c["G"])],
                              ema_dw=[pq.ema_dw[g].hidden.clone() for g in range(c["G"])],
                              ema_cs=[pq.ema_cluster_size[g].hidden.clone() for g in range(c["G"])]))
        out["cases"].append(dict(cfg=c, steps=steps))
    path = os.path.join(ROOT, "tests", "golden", "ema_vectors.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
