"""Generate tests/golden/metrics_vectors.pt: LVD and diversity of the UNMODIFIED reference (evaluation/metrics.py:79-110)
on seeded inputs (TEST INFRASTRUCTURE; authoring container only: python -m oracle.make_golden_metrics)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH   # noqa: E402


def make_inputs(seed=51, B=6, T=120, J=22, D=3):
    g = torch.Generator().manual_seed(seed)
    gt = torch.cumsum(torch.randn(T, J, D, generator=g) * 0.05, 0)
    pr = gt[None] + torch.cumsum(torch.randn(B, T, J, D, generator=g) * 0.02, 1)
    return gt, pr


def main():
    RH.load_reference()
    from evaluation.metrics import LVD, diversity
    out = {"cases": []}
    for seed, (B, T, J, D) in ((51, (6, 120, 22, 3)), (52, (16, 300, 55, 3)), (53, (2, 7, 2, 6))):
        gt, pr = make_inputs(seed, B, T, J, D)
        out["cases"].append(dict(seed=seed, shape=(B, T, J, D), lvd=float(LVD(gt.clone(), pr.clone())),
                                 lvd_single=float(LVD(gt.clone(), pr[0].clone())),
                                 diversity=float(diversity(pr.reshape(B, T, J * D).numpy()))))
    path = os.path.join(ROOT, "tests", "golden", "metrics_vectors.pt")
    torch.save(out, path)
    print("wrote", path, out)


if __name__ == "__main__":
    main()
