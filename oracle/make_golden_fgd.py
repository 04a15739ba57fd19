"""Generate tests/golden/fgd_vectors.pt by running the UNMODIFIED reference FGD evaluator on CPU.

TEST INFRASTRUCTURE; run in the authoring container only (needs /root/reference):
    python -m oracle.make_golden_fgd
Stores a seeded EmbeddingNet state_dict (BatchNorm statistics and affines randomised so that folding is exercised),
small pose batches, the reference's features (EmbeddingNet.extract) and the reference's scores
(EmbeddingSpaceEvaluator.push_samples / get_scores with the same net in all three slots).
"""
import contextlib
import io
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH   # noqa: E402


def make_inputs(n, dim, seed=33):
    """The pose batches of the golden file: re-created from the seed by the tests (40 MB would not belong in git)."""
    g = torch.Generator().manual_seed(seed)
    real = torch.randn(n, dim, 90, generator=g) * 1.2 + 0.3
    gene = real + 0.35 * torch.randn(n, dim, 90, generator=g)
    return real, gene


def main():
    RH.load_reference()
    # compatibility shim (scipy >= 1.18 dropped sqrtm's `disp` argument that FGD.py:228 passes): same routine, old signature
    from scipy import linalg as _la
    _sqrtm = _la.sqrtm

    def sqrtm_compat(a, disp=True, blocksize=64):
        r = _sqrtm(a)
        return r if disp else (r, 0.0)

    _la.sqrtm = sqrtm_compat
    from evaluation.FGD import EmbeddingSpaceEvaluator
    from nets.embedding_net import EmbeddingNet
    torch.manual_seed(31)
    dim = 376
    net = EmbeddingNet(None, dim, 90).eval()
    g = torch.Generator().manual_seed(32)
    with torch.no_grad():
        for name, buf in net.named_buffers():
            if name.endswith("running_mean"):
                buf.copy_(torch.randn(buf.shape, generator=g) * 0.2)
            elif name.endswith("running_var"):
                buf.copy_(torch.rand(buf.shape, generator=g) + 0.5)
        for name, prm in net.named_parameters():
            if ".1.weight" in name or ".4.weight" in name:          # BatchNorm affines
                if prm.dim() == 1:
                    prm.copy_(torch.rand(prm.shape, generator=g) + 0.5)
            if name.startswith("pose_encoder.out_net.") and prm.dim() == 2 or name.startswith("pose_encoder.fc_mu.weight"):
                prm.mul_(3.0)                                        # O(1) features instead of the ~1e-2 of a fresh init
    sd = {k: v.clone() for k, v in net.state_dict().items() if k.startswith("pose_encoder.")}
    n = 300                                                          # > 128 so that the covariances are full rank
    real, gene = make_inputs(n, dim)
    with torch.no_grad():
        f_real = net.extract(real)[0].transpose(0, 1).contiguous()   # extract returns feat.transpose(0, 1)
        f_gene = net.extract(gene)[0].transpose(0, 1).contiguous()
        ev = EmbeddingSpaceEvaluator(net, net, net, None, "cpu")
        for i in range(0, n, 100):
            ev.push_samples(gene[i:i + 100], real[i:i + 100])
        with contextlib.redirect_stdout(io.StringIO()):
            fgd, feat_dist = ev.get_scores("fbhe")
    out = {"torch_version": str(torch.__version__), "state_dict": sd, "n": n, "dim": dim, "input_seed": 33,
           "real_probe": real[:2, :4, :5].clone(), "gene_probe": gene[:2, :4, :5].clone(),
           "feat_real": f_real, "feat_gene": f_gene, "fgd": float(fgd), "feat_dist": float(feat_dist)}
    path = os.path.join(ROOT, "tests", "golden", "fgd_vectors.pt")
    torch.save(out, path)
    print("wrote", path, "fgd", fgd, "feat_dist", feat_dist, f_real.shape)


if __name__ == "__main__":
    main()
