"""Frechet gesture distance evaluator, mirroring evaluation/FGD.py's EmbeddingSpaceEvaluator.

* feature extraction: ``EmbeddingPlan`` (engine.py) runs PoseEncoderConv (nets/embedding_net.py:56-95) on the GEMM engine;
* statistics: ``feature_stats`` -> ``pt_feature_stats`` (fp64 mean / covariance, deterministic);
* distance: ``frechet_distance`` = calculate_frechet_distance (evaluation/FGD.py:190-257) on the host -- a 128 x 128
  matrix square root, host logic like in the reference.
CUDA tensors only; the library must be present (no CPU fallback).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


def feature_stats(feats: torch.Tensor):
    """feats [N, D] fp32 CUDA -> (mean [D], cov [D, D]) fp64 CUDA tensors; np.mean(axis=0) / np.cov(rowvar=False)."""
    if not feats.is_cuda:
        raise _lib.PtError("probtalk_b200.evaluation needs CUDA tensors (no CPU fallback)")
    assert feats.dim() == 2 and feats.dtype == torch.float32
    feats = feats.contiguous()
    N, D = feats.shape
    L = _lib.lib()
    nbytes = L.pt_feature_stats_workspace_bytes(D)
    if nbytes < 0:
        raise _lib.PtError("pt_feature_stats: feature dimension %d is not supported" % D)
    ws = torch.empty(nbytes // 8, dtype=torch.float64, device=feats.device)
    mean = torch.empty(D, dtype=torch.float64, device=feats.device)
    cov = torch.empty(D, D, dtype=torch.float64, device=feats.device)
    _lib.check(L.pt_feature_stats(feats.data_ptr(), N, D, mean.data_ptr(), cov.data_ptr(), ws.data_ptr(), nbytes,
                                  torch.cuda.current_stream().cuda_stream), "pt_feature_stats")
    return mean, cov


def lvd(gt_kps: torch.Tensor, pr_kps: torch.Tensor):
    """Landmark velocity difference, LVD(gt [T, J, D], pr [B, T, J, D]) of evaluation/metrics.py:79-94 (Batch_LVD branch,
    symmetrical = False, weight = False; the longer of the two is cut to the shorter).  Returns (mean float, per-sample
    fp64 CUDA tensor [B])."""
    if not (gt_kps.is_cuda and pr_kps.is_cuda):
        raise _lib.PtError("probtalk_b200.evaluation needs CUDA tensors (no CPU fallback)")
    if pr_kps.dim() == 3:
        pr_kps = pr_kps.unsqueeze(0)
    T = min(gt_kps.shape[0], pr_kps.shape[1])
    gt = gt_kps[:T].float().contiguous()
    pr = pr_kps[:, :T].float().contiguous()
    B, _, J, D = pr.shape
    assert gt.shape == (T, J, D)
    per = torch.empty(B, dtype=torch.float64, device=gt.device)
    mean = torch.empty(1, dtype=torch.float64, device=gt.device)
    _lib.check(_lib.lib().pt_lvd(gt.data_ptr(), pr.data_ptr(), B, T, J, D, per.data_ptr(), mean.data_ptr(),
                                 torch.cuda.current_stream().cuda_stream), "pt_lvd")
    return float(mean), per


def diversity(kps: torch.Tensor):
    """Mean pairwise L1 distance between the B samples kps [B, seq, dim] (evaluation/metrics.py:96-110).
    Returns (mean float, per-pair fp64 CUDA tensor [B (B - 1) / 2], pairs i < j with i outer)."""
    if not kps.is_cuda:
        raise _lib.PtError("probtalk_b200.evaluation needs CUDA tensors (no CPU fallback)")
    B = kps.shape[0]
    x = kps.float().reshape(B, -1).contiguous()
    per = torch.empty(B * (B - 1) // 2, dtype=torch.float64, device=x.device)
    mean = torch.empty(1, dtype=torch.float64, device=x.device)
    _lib.check(_lib.lib().pt_diversity(x.data_ptr(), B, x.shape[1], per.data_ptr(), mean.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream), "pt_diversity")
    return float(mean), per


def frechet_distance(mu1, sigma1, mu2, sigma2, eps=1e-6):
    """d^2 = |mu1 - mu2|^2 + Tr(S1 + S2 - 2 sqrt(S1 S2)), same steps and fallbacks as evaluation/FGD.py:190-257."""
    from scipy import linalg
    mu1, mu2 = np.atleast_1d(np.asarray(mu1, dtype=np.float64)), np.atleast_1d(np.asarray(mu2, dtype=np.float64))
    sigma1, sigma2 = np.atleast_2d(np.asarray(sigma1, dtype=np.float64)), np.atleast_2d(np.asarray(sigma2, dtype=np.float64))
    assert mu1.shape == mu2.shape and sigma1.shape == sigma2.shape
    diff = mu1 - mu2
    covmean = linalg.sqrtm(sigma1.dot(sigma2))
    if isinstance(covmean, tuple):
        covmean = covmean[0]
    if not np.isfinite(covmean).all():
        offset = np.eye(sigma1.shape[0]) * eps
        covmean = linalg.sqrtm((sigma1 + offset).dot(sigma2 + offset))
    if np.iscomplexobj(covmean):
        if not np.allclose(np.diagonal(covmean).imag, 0, atol=1e-3):
            raise ValueError("Imaginary component {}".format(np.max(np.abs(covmean.imag))))
        covmean = covmean.real
    return float(diff.dot(diff) + np.trace(sigma1) + np.trace(sigma2) - 2 * np.trace(covmean))


class EmbeddingSpaceEvaluator:
    """push_samples / get_scores of evaluation/FGD.py:18-188 for one feature extractor.

    ``extract`` maps poses ``[n, dim, 90]`` (channel-first 90-frame chunks: what the reference's
    ``transpose(1, 2).unfold(1, 90, 90).flatten(0, 1)`` produces, test_holisticbody.py:279) to features ``[n, 128]`` on the
    GPU -- pass ``EmbeddingPlan(...).extract`` (engine.py).  ``extract`` may return a view of a buffer it reuses: the
    features are copied here before the next call."""

    def __init__(self, extract):
        self.extract = extract
        self.real_feat, self.gene_feat = [], []

    def reset(self):
        self.real_feat, self.gene_feat = [], []

    def push_samples(self, generated_poses, real_poses):
        # the reference's `extract` returns feat.transpose(0, 1) (embedding_net.py:150): a push contributes a [128, n] block
        # .clone(): `.t().contiguous()` of a one-chunk [1, 128] result is a no-copy view of the extractor's buffer
        self.real_feat.append(self.extract(real_poses.float()).t().clone())
        self.gene_feat.append(self.extract(generated_poses.float()).t().clone())

    def get_scores(self):
        """-> (frechet_dist, feat_dist) exactly as FGD.py:155-188 computes them: the stacked [128 * pushes, n] matrices are
        handed to mean / cov with rowvar=False, i.e. statistics over feature-dimension rows with the n samples of a push
        as columns (all pushes must have the same n <= 128); feat_dist = mean over rows of the L1 row distance."""
        gene, real = torch.cat(self.gene_feat), torch.cat(self.real_feat)
        gm, gs = feature_stats(gene)
        rm, rs = feature_stats(real)
        try:
            fgd = frechet_distance(gm.cpu().numpy(), gs.cpu().numpy(), rm.cpu().numpy(), rs.cpu().numpy())
        except ValueError:
            fgd = 1e10
        r, g = real.cpu().numpy(), gene.cpu().numpy()       # host arithmetic as in the reference (FGD.py:181-186)
        feat_dist = float(np.mean(np.sum(np.absolute(r - g), axis=1)))
        return fgd, feat_dist
