"""CPU restatement of the reference's FGD evaluator path (TEST INFRASTRUCTURE -- only tests/ may import this).

  pose_encoder_features   PoseEncoderConv.forward in eval mode with variational_encoding=False
                          (nets/embedding_net.py:56-95; ConvNormRelu :29-53; EmbeddingNet.extract :147-150)
  feature_stats           np.mean(axis=0), np.cov(rowvar=False)          (evaluation/FGD.py:172-176)
  frechet_distance        calculate_frechet_distance                      (evaluation/FGD.py:190-257)
  scores                  EmbeddingSpaceEvaluator.get_scores              (evaluation/FGD.py:155-188)
Pinned by tests/golden/fgd_vectors.pt, generated from the unmodified reference by oracle/make_golden_fgd.py.
"""
import numpy as np
import torch
import torch.nn.functional as F


def _bn(x, sd, p):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"], False, 0.0, 1e-5)


def pose_encoder_features(sd, poses, prefix="pose_encoder."):
    """sd: EmbeddingNet state_dict; poses [n, dim, 90] fp32 -> mu [n, 128]."""
    g = lambda k: sd[prefix + k]
    x = poses
    for i, stride in ((0, 1), (1, 1), (2, 2)):
        x = F.conv1d(x, g("net.%d.0.weight" % i), g("net.%d.0.bias" % i), stride=stride)
        x = F.leaky_relu(_bn(x, sd, prefix + "net.%d.1" % i), 0.2)
    x = F.conv1d(x, g("net.3.weight"), g("net.3.bias")).flatten(1)
    x = F.linear(x, g("out_net.0.weight"), g("out_net.0.bias"))
    x = F.leaky_relu(_bn(x, sd, prefix + "out_net.1"), 1.0)      # nn.LeakyReLU(True): negative_slope = 1.0
    x = F.linear(x, g("out_net.3.weight"), g("out_net.3.bias"))
    x = F.leaky_relu(_bn(x, sd, prefix + "out_net.4"), 1.0)
    x = F.linear(x, g("out_net.6.weight"), g("out_net.6.bias"))
    return F.linear(x, g("fc_mu.weight"), g("fc_mu.bias"))


def feature_stats(feats):
    feats = np.asarray(feats)
    return np.mean(feats, axis=0), np.cov(feats, rowvar=False)


def frechet_distance(mu1, sigma1, mu2, sigma2, eps=1e-6):
    from scipy import linalg
    diff = mu1 - mu2
    covmean = linalg.sqrtm(sigma1.dot(sigma2))
    if isinstance(covmean, tuple):
        covmean = covmean[0]
    if not np.isfinite(covmean).all():
        offset = np.eye(sigma1.shape[0]) * eps
        covmean = linalg.sqrtm((sigma1 + offset).dot(sigma2 + offset))
    if np.iscomplexobj(covmean):
        if not np.allclose(np.diagonal(covmean).imag, 0, atol=1e-3):
            raise ValueError("Imaginary component")
        covmean = covmean.real
    return float(diff.dot(diff) + np.trace(sigma1) + np.trace(sigma2) - 2 * np.trace(covmean))


def stack_pushes(batches):
    """What EmbeddingSpaceEvaluator accumulates: `extract` returns feat.transpose(0, 1) (embedding_net.py:150), so every
    push contributes a [128, n] block and np.vstack gives [128 * pushes, n]: the ROWS the statistics run over are feature
    dimensions and the COLUMNS are the samples of a push (FGD.py:52-71, 168-169).  Reproduced as is."""
    return np.vstack([np.asarray(b).T for b in batches])


def scores(gene_batches, real_batches):
    """gene_batches / real_batches: lists of per-push feature arrays [n, 128] (equal n) -> (fgd, feat_dist)."""
    gene, real = stack_pushes(gene_batches), stack_pushes(real_batches)
    gm, gs = feature_stats(gene)
    rm, rs = feature_stats(real)
    fgd = frechet_distance(gm, gs, rm, rs)
    feat_dist = float(np.mean(np.sum(np.absolute(real - gene), axis=1)))
    return fgd, feat_dist


def lvd(gt_kps, pr_kps):
    """LVD of evaluation/metrics.py:79-94 for gt [T, J, D] and pr [B, T, J, D] (torch tensors): the Batch_LVD branch
    (:26-77) with symmetrical = False, weight = False, which is how the evaluator calls it
    (scripts/test_holisticbody.py:120,130)."""
    import torch
    gt_kps, pr_kps = gt_kps.squeeze(), pr_kps.squeeze()
    if pr_kps.dim() == 3:
        gt_v = (gt_kps[1:] - gt_kps[:-1]).norm(p=2, dim=-1)
        pr_v = (pr_kps[1:] - pr_kps[:-1]).norm(p=2, dim=-1)
        return float((pr_v - gt_v).abs().sum(dim=-1).mean())
    length = min(gt_kps.shape[0], pr_kps.shape[1])
    gt_kps, pr_kps = gt_kps[:length], pr_kps[:, :length]
    gt_v = (gt_kps[1:] - gt_kps[:-1]).norm(p=2, dim=-1)
    pr_v = (pr_kps[:, 1:] - pr_kps[:, :-1]).norm(p=2, dim=-1)
    w = 1 / gt_v.shape[0]
    return float(((pr_v - gt_v).abs().sum(dim=-1) * w).sum(dim=-1).mean())


def diversity(kps):
    """diversity of evaluation/metrics.py:96-110: kps [bs, seq, dim] (numpy), mean over pairs of mean |x_i - x_j|."""
    dis = []
    for i in range(kps.shape[0]):
        for j in range(i + 1, kps.shape[0]):
            dis.append(np.mean(np.abs(kps[i] - kps[j])))
    return float(np.mean(dis))
