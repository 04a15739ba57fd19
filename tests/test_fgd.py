"""FGD evaluator: oracle vs the reference's golden vectors (CPU); feature-statistics kernel,
feature extractor plan and evaluator vs oracle / golden (GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import fgd_oracle as FO
from oracle.make_golden_fgd import make_inputs

GOLD = os.path.join(os.path.dirname(__file__), "golden", "fgd_vectors.pt")


@pytest.fixture(scope="module")
def gold():
    g = torch.load(GOLD, weights_only=False)
    g["real"], g["gene"] = make_inputs(g["n"], g["dim"], g["input_seed"])
    assert torch.equal(g["real"][:2, :4, :5], g["real_probe"]) and torch.equal(g["gene"][:2, :4, :5], g["gene_probe"])
    return g


def test_oracle_features_match_reference_golden(gold):
    with torch.no_grad():
        fr = FO.pose_encoder_features(gold["state_dict"], gold["real"])
        fg = FO.pose_encoder_features(gold["state_dict"], gold["gene"])
    # same fp32 torch ops as nets/embedding_net.py (functional instead of module form)
    assert (fr - gold["feat_real"]).abs().max() < 2e-5 and (fg - gold["feat_gene"]).abs().max() < 2e-5


def test_oracle_scores_match_reference_golden(gold):
    fgd, fd = FO.scores([gold["feat_gene"][i:i + 100].numpy() for i in range(0, 300, 100)],
                        [gold["feat_real"][i:i + 100].numpy() for i in range(0, 300, 100)])
    assert abs(fgd - gold["fgd"]) < 1e-6 * max(1.0, abs(gold["fgd"]))
    assert abs(fd - gold["feat_dist"]) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("N,D", [(300, 128), (2, 128), (1000, 100), (4097, 7), (20000, 128)])
def test_feature_stats_kernel_vs_numpy(N, D):
    from probtalk_b200 import evaluation as EV
    g = torch.Generator().manual_seed(N + D)
    x = torch.randn(N, D, generator=g) * 2.0 + torch.linspace(-3, 3, D)
    mean, cov = EV.feature_stats(x.cuda())
    rm, rc = FO.feature_stats(x.numpy().astype(np.float64))
    # fp64 both sides; numpy sums pairwise, the kernel in fixed chunks: agreement to ~1e-12 relative
    assert np.abs(mean.cpu().numpy() - rm).max() < 1e-11
    assert np.abs(cov.cpu().numpy().reshape(D, D) - np.atleast_2d(rc)).max() < 1e-10
    mean2, cov2 = EV.feature_stats(x.cuda())
    assert torch.equal(mean, mean2) and torch.equal(cov, cov2)          # deterministic reduction order


@pytest.mark.gpu
def test_feature_stats_rejects_bad_arguments():
    from probtalk_b200 import _lib, evaluation as EV
    with pytest.raises(_lib.PtError):
        EV.feature_stats(torch.randn(10, 129).cuda())
    with pytest.raises(_lib.PtError):
        EV.feature_stats(torch.randn(1, 16).cuda())


@pytest.mark.gpu
def test_embedding_plan_and_evaluator_vs_golden(gold):
    from probtalk_b200 import engine as E, evaluation as EV
    W = E.EmbeddingWeights(gold["state_dict"], "cuda")
    plan = E.EmbeddingPlan(W, 128)
    feats = torch.cat([plan.extract(gold["real"][i:i + 128].cuda()).clone() for i in range(0, gold["n"], 128)])
    # tolerance: bf16 operands / fp32 accumulate through 4 convolutions and 4 linear layers, features of std 1.8
    err = (feats.cpu() - gold["feat_real"]).abs()
    assert err.max() < 0.12 and err.mean() < 0.02
    ev = EV.EmbeddingSpaceEvaluator(plan.extract)
    for i in range(0, gold["n"], 100):
        ev.push_samples(gold["gene"][i:i + 100].cuda(), gold["real"][i:i + 100].cuda())
    fgd, fd = ev.get_scores()
    assert abs(fd - gold["feat_dist"]) < 0.02 * gold["feat_dist"]
    assert abs(fgd - gold["fgd"]) < 0.05 * gold["fgd"]
    # the statistics + Frechet half alone, fed the reference's own features, reproduces the reference's number
    stack = lambda f: torch.cat([f[i:i + 100].t() for i in range(0, 300, 100)]).contiguous().cuda()
    gm, gs = EV.feature_stats(stack(gold["feat_gene"]))
    rm, rs = EV.feature_stats(stack(gold["feat_real"]))
    exact = EV.frechet_distance(gm.cpu().numpy(), gs.cpu().numpy(), rm.cpu().numpy(), rs.cpu().numpy())
    assert abs(exact - gold["fgd"]) < 1e-6 * max(1.0, gold["fgd"])


@pytest.mark.gpu
def test_evaluator_single_chunk_push_does_not_alias(gold):
    """A push with ONE 90-frame chunk (any 90..179-frame clip): the real features must survive the generated-feature
    extraction that follows (the extractor reuses its output buffer)."""
    from probtalk_b200 import engine as E, evaluation as EV
    plan = E.EmbeddingPlan(E.EmbeddingWeights(gold["state_dict"], "cuda"), 4)
    ev = EV.EmbeddingSpaceEvaluator(plan.extract)
    for i in range(3):
        ev.push_samples(gold["gene"][i:i + 1].cuda(), gold["real"][i:i + 1].cuda())
    for i in range(3):
        fr = plan.extract(gold["real"][i:i + 1].cuda()).clone()
        fg = plan.extract(gold["gene"][i:i + 1].cuda()).clone()
        assert torch.equal(ev.real_feat[i], fr.t()) and torch.equal(ev.gene_feat[i], fg.t())
        assert not torch.equal(ev.real_feat[i], ev.gene_feat[i])


METRICS_GOLD = os.path.join(os.path.dirname(__file__), "golden", "metrics_vectors.pt")


def test_oracle_lvd_diversity_match_reference_golden():
    from oracle.make_golden_metrics import make_inputs
    for c in torch.load(METRICS_GOLD, weights_only=False)["cases"]:
        gt, pr = make_inputs(c["seed"], *c["shape"])
        B, T, J, D = c["shape"]
        assert abs(FO.lvd(gt, pr) - c["lvd"]) <= 1e-6 * max(1.0, abs(c["lvd"]))
        assert abs(FO.lvd(gt, pr[0]) - c["lvd_single"]) <= 1e-6 * max(1.0, abs(c["lvd_single"]))
        assert abs(FO.diversity(pr.reshape(B, T, J * D).numpy()) - c["diversity"]) <= 1e-6 * max(1.0, c["diversity"])


@pytest.mark.gpu
def test_lvd_diversity_kernels_vs_reference_golden():
    """pt_lvd / pt_diversity (fp32 differences and norms, fp64 sums) against the reference's own numbers."""
    from oracle.make_golden_metrics import make_inputs
    from probtalk_b200 import evaluation as EV
    for c in torch.load(METRICS_GOLD, weights_only=False)["cases"]:
        gt, pr = make_inputs(c["seed"], *c["shape"])
        B, T, J, D = c["shape"]
        v, per = EV.lvd(gt.cuda(), pr.cuda())
        assert abs(v - c["lvd"]) <= 2e-6 * max(1.0, abs(c["lvd"])) and per.shape == (B,)
        v1, _ = EV.lvd(gt.cuda(), pr[0].cuda())
        assert abs(v1 - c["lvd_single"]) <= 2e-6 * max(1.0, abs(c["lvd_single"]))
        d, pairs = EV.diversity(pr.reshape(B, T, J * D).cuda())
        assert abs(d - c["diversity"]) <= 2e-6 * max(1.0, c["diversity"]) and pairs.numel() == B * (B - 1) // 2
        d2, pairs2 = EV.diversity(pr.reshape(B, T, J * D).cuda())
        assert d == d2 and torch.equal(pairs, pairs2)              # deterministic reduction order
    from probtalk_b200 import _lib
    with pytest.raises(_lib.PtError):
        EV.diversity(torch.zeros(1, 10, 3).cuda())
