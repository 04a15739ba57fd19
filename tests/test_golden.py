"""Oracle vs the committed golden vectors (outputs of the UNMODIFIED reference, oracle/make_golden.py).
Runs anywhere (CPU, no reference tree needed)."""
import os

import pytest
import torch

from oracle import probtalk_oracle as O
from probtalk_b200 import synthetic as S

GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.pt")


@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD, map_location="cpu", weights_only=False)


@pytest.fixture(scope="module")
def states(gold):
    s = gold["seeds"]
    return S.make_vq_state(s["vq"]), S.make_predictor_state(s["pred"]), S.make_refiner_state(s["ref"])


def noise(T, B, N, seed):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(T, B, N, generator=g)
    gum = -torch.log(-torch.log(torch.rand(T, B, N, generator=g).clamp_(1e-10, 1 - 1e-7)))
    return u, gum


def test_cfg1_vqvae(gold, states):
    c = gold["cfg1"]
    x = S.synthetic_poses(1, 300, c["poses_seed"])
    z, idx, rec = O.vqvae_forward(x, states[0])
    assert torch.allclose(z, c["z"], atol=2e-5, rtol=1e-5)
    assert torch.equal(idx, c["codes"].long())
    assert torch.allclose(rec, c["recon"], atol=2e-5, rtol=1e-5)


def test_cfg2_pq_layer(gold, states):
    c = gold["cfg2"]
    g = torch.Generator().manual_seed(c["seed"])
    for case in c["cases"]:
        z = torch.randn(c["B"], 512, c["W"], generator=g) * case["scale"]
        assert torch.equal(O.pq_search(z, O.pq_codebooks(states[0])), case["codes"].long())


def test_cfg3_logits(gold, states):
    c = gold["cfg3"]
    g = torch.Generator().manual_seed(c["seed"])
    tokens = torch.randint(0, 129, (3, 22, 4), generator=g, dtype=torch.int32)
    cond = torch.randn(3, 768, 22, generator=g)
    out = O.tokens_to_logits(tokens, c["label"], cond, O._sub(states[1]["Predictor"], "decoder"))
    assert torch.allclose(out, c["logits"], atol=5e-5, rtol=1e-4)


def test_cfg4_predictor_refiner(gold, states):
    c = gold["cfg4"]
    vq, pst, rst = states
    B, T = c["B"], c["T"]
    aud = S.synthetic_audio(1, T, c["aud_seed"]).repeat(B, 1, 1)
    gt = S.synthetic_poses(B, T, c["gt_seed"])
    mask = torch.zeros(B, 1, T)
    mask[:, :, :30] = 1
    u, gum = noise(8, B, 88, c["noise_seed"])
    ids = torch.tensor([c["id"]]).repeat(B)
    poses, codes = O.predictor_infer_on_batch(aud, gt, mask, ids, pst, vq, u, gum)
    assert torch.equal(codes.long(), c["codes"].long())
    assert torch.allclose(poses, c["poses"], atol=5e-5, rtol=1e-4)
    out = O.refiner_infer_on_batch(aud, gt, mask, torch.tensor([c["id"]]), poses, rst["HFTransformer"])
    assert torch.allclose(out, c["refined"], atol=1e-4, rtol=1e-4)
