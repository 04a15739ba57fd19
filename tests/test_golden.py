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


def test_window_oracle_replays_reference_window_vectors():
    """oracle/window_oracle.py against tests/golden/window_vectors.pt (the reference's own loops, made by
    oracle/make_golden_windows.py): runs anywhere, no reference tree needed."""
    import os
    from oracle import make_golden_windows as GW
    from oracle import window_oracle as WO
    from probtalk_b200 import synthetic as S
    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "window_vectors.pt"), map_location="cpu",
                      weights_only=False)
    s = gold["seeds"]
    vq, pst, rst = S.make_vq_state(s["vq"]), S.make_predictor_state(s["pred"]), S.make_refiner_state(s["ref"])
    c = gold["eval"]
    aud1, gt1, mask, noise = GW.eval_inputs()
    B = c["B"]
    ids = torch.tensor([c["id"]]).repeat(B)
    out, codes = WO.windows_eval(aud1.repeat(B, 1, 1), gt1.repeat(B, 1, 1), mask, ids, pst, vq, rst["HFTransformer"], noise=noise)
    assert all(torch.equal(a.long(), b.long()) for a, b in zip(codes, c["codes"]))
    assert torch.allclose(out, c["poses"], atol=2e-4, rtol=1e-4)
    d = gold["demo"]
    ids = torch.tensor([d["id"]]).repeat(d["B"])
    aud1, noise = GW.demo_inputs()
    # teacher-forced with the reference's per-window codes (cheap: no sampler iterations)
    poses, all_codes, _ = WO.windows_demo(aud1.repeat(d["B"], 1, 1), ids, pst, vq, rst["HFTransformer"],
                                          codes=[x.long() for x in d["window_codes"]])
    assert torch.equal(all_codes.long(), d["codes"].long())
    assert torch.allclose(poses, d["poses"], atol=2e-4, rtol=1e-4)
