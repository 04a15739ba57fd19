"""Pin the oracle (oracle/probtalk_oracle.py) against the UNMODIFIED reference run on CPU.

Only possible where /root/reference exists (the authoring container); elsewhere the committed
golden vectors (tests/test_golden.py) carry the pin.  Both sides get the same synthetic
checkpoints (probtalk_b200.synthetic) through the reference's own load_state_dict."""
import numpy as np
import pytest
import torch

from oracle import probtalk_oracle as O
from oracle import ref_harness as RH
from probtalk_b200 import synthetic as S

pytestmark = pytest.mark.skipif(not RH.reference_available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def states():
    torch.manual_seed(0)
    return S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)


@pytest.fixture(scope="module")
def ref_vq(states):
    return RH.build_vqt(states[0])


@pytest.fixture(scope="module")
def ref_pred(states):
    return RH.build_predictor(states[0], states[1])


@pytest.fixture(scope="module")
def ref_refiner(states):
    return RH.build_refiner(states[0], states[2])


def test_vqvae_encode_quantize_decode(states, ref_vq):
    vq = states[0]
    x = S.synthetic_poses(2, 300, 1001)
    with torch.no_grad():
        z_ref, _ = ref_vq.VQ.encoder(x)
        qo = ref_vq.VQ.vq_layer(z_ref)
        rec_ref = ref_vq.VQ.decode(qo.quantized, None)
        z, idx, rec = O.vqvae_forward(x, vq)
    assert z.shape == (2, 512, 37) and rec.shape == (2, 376, 296)
    assert torch.allclose(z, z_ref, atol=2e-5, rtol=1e-5)
    assert torch.equal(idx, qo.loss)            # eval: indices are returned in the .loss field (vqvae_modules.py:491)
    assert torch.allclose(rec, rec_ref, atol=2e-5, rtol=1e-5)
    # the quantized tensor is the gathered codebook rows, [B, 512, T']
    e = O.pq_gather(idx, O.pq_codebooks(vq)).transpose(1, 2)
    assert torch.equal(e, qo.quantized)


def test_pq_search_matches_reference_layer(states, ref_vq):
    g = torch.Generator().manual_seed(1002)
    for scale in (1.0, 0.1):
        z = torch.randn(32, 512, 22, generator=g) * scale
        with torch.no_grad():
            qo = ref_vq.VQ.vq_layer(z)
        idx = O.pq_search(z, O.pq_codebooks(states[0]))
        assert torch.equal(idx, qo.loss)
        idx64, gap = O.pq_search_fp64(z, O.pq_codebooks(states[0]))
        bad = idx64 != idx
        assert bool((gap[bad] < O.pq_tie_eps(z, O.pq_codebooks(states[0]))[bad]).all())


def test_audio_encoder_and_context(states, ref_pred):
    pst = states[1]
    aud = S.synthetic_audio(3, 180, 1003)
    gt = S.synthetic_poses(3, 180, 1004)
    mask = torch.zeros(3, 1, 180)
    mask[:, :, :30] = 1
    with torch.no_grad():
        a_ref = ref_pred.AudEnc(aud)
        a = O.condition_encoder(aud, pst["AudEnc"])
        assert torch.allclose(a, a_ref, atol=1e-5, rtol=1e-5)
        P = ref_pred.Predictor
        mm = P.motion_embedding(torch.cat([gt * mask, mask], 1))
        for layer in P.enc:
            mm = layer(mm)
        cond_ref = torch.cat([mm, a_ref], 1)
        _, cond = O.predictor_condition(aud, gt, mask, pst)
    assert cond.shape == (3, 768, 22)
    assert torch.allclose(cond, cond_ref, atol=2e-5, rtol=1e-5)


def test_tokens_to_logits(states, ref_pred):
    pst = states[1]
    g = torch.Generator().manual_seed(1005)
    tokens = torch.randint(0, 129, (3, 22, 4), generator=g, dtype=torch.int32)
    label = torch.tensor([0, 3, 1])
    cond = torch.randn(3, 768, 22, generator=g)
    with torch.no_grad():
        ref = ref_pred.Predictor.decoder.tokens_to_logits(tokens, label, cond)
        out = O.tokens_to_logits(tokens, label, cond, O._sub(pst["Predictor"], "decoder"))
    assert out.shape == (3, 22, 4, 128)
    assert torch.allclose(out, ref, atol=5e-5, rtol=1e-4)
    # the synthetic weights make the logits depend on the condition (SURVEY 8c caveat 1)
    with torch.no_grad():
        out2 = O.tokens_to_logits(tokens, label, cond * 0.5, O._sub(pst["Predictor"], "decoder"))
    assert (out - out2).abs().max() > 1e-2


def _patched_sampler(u_all, gum_all):
    """Replace the two RNG draws of sample_good by injected noise (the rest of the loop is the reference's)."""
    import torch.distributions as D
    state = {"t": 0}
    orig_cat, orig_gum = D.categorical.Categorical.sample, D.gumbel.Gumbel.sample

    def cat_sample(self, sample_shape=torch.Size()):
        return O.categorical_from_uniform(self.logits, u_all[state["t"]])

    def gum_sample(self, sample_shape=torch.Size()):
        t = state["t"]
        state["t"] += 1
        return gum_all[t]

    D.categorical.Categorical.sample = cat_sample
    D.gumbel.Gumbel.sample = gum_sample

    def restore():
        D.categorical.Categorical.sample = orig_cat
        D.gumbel.Gumbel.sample = orig_gum
    return restore


def _noise(T, B, N, seed):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(T, B, N, generator=g)
    gum = -torch.log(-torch.log(torch.rand(T, B, N, generator=g).clamp_(1e-10, 1 - 1e-7)))
    return u, gum


def test_sample_good_with_injected_noise(states, ref_pred):
    pst = states[1]
    g = torch.Generator().manual_seed(1006)
    B = 3
    cond = torch.randn(B, 768, 22, generator=g)
    label = torch.tensor([2, 0, 1])
    u, gum = _noise(8, B, 88, 1007)
    restore = _patched_sampler(u, gum)
    try:
        with torch.no_grad():
            ref = ref_pred.Predictor.decoder.sample_good(label, cond, None, T=8)
    finally:
        restore()
    ids = O.sample_good(label, cond, O._sub(pst["Predictor"], "decoder"), 8, u, gum)
    assert ids.shape == (B, 22, 4)
    assert int((ids == O.MASK_ID).sum()) == 0
    assert torch.equal(ids, ref.to(torch.int32))


def test_predictor_and_refiner_infer_on_batch(states, ref_pred, ref_refiner):
    vq, pst, rst = states
    B, T = 2, 180
    aud1 = S.synthetic_audio(1, T, 1008)
    gt = S.synthetic_poses(B, T, 1009)
    mask = torch.zeros(B, 1, T)
    mask[:, :, :30] = 1
    u, gum = _noise(8, B, 88, 1010)
    restore = _patched_sampler(u, gum)
    try:
        with torch.no_grad():
            ref_poses, ref_codes, _ = ref_pred.infer_to_code(aud=aud1, text=aud1, id=1, B=B, gt_poses=gt, mask=mask)
    finally:
        restore()
    aud = aud1.repeat(B, 1, 1)
    ids = torch.tensor([1]).repeat(B)
    poses, codes = O.predictor_infer_on_batch(aud, gt, mask, ids, pst, vq, u, gum)
    assert torch.equal(codes, ref_codes.to(torch.int32))
    assert torch.allclose(poses, ref_poses, atol=5e-5, rtol=1e-4)
    with torch.no_grad():
        ref_out, _ = ref_refiner.infer_on_batch(aud=aud1, B=B, id=torch.tensor([1]), gt_poses=gt, mask=mask,
                                                pred_poses=ref_poses)
    out = O.refiner_infer_on_batch(aud, gt, mask, torch.tensor([1]), poses, rst["HFTransformer"])
    assert out.shape == (B, 376, T)
    assert torch.allclose(out, ref_out, atol=1e-4, rtol=1e-4)


def test_window_loops_match_reference_scripts(states, ref_pred, ref_refiner):
    """oracle/window_oracle.py against LITERAL transcriptions of the script loops (scripts/test_holisticbody.py:239-262,
    scripts/demo.py:116-181) driving the unmodified reference wrappers / All_In_One_Model.__call__, sampler noise
    injected per window.  t = 300: windows of 180 and 150 frames (the second with 30 known frames of feedback)."""
    from oracle import make_golden_windows as GW
    from oracle import window_oracle as WO
    vq, pst, rst = states
    gen = GW.reference_generator(ref_pred, ref_refiner)
    c = GW.EVAL
    aud1, gt1, mask, noise = GW.eval_inputs()
    with torch.no_grad():
        ref = GW.ref_windows_eval(gen, aud1, aud1, gt1, mask, torch.tensor([c["id"]]), c["B"], noise)
    B = c["B"]
    ids = torch.tensor([c["id"]]).repeat(B)
    out, codes = WO.windows_eval(aud1.repeat(B, 1, 1), gt1.repeat(B, 1, 1), mask, ids, pst, vq, rst["HFTransformer"], noise=noise)
    assert out.shape == ref.shape == (B, 376, c["t"])
    assert [tuple(x.shape) for x in codes] == [(B, 22, 4), (B, 18, 4)]
    assert torch.allclose(out, ref, atol=2e-4, rtol=1e-4)
    assert torch.equal(out[:, :, c["known"][0]:c["known"][1]], gt1.repeat(B, 1, 1)[:, :, c["known"][0]:c["known"][1]])
    # teacher-forcing the codes reproduces the same poses (what the GPU parity test relies on)
    out_tf, _ = WO.windows_eval(aud1.repeat(B, 1, 1), gt1.repeat(B, 1, 1), mask, ids, pst, vq, rst["HFTransformer"], codes=codes)
    assert torch.equal(out_tf, out)
    d = GW.DEMO
    aud1, noise = GW.demo_inputs()
    with torch.no_grad():
        ref_poses, ref_codes, _ = GW.ref_windows_demo(gen, aud1, torch.tensor([d["id"]]), d["B"], noise)
    ids = torch.tensor([d["id"]]).repeat(d["B"])
    poses, all_codes, _ = WO.windows_demo(aud1.repeat(d["B"], 1, 1), ids, pst, vq, rst["HFTransformer"], noise=noise)
    assert torch.equal(all_codes.long(), ref_codes.long()) and all_codes.shape == (d["B"], 22 + 18 - 3, 4)
    assert torch.allclose(poses, ref_poses, atol=2e-4, rtol=1e-4)
