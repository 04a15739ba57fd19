"""wav2vec2 audio features on the B200 against the fp32 model the reference calls.

Reference path: scripts/demo.py:105-111 -- Wav2Vec2Processor normalisation, transformers.Wav2Vec2Model(...).last_hidden_state,
data_utils/utils.py:14-16 linear interpolation to 30 fps.  The checkpoint (facebook/wav2vec2-base-960h) cannot be fetched
here, so the model is the same architecture with seeded, fan-in scaled random weights (synthetic.make_wav2vec2).  The
oracle is the transformers implementation itself, run in fp32 on the host cores.

Stated bounds.  The path computes in bf16 (fp32 accumulation / statistics / softmax); every stage ends in a LayerNorm or
GroupNorm, outputs are O(1).  Observed on B200 (printed by the tests): see BOUND below = 1.5 x observed.
"""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# stage -> (max-abs, mean-abs) bound on |ours - fp32 model|; the values are O(1) (unit variance after each norm)
BOUND = {
    # observed (B, L) = (1, 48000) / (2, 20013) / (1, 400000):
    #   conv0 .031 / .00055   features .041 / .0015   projected .039 / .0062   hidden0 .074 / .0065
    #   last_hidden_state .099 / .0075   aud .089 / .0073        (|reference| mean: 0.39, 0.22, 0.86, 0.78, 0.78, 0.78)
    "conv0": (0.047, 0.00085), "features": (0.062, 0.0023), "projected": (0.059, 0.0093), "hidden0": (0.11, 0.0097),
    "last_hidden_state": (0.15, 0.0113), "aud": (0.134, 0.011),
}


@pytest.fixture(scope="module")
def hf_model():
    from probtalk_b200 import synthetic as S
    return S.make_wav2vec2(11)


def _normalise(w):
    """Wav2Vec2FeatureExtractor.zero_mean_unit_var_norm (per utterance)."""
    w = w.double()
    return ((w - w.mean(-1, keepdim=True)) / torch.sqrt(w.var(-1, unbiased=False, keepdim=True) + 1e-7)).float()


# ------------------------------------------------------------------------------------------------ kernels
@pytest.mark.parametrize("B,L", [(1, 16000), (3, 4013), (2, 400), (1, 160007)])
def test_frontend_kernel(B, L):
    """normalisation + conv0 + GroupNorm + GELU in one pass vs fp64 torch."""
    from probtalk_b200 import ops, synthetic as S
    g = torch.Generator().manual_seed(L)
    wave = S.synthetic_waveform(B, L, L) * 0.3 + 0.01
    w0 = torch.randn(512, 1, 10, generator=g) * math.sqrt(2.0 / 10)
    gam, bet = torch.rand(512, generator=g) + 0.5, torch.randn(512, generator=g) * 0.05
    x = _normalise(wave).double()
    y = F.conv1d(x[:, None], w0.double(), stride=5)
    ref = F.gelu(F.group_norm(y, 512, gam.double(), bet.double(), 1e-5)).transpose(1, 2)
    out = ops.w2v_frontend(wave.cuda(), w0.reshape(512, 10).cuda().contiguous(), gam.cuda(), bet.cuda())
    assert out.shape == ref.shape
    err = (out.double().cpu() - ref).abs()
    tol = 2.0 ** -8 * ref.abs().clamp_min(1.0)                  # the bf16 rounding of the output
    print("frontend B=%d L=%d: max err %.3e (|ref| max %.2f)" % (B, L, err.max().item(), ref.abs().max().item()))
    assert bool((err <= tol + 2e-4).all())


@pytest.mark.parametrize("B,T,C,Tout,dt", [(2, 99, 768, 59, torch.bfloat16), (1, 2999, 768, 1800, torch.float32), (3, 7, 40, 23, torch.float32),
                                           (1, 50, 768, 50, torch.bfloat16), (2, 1, 128, 4, torch.float32)])
def test_interp_kernel(B, T, C, Tout, dt):
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(T)
    rows = torch.randn(B, T, C, device="cuda", generator=g).to(dt)
    out = ops.w2v_interp(rows, Tout)
    ref = F.interpolate(rows.float().transpose(1, 2), size=Tout, mode="linear")
    assert out.shape == ref.shape
    assert (out - ref).abs().max().item() < 2e-6 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("M,C", [(1000, 768), (77, 512), (5, 128), (300, 1024)])
def test_layernorm_rows(M, C):
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + C)
    x = (torch.randn(M, C, device="cuda", generator=g) * 2 + 0.3).bfloat16()
    r = torch.randn(M, C, device="cuda", generator=g).bfloat16()
    gam, bet = torch.rand(C, device="cuda", generator=g) + 0.5, torch.randn(C, device="cuda", generator=g) * 0.1
    for res in (None, r):
        ref = F.layer_norm(x.double() + (0 if res is None else res.double()), (C,), gam.double(), bet.double(), 1e-5)
        out = ops.layernorm_rows(x, res, gam, bet)
        assert (out.double() - ref).abs().max().item() < 2.0 ** -8 * ref.abs().max().item() + 1e-5
    y = x.clone()
    ops.layernorm_rows(y, r, gam, bet, out=y)                  # in place
    assert torch.equal(y, ops.layernorm_rows(x, r, gam, bet))


@pytest.mark.parametrize("B,Hh,Nq,Nk", [(1, 12, 1249, 1249), (2, 12, 300, 1100), (1, 8, 2999, 2999), (2, 3, 1030, 65), (1, 2, 1000, 1000)])
def test_attention_streamed(B, Hh, Nq, Nk):
    """pt_attention beyond what fits on chip: keys streamed through shared memory, online softmax."""
    from probtalk_b200 import ops
    D = 64
    g = torch.Generator(device="cuda").manual_seed(Nq + Nk)
    q = torch.randn(B * Nq, Hh * D, device="cuda", generator=g).bfloat16()
    kv = torch.randn(B * Nk, 2 * Hh * D, device="cuda", generator=g).bfloat16()
    out = ops.attention(q, kv[:, :Hh * D], kv[:, Hh * D:], B, Nq, Nk, heads=Hh)
    qf = q.float().view(B, Nq, Hh, D).transpose(1, 2)
    kf = kv[:, :Hh * D].float().view(B, Nk, Hh, D).transpose(1, 2)
    vf = kv[:, Hh * D:].float().view(B, Nk, Hh, D).transpose(1, 2)
    ref = F.scaled_dot_product_attention(qf, kf, vf).transpose(1, 2).reshape(B * Nq, Hh * D)
    err = (out.float() - ref).abs().max().item()
    print("streamed attention Nq=%d Nk=%d: max err %.3e" % (Nq, Nk, err))
    assert err < 0.02                                           # bf16 probabilities and outputs, values ~ N(0, 1/sqrt(Nk)..1)


def test_gemm_gelu_strided_taps():
    """The feature encoder's k3 / k2 stride-2 valid convolutions + GELU as GEMMs over token-major rows."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    B, T, Cc = 2, 101, 512
    P = T + 1
    x = torch.zeros(B, P, Cc, device="cuda")
    x[:, :T] = torch.randn(B, T, Cc, device="cuda", generator=g)
    xb = x.bfloat16()
    for k in (3, 2):
        w = torch.randn(512, Cc, k, device="cuda", generator=g) * math.sqrt(2.0 / (Cc * k))
        wm = w.permute(0, 2, 1).reshape(512, -1).contiguous().bfloat16()
        To = (T - k) // 2 + 1
        out = torch.zeros(B, To, 512, device="cuda", dtype=torch.bfloat16)
        ops.gemm(xb.view(-1), wm, out.view(-1), M=B * P // 2, C_=Cc, taps=k, s=2, base_row=0, rows_per_sample=P // 2,
                 t_valid=To, oss=To * 512, ors=512, act=ops.ACT_GELU)
        ref = F.gelu(F.conv1d(xb.float()[:, :T].transpose(1, 2), wm.float().view(512, k, Cc).permute(0, 2, 1), stride=2))
        assert (out.float().transpose(1, 2) - ref).abs().max().item() < 0.02


# ------------------------------------------------------------------------------------------------ the whole path
def _hf_stages(model, wave):
    x = _normalise(wave)
    with torch.no_grad():
        conv0 = model.feature_extractor.conv_layers[0](x[:, None]).transpose(1, 2)
        feats = model.feature_extractor(x).transpose(1, 2)          # (the output's `extract_features` is already normalised)
        proj, _ = model.feature_projection(feats)
        o = model(x, output_hidden_states=True)
    return {"conv0": conv0, "features": feats, "projected": proj, "hidden0": o.hidden_states[0],
            "last_hidden_state": o.last_hidden_state}


@pytest.mark.parametrize("B,L", [(1, 48000), (2, 20013), (1, 400000)])
def test_wav2vec2_against_transformers(hf_model, B, L):
    from probtalk_b200 import audio as A, synthetic as S
    wave = S.synthetic_waveform(B, L, 100 + B)
    ref = _hf_stages(hf_model, wave)
    W = A.Wav2VecWeights(hf_model.state_dict(), "cuda")
    plan = A.Wav2VecPlan(W, B, L)
    assert plan.T == ref["last_hidden_state"].shape[1] and plan.t_out == int(L / 16000 * 30)
    got = {"conv0": plan.run_until(wave, "conv0"), "features": plan.run_until(wave, "conv6"),
           "projected": plan.run_until(wave, "projected"), "hidden0": plan.run_until(wave, "hidden0")}
    aud = plan.run(wave.cuda() if B == 1 else wave)                      # device and host inputs
    got["last_hidden_state"] = plan.last_hidden_state()
    got["aud"] = aud
    ref["aud"] = F.interpolate(ref["last_hidden_state"].transpose(1, 2), size=plan.t_out, mode="linear")
    assert aud.shape == (B, 768, plan.t_out) and aud.dtype == torch.float32
    for name in ("conv0", "features", "projected", "hidden0", "last_hidden_state", "aud"):
        d = (got[name].float().cpu() - ref[name]).abs()
        print("w2v B=%d L=%d %-18s max %.4f mean %.5f (|ref| mean %.3f)" % (B, L, name, d.max().item(), d.mean().item(),
                                                                           ref[name].abs().mean().item()))
    for name in ("conv0", "features", "projected", "hidden0", "last_hidden_state", "aud"):
        d = (got[name].float().cpu() - ref[name]).abs()
        assert d.max().item() < BOUND[name][0] and d.mean().item() < BOUND[name][1], name
    # the interpolation itself is exact given our own encoder output
    own = F.interpolate(got["last_hidden_state"].transpose(1, 2), size=plan.t_out, mode="linear")
    assert (aud - own).abs().max().item() < 1e-5


def test_features_facade_feeds_the_pipeline(hf_model):
    """Wav2Vec2Features(audio) has the layout the predictor / refiner take as `aud` ([1, 768, t] at 30 fps), plans are cached."""
    from probtalk_b200 import audio as A, synthetic as S
    feats = A.Wav2Vec2Features.from_hf(hf_model)
    wave = S.synthetic_waveform(1, 96000, 9)[0]
    aud = feats(wave.numpy(), sampling_rate=16000)
    assert aud.shape == (1, 768, 180) and bool(torch.isfinite(aud).all())
    again = feats(wave, sampling_rate=16000)
    assert len(feats._plans) == 1 and torch.equal(aud, again)
    with pytest.raises(ValueError):
        feats(wave, sampling_rate=22050)
    with pytest.raises(ValueError):
        feats(wave[:300])


def test_demo_from_waveform(hf_model):
    """scripts/demo.py::infer end to end on the device: waveform -> wav2vec2 features -> predictor / decode / refiner windows ->
    SMPL-X rows, equal to the composition of its separately parity-tested parts."""
    from probtalk_b200 import audio as A, driver as D, postprocess as P, synthetic as S
    pipe = D.HolisticPipeline(S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5))
    feats = A.Wav2Vec2Features.from_hf(hf_model)
    wave = S.synthetic_waveform(1, 16000 * 11, 21)[0]                       # 11 s -> t = 330 frames: 2 + 2 windows
    g = torch.Generator(device="cuda").manual_seed(3)
    aud = feats(wave)
    assert aud.shape == (1, 768, 330)
    n_tok = [pipe.predictor(1, e - s).N for s, e in D.demo_window_bounds(330, 180)]
    noise = [(torch.rand(8, 1, n, device="cuda", generator=g),
              -torch.log(-torch.log(torch.rand(8, 1, n, device="cuda", generator=g).clamp_(1e-10, 1 - 1e-7)))) for n in n_tok]
    smplx, poses, codes = pipe.infer_from_audio(feats, wave, [2], noise=noise)
    ref_poses, ref_codes = pipe.windows_demo(aud, torch.tensor([2], dtype=torch.int32, device="cuda"), noise=noise)
    assert torch.equal(codes, ref_codes) and torch.equal(poses, ref_poses)
    assert poses.shape == (1, 376, 330) and smplx.shape == (1, 330, 265) and bool(torch.isfinite(smplx).all())
    assert torch.equal(smplx, P.poses_to_smplx(ref_poses))


def test_feature_error_downstream(hf_model):
    """What the bf16 wav2vec2 path costs downstream: the same windows (teacher-forced codes, so that both runs decode the same
    tokens) driven by our features and by the fp32 transformers features.  The refined poses differ by less than the
    pipeline's own bf16 tolerance (refiner max-abs 0.038 / mean 0.0058)."""
    from probtalk_b200 import audio as A, driver as D, synthetic as S
    pipe = D.HolisticPipeline(S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5))
    feats = A.Wav2Vec2Features.from_hf(hf_model)
    wave = S.synthetic_waveform(1, 16000 * 11, 22)
    aud = feats(wave)
    with torch.no_grad():
        hid = hf_model(_normalise(wave)).last_hidden_state
    aud_ref = F.interpolate(hid.transpose(1, 2), size=aud.shape[-1], mode="linear").cuda()
    ids = torch.tensor([1], dtype=torch.int32, device="cuda")
    poses, codes = pipe.windows_demo(aud, ids)
    # teacher-force both runs with the codes of the first: window w covers code steps [19 w, 19 w + 22) of the concatenation
    forced = []
    for w, (s0, e0) in enumerate(D.demo_window_bounds(aud.shape[-1], 180)):
        n = pipe.predictor(1, e0 - s0).N // codes.shape[-1]               # token steps of this window
        forced.append(codes[:, 19 * w:19 * w + n].contiguous())
    a, _ = pipe.windows_demo(aud, ids, force_codes=forced)
    b, _ = pipe.windows_demo(aud_ref, ids, force_codes=forced)
    d = (a - b).abs()
    print("downstream of the feature error: poses max %.4f mean %.5f (|pose| mean %.3f)" % (d.max().item(), d.mean().item(), b.abs().mean().item()))
    assert d.max().item() < 0.038 and d.mean().item() < 0.0058


def test_gemm_a_col_table_grouped_conv():
    """pt_gemm_desc.a_col_table: one launch whose 128-column output tiles read different channel slabs of A -- a grouped
    k = 5 'same' convolution (8 groups of 48 channels, 384 -> 384) against F.conv1d."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(9)
    B, T, Cc, G, K = 2, 70, 384, 8, 5
    cg = Cc // G
    x = torch.randn(B, Cc, T, device="cuda", generator=g)
    w = torch.randn(Cc, cg, K, device="cuda", generator=g) / (cg * K) ** 0.5
    bias = torch.randn(Cc, device="cuda", generator=g)
    H = K // 2
    Tp = T + 2 * H
    xf = torch.zeros(B, Tp, Cc, device="cuda")
    xf[:, H:H + T] = x.transpose(1, 2)
    xb = xf.bfloat16()
    offs = (0, 96, 192)                              # tile j = outputs [128 j, 128 j + 128): groups 0-2, 2-5, 5-7; slab of 192 channels
    wm = torch.zeros(Cc, K, 192, device="cuda")
    for n in range(Cc):
        j, grp = n // 128, n // cg
        lo = cg * grp - offs[j]
        assert 0 <= lo and lo + cg <= 192
        wm[n, :, lo:lo + cg] = w[n].t()
    wmat = wm.reshape(Cc, K * 192).bfloat16().contiguous()
    out = torch.zeros(B, T, Cc, device="cuda", dtype=torch.bfloat16)
    ops.gemm(xb.view(-1), wmat, out.view(-1), M=B * Tp, C_=192, taps=K, base_row=0, rows_per_sample=Tp, t_valid=T,
             oss=T * Cc, ors=Cc, bias=bias, a_pitch=Cc, a_rows=B * Tp, a_col_table=list(offs))
    ref = F.conv1d(xb.float()[:, H:H + T].transpose(1, 2), w.bfloat16().float(), bias, padding=H, groups=G)
    assert (out.float().transpose(1, 2) - ref).abs().max().item() < 0.03
    with pytest.raises(Exception):                   # offset + C beyond the row pitch
        ops.gemm(xb.view(-1), wmat, out.view(-1), M=B * Tp, C_=192, taps=K, base_row=0, rows_per_sample=Tp, t_valid=T,
                 oss=T * Cc, ors=Cc, bias=bias, a_pitch=Cc, a_rows=B * Tp, a_col_table=[0, 96, 200])
