"""The tolerance-tight arithmetic mode ("bf16x3": split-bf16 operands, three tensor-core passes, fp32 accumulation,
fp32 attention / LayerNorm) against fp64 torch, the oracle and the reference-generated golden vectors.

Stated bounds: poses / logits  max-abs <= 1e-3, mean-abs <= 1e-4 (normalised pose units) for identical
codes / tokens;  VQ.encode  >= 99.9 % of the PQ codes equal to the fp32 reference's, every mismatch a near-tie whose
fp64 top-2 gap is explained by the difference between our z and the reference's.
The PQ-VAE ENCODER always runs in this mode (engine.VQVAEWeights), whatever precision the rest of the path uses."""
import math
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import probtalk_oracle as O
from probtalk_b200 import synthetic as S

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.pt")

POSE_MAX, POSE_MEAN = 1e-3, 1e-4


def _split(x):
    from probtalk_b200 import ops
    return ops.split_bf16(x)


# ------------------------------------------------------------------------------------------------ kernels
@pytest.mark.parametrize("M,N,K", [(1000, 256, 512), (333, 384, 1024), (4100, 1536, 512), (19000, 1024, 512), (40000, 512, 2048)])
def test_gemm_split_plain(M, N, K):
    """A_hi W_hi + A_hi W_lo + A_lo W_hi vs an fp64 product of the fp32 operands: relative error ~2^-16 per term."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    bias = torch.randn(N, device="cuda", generator=g)
    ah, al = _split(a)
    ws = ops.split_weight(w)
    ref = (a.double() @ w.double().t() + bias.double())
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(ah, ws, out, M=M, C_=K, bias=bias, ors=N, a_lo=al)
    # outputs ~ N(0, 1): 2^-17 per operand split, plus the tensor core's fp32 accumulation (truncating adds, so the
    # error grows with the number of K steps: observed 3e-5 at K = 512, 7.4e-5 at K = 2048)
    tol = 4e-5 * max(1.0, K / 768.0)
    err = (out.double() - ref).abs().max().item()
    print("split GEMM M=%d N=%d K=%d: max abs err %.2e (outputs ~N(0,1))" % (M, N, K, err))
    assert err < tol
    oh = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    ol = torch.full_like(oh, float("nan"))
    ops.gemm(ah, ws, oh, M=M, C_=K, bias=bias, ors=N, a_lo=al, out_lo=ol, act=ops.ACT_RELU)
    got = oh.double() + ol.double()
    assert (got - ref.clamp_min(0)).abs().max() < tol + 2.0 ** -16 * ref.abs().max().item()
    # the pair is a proper split: lo is at most half a bf16 ulp of hi
    assert bool((ol.float().abs() <= oh.float().abs() * 2.0 ** -8 + 1e-30).all())


def test_gemm_split_conv_gate_residual():
    """k3 / k4-stride-2 / transposed convolutions, the gated epilogue and a split residual in split arithmetic."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(12)
    B, T, Cin, Cout, H = 5, 45, 128, 256, 1
    x = torch.randn(B, Cin, T, device="cuda", generator=g)
    Tp = T + 2 * H + ((T + 2 * H) % 2)
    xf = torch.zeros(B, Tp, Cin, device="cuda")
    xf[:, H:H + T] = x.transpose(1, 2)
    xh, xl = _split(xf)
    bias = torch.randn(Cout, device="cuda", generator=g)
    res = torch.randn(B, Tp, Cout, device="cuda", generator=g)
    rh, rl = _split(res)
    tapmat = lambda w: ops.split_weight(w.permute(0, 2, 1).reshape(w.shape[0], -1).contiguous())
    # k3 s1 p1 + split residual + LeakyReLU -> split output
    w3 = torch.randn(Cout, Cin, 3, device="cuda", generator=g) / (3 * Cin) ** 0.5
    oh, ol = torch.zeros(B, Tp, Cout, device="cuda", dtype=torch.bfloat16), torch.zeros(B, Tp, Cout, device="cuda", dtype=torch.bfloat16)
    ops.gemm(xh.view(-1), tapmat(w3), oh.view(-1), M=B * Tp, C_=Cin, taps=3, base_row=H - 1, rows_per_sample=Tp, t_valid=T,
             oss=Tp * Cout, ors=Cout, out_offset=H * Cout, bias=bias, act=ops.ACT_LEAKY, a_lo=xl.view(-1), out_lo=ol.view(-1),
             residual=rh.view(-1), residual_lo=rl.view(-1), rss=Tp * Cout, rrs=Cout, res_offset=H * Cout)
    ref = F.leaky_relu(F.conv1d(x.double(), w3.double(), bias.double(), padding=1)
                       + (rh.double() + rl.double())[:, H:H + T].transpose(1, 2), 0.2)
    got = (oh.double() + ol.double())[:, H:H + T].transpose(1, 2)
    assert (got - ref).abs().max() < 1e-4
    assert float(oh[:, :H].abs().max()) == 0 and float(ol[:, H + T:].abs().max()) == 0          # halo rows untouched
    # k4 s2 p1 -> fp32 output
    w4 = torch.randn(Cout, Cin, 4, device="cuda", generator=g) / (4 * Cin) ** 0.5
    To = (T - 2) // 2 + 1
    out2 = torch.zeros(B, To + 2, Cout, device="cuda")
    ops.gemm(xh.view(-1), tapmat(w4), out2.view(-1), M=B * (Tp // 2), C_=Cin, taps=4, s=2, base_row=0,
             rows_per_sample=Tp // 2, t_valid=To, oss=(To + 2) * Cout, ors=Cout, out_offset=Cout, bias=bias, a_lo=xl.view(-1))
    ref2 = F.conv1d(x.double(), w4.double(), bias.double(), stride=2, padding=1)
    assert (out2[:, 1:1 + To].transpose(1, 2).double() - ref2).abs().max() < 4e-5
    # transposed k4 s2 p1, channel-first fp32 output
    wt = torch.randn(Cin, Cout, 4, device="cuda", generator=g) / (2 * Cin) ** 0.5
    outT = torch.zeros(B, Cout, 2 * T, device="cuda")
    for ph, taps in enumerate(((3, 1), (2, 0))):
        wk = ops.split_weight(torch.cat([wt[:, :, taps[0]].t(), wt[:, :, taps[1]].t()], dim=1).contiguous())
        ops.gemm(xh.view(-1), wk, outT.view(-1), M=B * Tp, C_=Cin, taps=2, base_row=H - 1 + ph, rows_per_sample=Tp,
                 t_valid=T, oss=Cout * 2 * T, ors=2, ocs=2 * T, out_offset=ph, bias=bias, a_lo=xl.view(-1))
    refT = F.conv_transpose1d(x.double(), wt.double(), bias.double(), stride=2, padding=1)
    assert (outT.double() - refT).abs().max() < 4e-5
    # gated epilogue: interleaved (a, b) columns -> tanh(a) sigmoid(b), split output
    M, Cc = 700, 256
    a = torch.randn(M, Cc, device="cuda", generator=g)
    w = torch.randn(2 * Cc, Cc, device="cuda", generator=g) / Cc ** 0.5
    b2 = torch.randn(2 * Cc, device="cuda", generator=g)
    wi = torch.stack([w[:Cc], w[Cc:]], dim=1).reshape(2 * Cc, Cc).contiguous()
    bi = torch.stack([b2[:Cc], b2[Cc:]], dim=1).reshape(-1).contiguous()
    ah, al = _split(a)
    gh, gl = torch.zeros(M, Cc, device="cuda", dtype=torch.bfloat16), torch.zeros(M, Cc, device="cuda", dtype=torch.bfloat16)
    ops.gemm(ah, ops.split_weight(wi), gh, M=M, C_=Cc, bias=bi, act=ops.ACT_GATE, ors=Cc, a_lo=al, out_lo=gl)
    h = a.double() @ w.double().t() + b2.double()
    refg = torch.tanh(h[:, :Cc]) * torch.sigmoid(h[:, Cc:])
    assert ((gh.double() + gl.double()) - refg).abs().max() < 3e-5


@pytest.mark.parametrize("Nq,Nk", [(88, 88), (88, 22), (180, 180), (5, 3), (72, 18), (150, 150), (17, 256)])
def test_attention_f32(Nq, Nk):
    from probtalk_b200 import ops
    B, Hh, D = 3, 8, 64
    g = torch.Generator(device="cuda").manual_seed(Nq * 7 + Nk)
    q = torch.randn(B * Nq, Hh * D, device="cuda", generator=g)
    kv = torch.randn(B * Nk, 2 * Hh * D, device="cuda", generator=g)
    oh, ol = ops.attention_f32(q, kv[:, :Hh * D], kv[:, Hh * D:], B, Nq, Nk)
    qf = q.double().view(B, Nq, Hh, D).transpose(1, 2)
    kf = kv[:, :Hh * D].double().view(B, Nk, Hh, D).transpose(1, 2)
    vf = kv[:, Hh * D:].double().view(B, Nk, Hh, D).transpose(1, 2)
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) / math.sqrt(D), -1) @ vf).transpose(1, 2).reshape(B * Nq, Hh * D)
    assert ((oh.double() + ol.double()) - ref).abs().max() < 5e-5


def test_gather_conv1x1_vs_oracle():
    """quantize_all -> aft_vq_conv as a sum of G table rows: against the oracle's pq_gather followed by the 1x1 conv."""
    from probtalk_b200 import engine as E, ops
    vq = S.make_vq_state(0)
    W = E.VQVAEWeights(vq, "cuda", precision="bf16x3")
    g = torch.Generator().manual_seed(21)
    idx = torch.randint(0, 128, (3, 37, 4), generator=g)
    lat = O.pq_gather(idx, O.pq_codebooks(vq)).transpose(1, 2).contiguous()              # [B, 512, T']
    ref = F.conv1d(lat.double(), vq["decoder.aft_vq_conv.weight"].double(), vq["decoder.aft_vq_conv.bias"].double())
    ref = ref.transpose(1, 2)                                                          # [B, T', 1024]
    for dt in (torch.int64, torch.int32):
        hi, lo, bad = ops.pq_gather_conv1x1(idx.to(dt).cuda(), W.aft_table, W.aft_bias, split=True)
        assert int(bad) == 0
        got = hi.double().cpu() + lo.double().cpu()
        assert (got - ref).abs().max() < 2e-5 * max(1.0, ref.abs().max().item())       # 2^-16 of the value
        hi16, _, _ = ops.pq_gather_conv1x1(idx.to(dt).cuda(), W.aft_table, W.aft_bias, split=False)
        assert torch.equal(hi16, hi)
        # bf16 plane = correctly rounded fp32 sum (1 ulp of bf16 at most from the fp64 answer)
        assert (hi.double().cpu() - ref).abs().max() <= ref.abs().max().item() * 2.0 ** -8
    bad_idx = idx.clone()
    bad_idx[0, 0, 0], bad_idx[1, 2, 3] = 128, -1
    _, _, bad = ops.pq_gather_conv1x1(bad_idx.cuda(), W.aft_table, W.aft_bias)
    assert int(bad) == 2


# ------------------------------------------------------------------------------------------------ pipeline
@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD, map_location="cpu", weights_only=False)


@pytest.fixture(scope="module")
def states(gold):
    s = gold["seeds"]
    return S.make_vq_state(s["vq"]), S.make_predictor_state(s["pred"]), S.make_refiner_state(s["ref"])


@pytest.fixture(scope="module")
def weights(states):
    from probtalk_b200 import engine as E
    P = "bf16x3"
    return (E.VQVAEWeights(states[0], "cuda", precision=P), E.PredictorWeights(states[1], "cuda", P),
            E.RefinerWeights(states[2], "cuda", P))


def noise(T, B, N, seed):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(T, B, N, generator=g)
    gum = -torch.log(-torch.log(torch.rand(T, B, N, generator=g).clamp_(1e-10, 1 - 1e-7)))
    return u, gum


def _code_check(codes, ref_codes, z, z_ref, cbs, min_agree=0.999):
    """>= 99.9 % of the codes equal the reference's; each mismatch is a near-tie of the REFERENCE's own search whose fp64
    top-2 gap is no larger than what our z error can move (2 |dz| max|c|, plus the fp32 tie band of the search itself)."""
    diff = codes != ref_codes
    agree = 1.0 - diff.float().mean().item()
    assert agree >= min_agree, "only %.4f of the codes agree with the reference" % agree
    if bool(diff.any()):
        _, gap_ref = O.pq_search_fp64(z_ref, cbs)
        G, D = len(cbs), cbs[0].shape[1]
        dz = (z - z_ref).permute(0, 2, 1).reshape(z.shape[0], -1, G, D).norm(dim=-1)          # per (step, group) |dz|
        cmax = max(cb.norm(dim=1).max().item() for cb in cbs)
        eps = O.pq_tie_eps(z_ref, cbs)
        assert bool((gap_ref[diff] <= 4.0 * dz[diff] * cmax + eps[diff]).all())
    return agree


def test_cfg1_encode_index_exact_and_decode_tight(gold, states, weights):
    """BASELINE config 1 in the tight mode: z within ~1e-5 relative of the reference's, codes (near-)identical, decoded
    poses inside SURVEY's 1e-3 / 1e-4."""
    from probtalk_b200 import engine as E
    c = gold["cfg1"]
    x = S.synthetic_poses(1, 300, c["poses_seed"])
    enc = E.EncoderPlan(weights[0], 1, 300)
    codes = enc.run(x.cuda()).cpu()
    z = enc.z_channel_first().cpu()
    dz = (z - c["z"]).abs()
    print("tight encoder: max |dz| %.2e (max |z| %.2f), mean |dz| %.2e (mean |z| %.2f)"
          % (dz.max().item(), c["z"].abs().max().item(), dz.mean().item(), c["z"].abs().mean().item()))
    # 22 chained convolutions with K up to 4096: observed 5.6e-5 of max |z| / 5.3e-5 of mean |z| (bf16: 7e-3)
    assert dz.max() < 1.5e-4 * c["z"].abs().max().item() and dz.mean() < 1.5e-4 * c["z"].abs().mean().item(), \
        (dz.max().item(), dz.mean().item())
    _code_check(codes, c["codes"].long(), z, c["z"], O.pq_codebooks(states[0]))
    rec = E.DecoderPlan(weights[0], 1, enc.Tq).run(codes=c["codes"].long().cuda()).cpu()
    err = (rec - c["recon"]).abs()
    print("tight decode: max %.2e mean %.2e" % (err.max().item(), err.mean().item()))
    assert err.max() < POSE_MAX and err.mean() < POSE_MEAN, (err.max().item(), err.mean().item())


@pytest.mark.parametrize("scale", [1.0, 0.02])
def test_encode_near_tie_stress(states, scale):
    """Inputs scaled down so that z collapses towards the bias-only latent and the nearest-code gaps shrink (the
    near-tie stress of SURVEY cfg 2, here THROUGH the encoder): 64 clips x 37 steps x 4 groups = 9472 codes."""
    from probtalk_b200 import engine as E
    W = E.VQVAEWeights(states[0], "cuda")                     # default precision: the encoder is split-bf16 regardless
    B, T = 64, 300
    x = S.synthetic_poses(B, T, 4242) * scale
    z_ref, idx_ref, _ = O.vqvae_forward(x, states[0])
    enc = E.EncoderPlan(W, B, T)
    codes = enc.run(x.cuda()).cpu()
    z = enc.z_channel_first().cpu()
    agree = _code_check(codes, idx_ref, z, z_ref, O.pq_codebooks(states[0]))
    print("encode near-tie stress scale %g: agreement %.5f, max |dz| %.2e" % (scale, agree, (z - z_ref).abs().max().item()))


def test_cfg3_logits_tight(gold, states, weights):
    from probtalk_b200 import engine as E
    c = gold["cfg3"]
    g = torch.Generator().manual_seed(c["seed"])
    tokens = torch.randint(0, 129, (3, 22, 4), generator=g, dtype=torch.int32)
    cond = torch.randn(3, 768, 22, generator=g)
    plan = E.PredictorPlan(weights[1], weights[0], 3, 180)
    plan.labels.copy_(c["label"].int())
    plan.cond.load_f32(cond.transpose(1, 2).cuda())
    for fn, args, _ in plan.plan_cond.steps[-2:]:
        assert fn(*args, torch.cuda.current_stream().cuda_stream) == 0
    plan.cur.copy_(tokens.view(3, -1))
    plan.plan_iter.run()
    err = (plan.logits.view(3, 22, 4, 128).cpu() - c["logits"]).abs()
    print("tight logits: max %.2e mean %.2e" % (err.max().item(), err.mean().item()))
    assert err.max() < POSE_MAX and err.mean() < POSE_MEAN, (err.max().item(), err.mean().item())


def test_cfg4_predictor_refiner_tight(gold, states, weights):
    """config 4 slice in the tight mode, sampler noise injected: with fp32-grade logits the free-running sampler follows
    the reference's decisions (>= 99 % of the codes), identical codes give poses inside 1e-3 / 1e-4, and the refiner
    on the reference's predictor output matches the reference's refiner inside the same bound."""
    from probtalk_b200 import engine as E
    c = gold["cfg4"]
    vq, pst, rst = states
    B, T = c["B"], c["T"]
    aud = S.synthetic_audio(1, T, c["aud_seed"]).repeat(B, 1, 1)
    gt = S.synthetic_poses(B, T, c["gt_seed"])
    mask = torch.zeros(B, 1, T)
    mask[:, :, :30] = 1
    u, gum = noise(8, B, 88, c["noise_seed"])
    labels = torch.tensor([c["id"]] * B)
    plan = E.PredictorPlan(weights[1], weights[0], B, T)
    poses, codes = plan.run(aud.cuda(), gt.cuda(), mask.cuda(), labels.int().cuda(), noise=(u.cuda(), gum.cuda()))
    poses, codes = poses.cpu(), codes.cpu()
    agree = (codes.long() == c["codes"].long()).float().mean().item()
    print("free-running code agreement in the tight mode: %.4f" % agree)
    assert agree >= 0.99
    assert torch.equal(poses[:, :, :30], gt[:, :, :30])
    tf = O.predictor_decode(codes.long(), gt * mask, mask, vq, T)
    e = (poses - tf).abs()
    print("tight predictor poses (same codes): max %.2e mean %.2e" % (e.max().item(), e.mean().item()))
    assert e.max() < POSE_MAX and e.mean() < POSE_MEAN, (e.max().item(), e.mean().item())
    plan.cur.copy_(c["codes"].int().view(B, -1))
    e = (plan.run_decode().cpu() - c["poses"]).abs()
    assert e.max() < POSE_MAX and e.mean() < POSE_MEAN, (e.max().item(), e.mean().item())
    rp = E.RefinerPlan(weights[2], B, T, T)
    out = rp.run(aud.cuda(), gt.cuda(), mask.cuda(), c["poses"].cuda(), labels.int().cuda()).cpu()
    e = (out - c["refined"]).abs()
    print("tight refiner: max %.2e mean %.2e" % (e.max().item(), e.mean().item()))
    assert e.max() < POSE_MAX and e.mean() < POSE_MEAN, (e.max().item(), e.mean().item())
    assert torch.equal(out[:, :, :30], gt[:, :, :30])


def test_context_encoder_tight(states, weights):
    from probtalk_b200 import engine as E
    B, T = 3, 180
    aud, gt = S.synthetic_audio(B, T, 1003), S.synthetic_poses(B, T, 1004)
    mask = torch.zeros(B, 1, T)
    mask[:, :, 40:70] = 1
    plan = E.PredictorPlan(weights[1], weights[0], B, T)
    plan.aud.copy_(aud); plan.gt.copy_(gt); plan.mask.copy_(mask)
    plan.run_condition()
    _, cond_ref = O.predictor_condition(aud, gt, mask, states[1])
    cond = plan.cond.valid_f32().transpose(1, 2).cpu()
    e = (cond - cond_ref).abs()
    assert e.max() < 1e-3 and e.mean() < 1e-4, (e.max().item(), e.mean().item())


def test_wrapper_precision_switch(states):
    """args.precision = 'bf16x3' selects the tight mode through the reference-facing wrappers."""
    import types
    import probtalk_b200.nets as nets
    from bench import _config
    vq, pst, rst = states
    args = types.SimpleNamespace(gpu=0, infer=True, precision="bf16x3")
    vqt = nets.s2g_body_vqt(args, _config())
    vqt.load_state_dict({"VQ": vq})
    x = S.synthetic_poses(2, 300, 83)
    z_ref, idx_ref, rec_ref = O.vqvae_forward(x, vq)
    rec, _ = vqt.infer_on_audio(gt_poses=x.cuda(), aud=None, text=None, id=None, B=1)
    codes = vqt.get_codes(x.cuda()).cpu()
    assert (codes == idx_ref).float().mean() >= 0.999
    if bool((codes == idx_ref).all()):
        e = (rec.cpu() - rec_ref).abs()
        assert e.max() < POSE_MAX and e.mean() < POSE_MEAN
    with pytest.raises(ValueError):
        nets.s2g_body_vqt(types.SimpleNamespace(gpu=0, infer=True, precision="fp8"), _config())
