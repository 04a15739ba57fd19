"""Unit parity of the individual CUDA entry points (through the C ABI) against plain PyTorch fp32 on the same
(bf16-rounded) inputs.  These are floating-point kernels, so the checker is a torch fp32 reference; tolerances are
those of bf16 outputs (rel 2^-8) or fp32 accumulation-order noise."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _st():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1000, 256, 512), (333, 384, 1024), (4100, 1536, 512), (50, 2048, 512),
                                   (19000, 1024, 512), (40000, 512, 2048)])      # the last two take the cta_group::2 path
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_gemm_plain(M, N, K, out_dtype):
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=out_dtype)
    ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N)
    ref = a.float() @ w.float().t() + bias
    tol = 1e-4 if out_dtype == torch.float32 else 0.03
    assert (out.float() - ref).abs().max() < tol


def test_gemm_ragged_n_residual_and_acts():
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    M, K, N, Npad = 777, 256, 376, 384
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = torch.zeros(Npad, K, device="cuda")
    w[:N] = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    w = w.bfloat16()
    bias = torch.zeros(Npad, device="cuda")
    bias[:N] = torch.randn(N, device="cuda", generator=g)
    ref = a.float() @ w[:N].float().t() + bias[:N]
    for res_dtype in (None, torch.float32, torch.bfloat16):
        for out_dtype in (torch.float32, torch.bfloat16):
            for act, fn in ((ops.ACT_NONE, lambda x: x), (ops.ACT_RELU, torch.relu), (ops.ACT_LEAKY, lambda x: F.leaky_relu(x, 0.2))):
                res = None if res_dtype is None else torch.randn(M, N, device="cuda", generator=g).to(res_dtype)
                out = torch.full((M, N), 7.0, device="cuda", dtype=out_dtype)
                ops.gemm(a, w, out, M=M, C_=K, bias=bias, n_valid=N, ors=N, residual=res, rrs=N, act=act)
                r = fn(ref + (0 if res is None else res.float()))
                tol = 2e-4 if out_dtype == torch.float32 else 0.04
                assert (out.float() - r).abs().max() < tol, (res_dtype, out_dtype, act)


def test_gemm_conv_taps_strides_and_transposed():
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(2)
    B, T, Cin, Cout, H = 5, 45, 128, 256, 1
    x = torch.randn(B, Cin, T, device="cuda", generator=g)
    Tp = T + 2 * H + ((T + 2 * H) % 2)
    xb = torch.zeros(B, Tp, Cin, device="cuda", dtype=torch.bfloat16)
    xb[:, H:H + T] = x.transpose(1, 2).bfloat16()
    xr = xb[:, H:H + T].float().transpose(1, 2)
    bias = torch.randn(Cout, device="cuda", generator=g)
    # k3 s1 p1
    w3 = (torch.randn(Cout, Cin, 3, device="cuda", generator=g) / (3 * Cin) ** 0.5).bfloat16()
    out = torch.zeros(B, Tp, Cout, device="cuda", dtype=torch.bfloat16)
    ops.gemm(xb.view(-1), w3.permute(0, 2, 1).reshape(Cout, -1).contiguous(), out.view(-1), M=B * Tp, C_=Cin, taps=3,
             base_row=H - 1, rows_per_sample=Tp, t_valid=T, oss=Tp * Cout, ors=Cout, out_offset=H * Cout, bias=bias,
             act=ops.ACT_LEAKY)
    ref = F.leaky_relu(F.conv1d(xr, w3.float(), bias, padding=1), 0.2)
    assert (out[:, H:H + T].float().transpose(1, 2) - ref).abs().max() < 0.04
    assert float(out[:, :H].abs().max()) == 0 and float(out[:, H + T:].abs().max()) == 0        # halo rows untouched
    # k4 s2 p1
    w4 = (torch.randn(Cout, Cin, 4, device="cuda", generator=g) / (4 * Cin) ** 0.5).bfloat16()
    To = (T - 2) // 2 + 1
    out2 = torch.zeros(B, To + 2, Cout, device="cuda")
    ops.gemm(xb.view(-1), w4.permute(0, 2, 1).reshape(Cout, -1).contiguous(), out2.view(-1), M=B * (Tp // 2), C_=Cin,
             taps=4, s=2, base_row=0, rows_per_sample=Tp // 2, t_valid=To, oss=(To + 2) * Cout, ors=Cout, out_offset=Cout,
             bias=bias)
    ref2 = F.conv1d(xr, w4.float(), bias, stride=2, padding=1)
    assert (out2[:, 1:1 + To].transpose(1, 2) - ref2).abs().max() < 2e-4
    # transposed k4 s2 p1, channel-first fp32 output
    wt = (torch.randn(Cin, Cout, 4, device="cuda", generator=g) / (2 * Cin) ** 0.5).bfloat16()
    outT = torch.zeros(B, Cout, 2 * T, device="cuda")
    for ph, taps in enumerate(((3, 1), (2, 0))):
        wk = torch.cat([wt[:, :, taps[0]].t(), wt[:, :, taps[1]].t()], dim=1).contiguous()
        ops.gemm(xb.view(-1), wk, outT.view(-1), M=B * Tp, C_=Cin, taps=2, base_row=H - 1 + ph, rows_per_sample=Tp,
                 t_valid=T, oss=Cout * 2 * T, ors=2, ocs=2 * T, out_offset=ph, bias=bias)
    refT = F.conv_transpose1d(xr, wt.float(), bias, stride=2, padding=1)
    assert (outT - refT).abs().max() < 2e-4


def test_gemm_gated_epilogue():
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(4)
    M, Cc = 700, 256
    a = torch.randn(M, Cc, device="cuda", generator=g).bfloat16()
    w = (torch.randn(2 * Cc, Cc, device="cuda", generator=g) / Cc ** 0.5).bfloat16()
    b2 = torch.randn(2 * Cc, device="cuda", generator=g)
    wi = torch.stack([w[:Cc], w[Cc:]], dim=1).reshape(2 * Cc, Cc).contiguous()
    bi = torch.stack([b2[:Cc], b2[Cc:]], dim=1).reshape(-1).contiguous()
    out = torch.zeros(M, Cc, device="cuda", dtype=torch.bfloat16)
    ops.gemm(a, wi, out, M=M, C_=Cc, bias=bi, act=ops.ACT_GATE, ors=Cc)
    h = a.float() @ w.float().t() + b2
    ref = torch.tanh(h[:, :Cc]) * torch.sigmoid(h[:, Cc:])
    assert (out.float() - ref).abs().max() < 0.01


@pytest.mark.parametrize("Nq,Nk", [(88, 88), (88, 22), (180, 180), (5, 3), (17, 65), (48, 48), (200, 300)])
def test_attention(Nq, Nk):
    from probtalk_b200 import _lib
    B, Hh, D = 3, 8, 64
    g = torch.Generator(device="cuda").manual_seed(Nq * 7 + Nk)
    q = torch.randn(B * Nq, Hh * D, device="cuda", generator=g).bfloat16()
    kv = torch.randn(B * Nk, 2 * Hh * D, device="cuda", generator=g).bfloat16()
    out = torch.zeros(B * Nq, Hh * D, device="cuda", dtype=torch.bfloat16)
    rc = _lib.lib().pt_attention(q.data_ptr(), Hh * D, kv.data_ptr(), 2 * Hh * D, kv.data_ptr() + 2 * Hh * D, 2 * Hh * D,
                                 out.data_ptr(), Hh * D, B, Hh, D, Nq, Nk, 1.0 / math.sqrt(D), _st())
    _lib.check(rc)
    qf = q.float().view(B, Nq, Hh, D).transpose(1, 2)
    kf = kv[:, :Hh * D].float().view(B, Nk, Hh, D).transpose(1, 2)
    vf = kv[:, Hh * D:].float().view(B, Nk, Hh, D).transpose(1, 2)
    ref = torch.softmax(qf @ kf.transpose(-1, -2) / math.sqrt(D), -1) @ vf
    ref = ref.transpose(1, 2).reshape(B * Nq, Hh * D)
    assert (out.float() - ref).abs().max() < 0.03


@pytest.mark.parametrize("Nq,Nk", [(88, 88), (88, 22), (180, 180)])
def test_attention_many_units_repeated(Nq, Nk):
    """Bench-sized launches (every persistent CTA walks ~28 units, all buffer rings wrap many times), repeated: guards
    the mbarrier protocol of the pipelined kernel (a parity-aliasing bug only showed at this size) and run-to-run
    determinism."""
    from probtalk_b200 import _lib
    B, Hh, D = 512, 8, 64
    g = torch.Generator(device="cuda").manual_seed(Nq + Nk)
    q = torch.randn(B * Nq, Hh * D, device="cuda", generator=g).bfloat16()
    kv = torch.randn(B * Nk, 2 * Hh * D, device="cuda", generator=g).bfloat16()
    outs = []
    for _ in range(6):
        out = torch.zeros(B * Nq, Hh * D, device="cuda", dtype=torch.bfloat16)
        _lib.check(_lib.lib().pt_attention(q.data_ptr(), Hh * D, kv.data_ptr(), 2 * Hh * D, kv.data_ptr() + 2 * Hh * D,
                                           2 * Hh * D, out.data_ptr(), Hh * D, B, Hh, D, Nq, Nk, 1.0 / math.sqrt(D), _st()))
        outs.append(out)
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o, outs[0])
    sl = slice(0, 16 * Nq)                                  # reference on the first 16 samples and the last one
    for lo, hi, blo, bhi in ((0, 16 * Nq, 0, 16), ((B - 1) * Nq, B * Nq, B - 1, B)):
        nb = bhi - blo
        qf = q[lo:hi].float().view(nb, Nq, Hh, D).transpose(1, 2)
        kf = kv[blo * Nk:bhi * Nk, :Hh * D].float().view(nb, Nk, Hh, D).transpose(1, 2)
        vf = kv[blo * Nk:bhi * Nk, Hh * D:].float().view(nb, Nk, Hh, D).transpose(1, 2)
        ref = (torch.softmax(qf @ kf.transpose(-1, -2) / math.sqrt(D), -1) @ vf).transpose(1, 2).reshape(nb * Nq, Hh * D)
        assert (outs[0][lo:hi].float() - ref).abs().max() < 0.03


def test_layernorm_kernels_and_embed():
    from probtalk_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(9)
    M, N = 1000, 88
    x = torch.randn(M, 512, device="cuda", generator=g) * 3 + 1
    gamma, beta = torch.rand(512, device="cuda", generator=g) + 0.5, torch.randn(512, device="cuda", generator=g)
    o32, o16 = torch.zeros(M, 512, device="cuda"), torch.zeros(M, 512, device="cuda", dtype=torch.bfloat16)
    _lib.check(L.pt_layernorm512(x.data_ptr(), 1, None, None, gamma.data_ptr(), beta.data_ptr(), o32.data_ptr(), o16.data_ptr(), None, M, _st()))
    xb16 = x.bfloat16()
    o16b = torch.zeros_like(o16)
    _lib.check(L.pt_layernorm512(xb16.data_ptr(), 0, None, None, gamma.data_ptr(), beta.data_ptr(), None, o16b.data_ptr(), None, M, _st()))
    assert (o16b.float() - F.layer_norm(xb16.float(), (512,), gamma, beta, 1e-5)).abs().max() < 0.05
    res16 = torch.randn(M, 512, device="cuda", generator=g).bfloat16()
    acc = res16.clone()                       # residual aliases the output, like the engine's x = LN(x + s)
    _lib.check(L.pt_layernorm512(xb16.data_ptr(), 0, acc.data_ptr(), None, gamma.data_ptr(), beta.data_ptr(), None, acc.data_ptr(), None, M, _st()))
    assert (acc.float() - F.layer_norm(xb16.float() + res16.float(), (512,), gamma, beta, 1e-5)).abs().max() < 0.05
    ref = F.layer_norm(x, (512,), gamma, beta, 1e-5)
    assert (o32 - ref).abs().max() < 2e-5 and (o16.float() - ref).abs().max() < 0.05
    # style modulation: labels per sample of 8 rows, 4 classes
    mod = torch.randn(4, 1024, device="cuda", generator=g) * 0.5
    labels = (torch.arange(M // 8, device="cuda") % 4).int()
    xb = x.bfloat16()
    out = torch.zeros(M, 512, device="cuda", dtype=torch.bfloat16)
    _lib.check(L.pt_ln_mod_silu512(xb.data_ptr(), 0, gamma.data_ptr(), beta.data_ptr(), mod.data_ptr(), labels.data_ptr(), 8,
                                   out.data_ptr(), None, M, _st()))
    m = mod[labels.long()].repeat_interleave(8, 0)
    y = F.layer_norm(xb.float(), (512,), gamma, beta, 1e-5) * (1 + m[:, :512]) + m[:, 512:]
    assert (out.float() - F.silu(y)).abs().max() < 0.06
    # embedding + fp64 positional table, rounded like (x + pe).to(float32)
    emb = torch.randn(129, 512, device="cuda", generator=g)
    pe = torch.randn(N, 512, device="cuda", generator=g, dtype=torch.float64)
    tok = torch.randint(0, 129, (M // N * N,), device="cuda", generator=g, dtype=torch.int32)
    e32 = torch.zeros(tok.numel(), 512, device="cuda")
    bad = torch.zeros(1, device="cuda", dtype=torch.int32)
    eh, el = torch.zeros_like(e32, dtype=torch.bfloat16), torch.zeros_like(e32, dtype=torch.bfloat16)
    _lib.check(L.pt_embed_tokens512(tok.data_ptr(), emb.data_ptr(), 129, pe.data_ptr(), N, e32.data_ptr(), eh.data_ptr(),
                                    el.data_ptr(), tok.numel(), bad.data_ptr(), _st()))
    ref = (emb[tok.long()].double() + pe.repeat(tok.numel() // N, 1)).float()
    assert torch.equal(e32, ref) and int(bad) == 0
    # split-bf16 pair: hi = bf16(v), lo = bf16(v - hi); hi + lo carries 16 mantissa bits
    assert torch.equal(eh, ref.bfloat16()) and torch.equal(el, (ref - ref.bfloat16().float()).bfloat16())
    assert ((eh.float() + el.float()) - ref).abs().max() <= ref.abs().max() * 2.0 ** -16
    # split outputs / residual of the LayerNorm kernels (the tolerance-tight mode): fp32 in, (hi, lo) residual and out
    rh, rl = res16, (torch.randn(M, 512, device="cuda", generator=g) * 2.0 ** -9).bfloat16()
    oh, ol = torch.zeros_like(o16), torch.zeros_like(o16)
    _lib.check(L.pt_layernorm512(x.data_ptr(), 1, rh.data_ptr(), rl.data_ptr(), gamma.data_ptr(), beta.data_ptr(), None,
                                 oh.data_ptr(), ol.data_ptr(), M, _st()))
    ref = F.layer_norm(x + rh.float() + rl.float(), (512,), gamma, beta, 1e-5)
    assert ((oh.float() + ol.float()) - ref).abs().max() <= 2.0 ** -16 * ref.abs().max()       # 16 mantissa bits
    _lib.check(L.pt_ln_mod_silu512(x.data_ptr(), 1, gamma.data_ptr(), beta.data_ptr(), mod.data_ptr(), labels.data_ptr(), 8,
                                   oh.data_ptr(), ol.data_ptr(), M, _st()))
    y = F.layer_norm(x, (512,), gamma, beta, 1e-5) * (1 + m[:, :512]) + m[:, 512:]
    assert ((oh.float() + ol.float()) - F.silu(y)).abs().max() <= 2.0 ** -15 * F.silu(y).abs().max()


def test_pack_and_blend_match_torch():
    from probtalk_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(10)
    B, Cc, T, Tpred = 3, 376, 180, 176
    src = torch.randn(B, Cc, T, device="cuda", generator=g)
    pred = torch.randn(B, Cc, Tpred, device="cuda", generator=g)
    mask = (torch.rand(B, 1, T, device="cuda", generator=g) < 0.3).float()
    H, Cp = 1, 384
    Tp = T + 2
    dst = torch.zeros(B, Tp, Cp, device="cuda", dtype=torch.bfloat16)
    _lib.check(L.pt_pack_rows(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), Tpred, dst.data_ptr(), None, B, Cc, T, Tp, H, Cp, 0, 2, _st()))
    interp = F.interpolate(pred, size=T, mode="linear")
    ref = src * mask + interp * (1 - mask)
    got = dst[:, H:H + T, :Cc].float().transpose(1, 2)
    assert (got - ref.bfloat16().float()).abs().max() < 0.02
    assert float(dst[:, :, Cc:].abs().max()) == 0 and float(dst[:, 0].abs().max()) == 0
    _lib.check(L.pt_pack_mask(mask.data_ptr(), dst.data_ptr(), B, T, Tp, H, Cp, Cc, 1, _st()))
    oh = F.one_hot(mask.long(), 2).squeeze(1).float()
    assert torch.equal(dst[:, H:H + T, Cc:Cc + 2].float(), oh)
    out = torch.zeros(B, Cc, T, device="cuda")
    _lib.check(L.pt_interp_blend(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), Tpred, out.data_ptr(), B, Cc, T, 2, _st()))
    assert (out - (src * mask * mask + interp * (1 - mask))).abs().max() < 1e-5


@pytest.mark.parametrize("B,Cc,T", [(5, 376, 180), (2, 129, 64), (3, 40, 36)])
def test_pack_and_blend_same_length_paths(B, Cc, T):
    """The slab pack kernel (modes 0/1/2) and the float4 blend path (pred already at T frames) vs torch."""
    from probtalk_b200 import _lib
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + Cc + T)
    src = torch.randn(B, Cc, T, device="cuda", generator=g)
    pred = torch.randn(B, Cc, T, device="cuda", generator=g)
    mask = (torch.rand(B, 1, T, device="cuda", generator=g) < 0.4).float()
    H, Cp, c_off = 2, ((Cc + 63) // 64) * 64 + 64, 8
    Tp = T + 5
    for mode, ref in ((0, src), (1, src * mask), (2, src * mask + pred * (1 - mask))):
        dst = torch.zeros(B, Tp, Cp, device="cuda", dtype=torch.bfloat16)
        lo = torch.zeros_like(dst)
        _lib.check(L.pt_pack_rows(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), T, dst.data_ptr(), lo.data_ptr(), B, Cc, T,
                                  Tp, H, Cp, c_off, mode, _st()))
        got = dst[:, H:H + T, c_off:c_off + Cc].float().transpose(1, 2)
        assert torch.equal(got, ref.bfloat16().float())
        got_lo = lo[:, H:H + T, c_off:c_off + Cc].float().transpose(1, 2)
        assert torch.equal(got_lo, (ref - ref.bfloat16().float()).bfloat16().float())      # the split pair's lo plane
        chk = dst.clone()
        chk[:, H:H + T, c_off:c_off + Cc] = 0
        assert float(chk.abs().max()) == 0                  # halo rows and padding channels untouched
    for pw in (1, 2):
        out = torch.zeros(B, Cc, T, device="cuda")
        _lib.check(L.pt_interp_blend(src.data_ptr(), mask.data_ptr(), pred.data_ptr(), T, out.data_ptr(), B, Cc, T, pw, _st()))
        ref = src * mask ** pw + pred * (1 - mask)
        assert (out - ref).abs().max() < 1e-6


@pytest.mark.parametrize("M,K", [(1000, 512), (128, 2048), (90112 // 8, 512), (90112 // 8 + 77, 2048), (22000, 1024)])
@pytest.mark.parametrize("mode", [1, 2])
def test_gemm_rownorm_cluster_epilogue(M, K, mode):
    """Fused LayerNorm epilogue (2-CTA cluster exchanging row statistics through DSMEM) vs torch."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + K + mode)
    N = 512
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    v = a.float() @ w.float().t() + bias
    if mode == 1:
        rows = 8
        na = torch.rand(4, N, device="cuda", generator=g) + 0.5
        nb = torch.randn(4, N, device="cuda", generator=g)
        labels = (torch.arange((M + rows - 1) // rows, device="cuda") % 4).int()
        out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N, row_norm=1, norm_a=na, norm_b=nb, labels=labels, label_rows=rows)
        cls = labels.long().repeat_interleave(rows)[:M]
        y = F.layer_norm(v, (N,), None, None, 1e-5) * na[cls] + nb[cls]
        ref = F.silu(y)
    else:
        na = torch.rand(1, N, device="cuda", generator=g) + 0.5
        nb = torch.randn(1, N, device="cuda", generator=g)
        x = torch.randn(M, N, device="cuda", generator=g).bfloat16()
        out = x.clone()                                   # residual aliases the output (x = LN(x + s))
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N, residual=out, rrs=N, row_norm=2, norm_a=na, norm_b=nb)
        ref = F.layer_norm(v + x.float(), (N,), na[0], nb[0], 1e-5)
    assert (out.float() - ref).abs().max() < 0.06


def test_gemm_rownorm_residual_many_launches():
    """Soak of the residual-through-the-tensor-core flavour (row_norm 2, bf16 residual): a few hundred back-to-back launches
    over many tiles per CTA must all complete (an earlier version dead-locked about once in 10^5 launches: a helper warp kept
    in step with the operand ring through mbarrier parity waits and could be lapped by the producer) and give the same bits."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(77)
    M, N, K = 90112, 512, 512
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    na, nb = torch.rand(1, N, device="cuda", generator=g) + 0.5, torch.randn(1, N, device="cuda", generator=g)
    x = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    outs = []
    for it in range(400):
        out = torch.empty_like(x)
        ops.gemm(a, w, out, M=M, C_=K, bias=bias, ors=N, residual=x, rrs=N, row_norm=2, norm_a=na, norm_b=nb)
        if it in (0, 399):
            outs.append(out)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1])
    ref = F.layer_norm(a[:256].float() @ w.float().t() + bias + x[:256].float(), (N,), na[0], nb[0], 1e-5)
    assert (outs[1][:256].float() - ref).abs().max() < 0.06
