"""Host-side execution engine: turns reference-layout checkpoints into launch plans.

Everything numerical happens in ``libprobtalk_b200.so``; this module only
  * folds / re-lays-out weights once (BatchNorm folding, residual-conv merging, tap-major bf16
    matrices, gate interleaving, per-class style modulation tables),
  * owns the activation buffers of a (batch, frames) bucket -- channels-last bf16 rows
    ``[B, Tp, C]`` with zero halo rows so that a k-tap convolution is a plain GEMM over shifted rows,
  * records the launch sequence (pre-built ``pt_gemm_desc`` structs) and replays it.

Reference call stacks being replaced are cited per builder.  DESIGN.md describes the data layout.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _lib
from ._lib import GemmDesc, check
from . import ops
from .ops import ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_GATE

BF16 = torch.bfloat16
F32 = torch.float32


def _ceil(x, m):
    return (x + m - 1) // m * m


def _stream():
    return torch.cuda.current_stream().cuda_stream


# ----------------------------------------------------------------------------------------------
# weight preparation
# ----------------------------------------------------------------------------------------------

def _bn_scale_shift(sd, prefix, eps=1e-5):
    g, b = sd[prefix + ".weight"].float(), sd[prefix + ".bias"].float()
    mu, var = sd[prefix + ".running_mean"].float(), sd[prefix + ".running_var"].float()
    scale = g / torch.sqrt(var + eps)
    return scale, b - mu * scale


class Weight:
    """bf16 [Npad, taps*Cp] matrix + fp32 bias [Npad] on the device."""

    def __init__(self, mat, bias, taps, Cp, n_valid):
        self.mat, self.bias, self.taps, self.Cp, self.n_valid = mat, bias, taps, Cp, n_valid
        self.N = mat.shape[0]


def _finish(wm, bias, taps, Cp, device):
    """wm fp32 [N, taps, Cin] -> padded bf16 matrix."""
    N, _, Cin = wm.shape
    Npad = _ceil(N, 128)
    full = torch.zeros(Npad, taps, Cp, dtype=F32, device=wm.device)
    full[:N, :, :Cin] = wm
    b = torch.zeros(Npad, dtype=F32, device=wm.device)
    if bias is not None:
        b[:N] = bias
    return Weight(full.reshape(Npad, taps * Cp).to(device=device, dtype=BF16).contiguous(), b.to(device), taps, Cp, N)


def prep_conv(w, b, device, bn=None, extra=None, gate=False, cin_pad=None):
    """Conv1d weight [Cout, Cin, k] (+ folded BatchNorm (scale, shift), + a same-geometry residual
    conv (w_res, b_res) summed in) -> tap-major matrix W[n, j*Cp + c] = w[n, c, j].
    gate=True interleaves the two halves of the output channels (a0, b0, a1, b1, ...)."""
    w, b = w.float(), b.float()
    if bn is not None:
        scale, shift = bn
        w = w * scale[:, None, None]
        b = b * scale + shift
    if extra is not None:
        w = w + extra[0].float()
        b = b + extra[1].float()
    Cout, Cin, k = w.shape
    if gate:
        h = Cout // 2
        w = torch.stack([w[:h], w[h:]], dim=1).reshape(Cout, Cin, k)
        b = torch.stack([b[:h], b[h:]], dim=1).reshape(Cout)
    Cp = cin_pad or _ceil(Cin, 64)
    return _finish(w.permute(0, 2, 1).contiguous(), b, k, Cp, device)


def prep_convT(w, b, device, bn=None, extra=None):
    """ConvTranspose1d k4 s2 p1 weight [Cin, Cout, 4] -> two phase matrices (even / odd outputs):
    out[2t]   = x[t-1] W[:,:,3] + x[t] W[:,:,1],   out[2t+1] = x[t] W[:,:,2] + x[t+1] W[:,:,0]."""
    w, b = w.float(), b.float()
    if bn is not None:
        scale, shift = bn
        w = w * scale[None, :, None]
        b = b * scale + shift
    if extra is not None:
        w = w + extra[0].float()
        b = b + extra[1].float()
    Cin, Cout, k = w.shape
    assert k == 4
    Cp = _ceil(Cin, 64)
    phases = []
    for taps in ((3, 1), (2, 0)):
        wm = torch.stack([w[:, :, taps[0]].t(), w[:, :, taps[1]].t()], dim=1)     # [Cout, 2, Cin]
        phases.append(_finish(wm.contiguous(), b, 2, Cp, device))
    return phases


def prep_linear(w, b, device, cin_pad=None):
    """nn.Linear / 1x1 conv weight [out, in] (or [out, in, 1])."""
    if w.dim() == 3:
        w = w[:, :, 0]
    Cp = cin_pad or _ceil(w.shape[1], 64)
    return _finish(w.float().unsqueeze(1).contiguous(), None if b is None else b.float(), 1, Cp, device)


# ----------------------------------------------------------------------------------------------
# activation buffers and the launch plan
# ----------------------------------------------------------------------------------------------

class RowBuf:
    """Channels-last activation [B, Tp, C]: H zero halo rows, T valid rows, zero tail (Tp even)."""

    def __init__(self, B, T, Cc, H=1, dtype=BF16, device="cuda"):
        self.B, self.T, self.C, self.H = B, T, Cc, H
        self.Tp = _ceil(T + 2 * H, 2)
        self.t = torch.zeros(B, self.Tp, Cc, dtype=dtype, device=device)

    def valid(self):
        return self.t[:, self.H:self.H + self.T]


class Plan:
    """A recorded sequence of C-ABI launches over fixed buffers; ``run()`` replays it on the current stream."""

    def __init__(self, device):
        self.device = device
        self.steps = []
        self.keep = []          # keeps tensors referenced by raw pointers alive
        self.launches = 0

    # -- generic recorders
    def call(self, fn, *args):
        self.steps.append((fn, args))
        self.launches += 1

    def run(self):
        st = _stream()
        L = _lib.lib()
        for fn, args in self.steps:
            if fn is None:
                args[0]()
                continue
            rc = fn(*args, st)
            if rc != 0:
                check(rc, fn.__name__)

    def host(self, f):
        """A host-side callback between launches (e.g. drawing sampler noise with torch)."""
        self.steps.append((None, (f,)))

    # -- GEMM recorders
    def gemm(self, a_t, a_rows, a_pitch, w: Weight, out_t, *, M, C_, s=1, base_row=0, rows_per_sample, t_valid,
             oss, ors, ocs=1, out_off=0, act=ACT_NONE, slope=0.2, residual=None, rss=0, rrs=0, res_off=0, a_off=0,
             n_valid=None, use_bias=True):
        d = GemmDesc()
        d.a = a_t.data_ptr() + 2 * a_off
        d.a_rows, d.a_pitch = a_rows, a_pitch
        d.w = w.mat.data_ptr()
        assert C_ == w.Cp, (C_, w.Cp)
        d.M, d.N, d.C, d.taps, d.s = M, w.N, C_, w.taps, s
        d.base_row, d.rows_per_sample, d.t_valid = base_row, rows_per_sample, t_valid
        d.n_valid = w.n_valid if n_valid is None else n_valid
        d.out = out_t.data_ptr() + out_t.element_size() * out_off
        d.oss, d.ors, d.ocs = oss, ors, ocs
        d.out_f32 = int(out_t.dtype == F32)
        d.bias = w.bias.data_ptr() if use_bias else None
        if residual is not None:
            d.residual = residual.data_ptr() + residual.element_size() * res_off
            d.res_f32 = int(residual.dtype == F32)
        else:
            d.residual, d.res_f32 = None, 0
        d.rss, d.rrs = rss, rrs
        d.act, d.slope = act, slope
        self.keep += [a_t, w, out_t, residual, d]
        self.call(_lib.lib().pt_gemm_bf16, C.byref(d))

    def conv(self, x: RowBuf, w: Weight, out: RowBuf, *, k, stride=1, pad=None, act=ACT_NONE, slope=0.2,
             residual: RowBuf = None, out_c_off=0):
        """Conv1d over rows: k taps, stride 1 or 2 (ConvNormRelu / GatedConv / 1x1 convs)."""
        pad = (k - 1) // 2 if pad is None else pad
        assert w.taps == k and w.Cp == x.C
        gate = act == ACT_GATE
        n_out = w.n_valid // 2 if gate else w.n_valid
        assert out_c_off + n_out <= out.C
        if stride == 1:
            assert x.H >= pad and out.T == x.T + 2 * pad - k + 1
            M, rps, base = x.B * x.Tp, x.Tp, x.H - pad
        else:
            assert stride == 2 and k == 4 and pad == 1 and x.Tp % 2 == 0 and (x.H - 1) % 2 == 0
            assert out.T == (x.T - 2) // 2 + 1
            M, rps, base = x.B * x.Tp // 2, x.Tp // 2, (x.H - 1) // 2
        kw = {}
        if residual is not None:
            assert residual.T == out.T and residual.C == n_out
            kw = dict(residual=residual.t, rss=residual.Tp * residual.C, rrs=residual.C, res_off=residual.H * residual.C)
        self.gemm(x.t, x.B * x.Tp, x.C, w, out.t, M=M, C_=x.C, s=stride, base_row=base, rows_per_sample=rps,
                  t_valid=out.T, oss=out.Tp * out.C, ors=out.C, out_off=out.H * out.C + out_c_off, act=act, slope=slope,
                  **kw)

    def convT(self, x: RowBuf, phases, out: RowBuf, *, act=ACT_NONE, slope=0.2):
        """ConvTranspose1d k4 s2 p1 as two phase GEMMs (even / odd output rows)."""
        assert out.T == 2 * x.T and x.H >= 1
        for ph, w in enumerate(phases):
            assert w.Cp == x.C
            self.gemm(x.t, x.B * x.Tp, x.C, w, out.t, M=x.B * x.Tp, C_=x.C, s=1, base_row=x.H - 1 + ph,
                      rows_per_sample=x.Tp, t_valid=x.T, oss=out.Tp * out.C, ors=2 * out.C,
                      out_off=(out.H + ph) * out.C, act=act, slope=slope)

    def linear(self, a_t, M, w: Weight, out_t, *, out_cols=None, act=ACT_NONE, residual=None, rss=0, rrs=0,
               rows_per_sample=None, out_off=0, a_pitch=None, a_off=0):
        """Plain row-major GEMM: out[m, :] = act(a[m, :] @ W^T + b (+ residual))."""
        ors = out_cols if out_cols is not None else w.n_valid
        rps = rows_per_sample or max(M, 1)
        self.gemm(a_t, M, a_pitch or w.Cp, w, out_t, M=M, C_=w.Cp, rows_per_sample=rps, t_valid=rps,
                  oss=rps * ors, ors=ors, out_off=out_off, act=act, residual=residual, rss=rss, rrs=rrs, a_off=a_off)


# ----------------------------------------------------------------------------------------------
# PQ-VAE  (reference: nets/inpainting/vqvae_1d_sc.py:197-324, nets/spg/vqvae_modules.py:107-245)
# ----------------------------------------------------------------------------------------------

class VQVAEWeights:
    """Folded weights of VQVAE_SC(376, 512, 128, num_hiddens=1024, num_residual_layers=2)."""

    def __init__(self, sd, device, groups=4):
        dev = device
        sd = {k: v.to(dev) for k, v in sd.items() if torch.is_tensor(v)}
        self.device = dev

        def cnr(p, sample="none"):
            bn = _bn_scale_shift(sd, p + ".norm")
            if sample == "none":
                return prep_conv(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn)
            ex = (sd[p + ".residual_layer.weight"], sd[p + ".residual_layer.bias"])
            if sample == "down":
                return prep_conv(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn, extra=ex)
            return prep_convT(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn, extra=ex)

        def stack(p):
            return [cnr(p + "._layers.0"), cnr(p + "._layers.1"), cnr(p)]      # last: conv + norm of the stack itself

        self.enc_project = cnr("encoder.project")
        self.enc_stacks = [stack("encoder._enc_%d" % i) for i in (1, 2, 3, 4)]
        self.enc_downs = [cnr("encoder._down_%d" % i, "down") for i in (1, 2, 3)]
        self.pre_vq = prep_linear(sd["encoder.pre_vq_conv.weight"], sd["encoder.pre_vq_conv.bias"], dev)
        self.aft_vq = prep_linear(sd["decoder.aft_vq_conv.weight"], sd["decoder.aft_vq_conv.bias"], dev)
        self.dec_stacks = [stack("decoder._dec_%d" % i) for i in (1, 2, 3, 4)]
        self.dec_ups = [cnr("decoder._up_%d" % i, "up") for i in (2, 3, 4)]
        self.dec_project = prep_linear(sd["decoder.project.weight"], sd["decoder.project.bias"], dev)
        self.codebooks = torch.stack([sd["vq_layer.embeddings_%d" % g].float() for g in range(groups)]).contiguous()
        self.groups = groups
        self.in_dim = sd["encoder.project.conv.weight"].shape[1]


def _stack_steps(plan, x, ws, B, device):
    """Res_CNR_Stack.forward (vqvae_modules.py:238-245): 2 x (conv+BN+LeakyReLU), conv+BN, + x, ReLU."""
    t1 = RowBuf(B, x.T, x.C, 1, device=device)
    t2 = RowBuf(B, x.T, x.C, 1, device=device)
    out = RowBuf(B, x.T, x.C, 1, device=device)
    plan.conv(x, ws[0], t1, k=3, act=ACT_LEAKY)
    plan.conv(t1, ws[1], t2, k=3, act=ACT_LEAKY)
    plan.conv(t2, ws[2], out, k=3, act=ACT_RELU, residual=x)
    return out


class EncoderPlan:
    """EncoderSC.forward + ProductQuantization search (vqvae_1d_sc.py:220-235, vqvae_modules.py:470-491).
    poses fp32 [B, 376, T] -> z fp32 rows [B, T', 512] -> codes int64 [B, T', G]."""

    def __init__(self, W: VQVAEWeights, B, T):
        dev = W.device
        self.W, self.B, self.T = W, B, T
        p = self.plan = Plan(dev)
        L = _lib.lib()
        self.poses = torch.zeros(B, W.in_dim, T, dtype=F32, device=dev)
        x0 = RowBuf(B, T, _ceil(W.in_dim, 64), 1, device=dev)
        p.call(L.pt_pack_rows, self.poses.data_ptr(), None, None, 0, x0.t.data_ptr(), B, W.in_dim, T, x0.Tp, x0.H,
               x0.C, 0, 0)
        h = RowBuf(B, T, 256, 1, device=dev)
        p.conv(x0, W.enc_project, h, k=3, act=ACT_LEAKY)
        chans = [256, 512, 1024]
        for i in range(4):
            h = _stack_steps(p, h, W.enc_stacks[i], B, dev)
            if i < 3:
                d = RowBuf(B, (h.T - 2) // 2 + 1, chans[i], 1, device=dev)
                p.conv(h, W.enc_downs[i], d, k=4, stride=2, pad=1, act=ACT_LEAKY)
                h = d
        self.Tq = h.T
        self.z = RowBuf(B, h.T, 512, 0, dtype=F32, device=dev)
        p.conv(h, W.pre_vq, self.z, k=1)
        self.codes = torch.zeros(B, self.Tq, W.groups, dtype=torch.int64, device=dev)
        zt = self.z.t
        p.call(L.pt_pq_search, zt.data_ptr(), self.z.Tp * 512, 1, 512, B, self.Tq, W.groups, 128, 128,
               W.codebooks.data_ptr(), self.codes.data_ptr(), None, None, None)
        p.keep += [x0]

    def run(self, poses):
        self.poses.copy_(poses)
        self.plan.run()
        return self.codes

    def z_channel_first(self):
        """z in the reference's [B, 512, T'] layout (a view)."""
        return self.z.valid().transpose(1, 2)


class DecoderPlan:
    """quantize_all + DecoderSC.forward (vqvae_modules.py:557-564, vqvae_1d_sc.py:259-274).
    codes [B, T', G] -> poses fp32 [B, 376, 8 T'] (channel-first, the reference layout)."""

    def __init__(self, W: VQVAEWeights, B, Tq, codes=None):
        dev = W.device
        self.W, self.B, self.Tq = W, B, Tq
        p = self.plan = Plan(dev)
        L = _lib.lib()
        self.codes = codes if codes is not None else torch.zeros(B, Tq, W.groups, dtype=torch.int64, device=dev)
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)
        e = RowBuf(B, Tq, 512, 0, device=dev)
        p.call(L.pt_pq_gather, self.codes.data_ptr(), int(self.codes.dtype == torch.int64), W.codebooks.data_ptr(),
               e.t.data_ptr(), 1, e.Tp * 512, 1, 512, B, Tq, W.groups, 128, 128, self.bad.data_ptr())
        h = RowBuf(B, Tq, 1024, 1, device=dev)
        p.conv(e, W.aft_vq, h, k=1)
        chans = [512, 256, 256]
        for i in range(4):
            h = _stack_steps(p, h, W.dec_stacks[i], B, dev)
            if i < 3:
                u = RowBuf(B, 2 * h.T, chans[i], 1, device=dev)
                p.convT(h, W.dec_ups[i], u, act=ACT_LEAKY)
                h = u
        self.T_out = h.T
        self.out = torch.zeros(B, W.in_dim, self.T_out, dtype=F32, device=dev)
        w = W.dec_project
        p.gemm(h.t, B * h.Tp, h.C, w, self.out, M=B * h.Tp, C_=h.C, base_row=h.H, rows_per_sample=h.Tp, t_valid=h.T,
               oss=W.in_dim * self.T_out, ors=1, ocs=self.T_out)
        p.keep += [e]

    def run(self, codes=None):
        if codes is not None:
            self.codes.copy_(codes)
        self.plan.run()
        return self.out
