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
from .ops import ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_GATE, ACT_GELU

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
    """bf16 [Npad, taps*Cp] matrix + fp32 bias [Npad] on the device.  split=True: the matrix is [Npad, 2*taps*Cp], every
    row its hi plane followed by its lo plane (w = hi + lo to 16 mantissa bits), for the split-bf16 arithmetic mode."""

    def __init__(self, mat, bias, taps, Cp, n_valid, cin=None, split=False):
        self.mat, self.bias, self.taps, self.Cp, self.n_valid = mat, bias, taps, Cp, n_valid
        self.N = mat.shape[0]
        self.cin = cin if cin is not None else Cp       # true (unpadded) input channels per tap
        self.split = split


def _finish(wm, bias, taps, Cp, device, split=False):
    """wm fp32 [N, taps, Cin] -> padded bf16 matrix (or the [hi | lo] pair of matrices)."""
    N, _, Cin = wm.shape
    Npad = _ceil(N, 128)
    full = torch.zeros(Npad, taps, Cp, dtype=F32, device=wm.device)
    full[:N, :, :Cin] = wm
    b = torch.zeros(Npad, dtype=F32, device=wm.device)
    if bias is not None:
        b[:N] = bias
    full = full.reshape(Npad, taps * Cp)
    hi = full.to(BF16)
    if split:
        lo = (full - hi.float()).to(BF16)
        mat = torch.cat([hi, lo], dim=1)
    else:
        mat = hi
    return Weight(mat.to(device).contiguous(), b.to(device), taps, Cp, N, Cin, split)


def prep_conv(w, b, device, bn=None, extra=None, gate=False, cin_pad=None, split=False):
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
    return _finish(w.permute(0, 2, 1).contiguous(), b, k, Cp, device, split)


def prep_convT(w, b, device, bn=None, extra=None, split=False):
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
        phases.append(_finish(wm.contiguous(), b, 2, Cp, device, split))
    return phases


def prep_linear(w, b, device, cin_pad=None, split=False):
    """nn.Linear / 1x1 conv weight [out, in] (or [out, in, 1])."""
    if w.dim() == 3:
        w = w[:, :, 0]
    Cp = cin_pad or _ceil(w.shape[1], 64)
    return _finish(w.float().unsqueeze(1).contiguous(), None if b is None else b.float(), 1, Cp, device, split)


# ----------------------------------------------------------------------------------------------
# activation buffers and the launch plan
# ----------------------------------------------------------------------------------------------

class RowBuf:
    """Channels-last activation [B, Tp, C]: H zero halo rows, T valid rows, zero tail (Tp even)."""

    def __init__(self, B, T, Cc, H=1, dtype=BF16, device="cuda", even=True, split=False):
        self.B, self.T, self.C, self.H = B, T, Cc, H
        # an even row count lets stride-2 convolutions view the buffer as [rows / 2, 2 C]; even=False (no halo, never
        # read with a stride) keeps exactly T rows per sample so that the buffer is a flat [B * T, C] matrix
        self.Tp = _ceil(T + 2 * H, 2) if even else T + 2 * H
        # split (tolerance-tight mode): two bf16 planes, value = hi + lo
        split = split and dtype == BF16
        self.planes = torch.zeros(2 if split else 1, B, self.Tp, Cc, dtype=dtype, device=device)
        self.t = self.planes[0]
        self.lo = self.planes[1] if split else None

    def valid(self):
        return self.t[:, self.H:self.H + self.T]

    def load_f32(self, x):
        """Set the valid rows from an fp32 tensor [B, T, C] (test / bring-up hook; the lo plane gets x - bf16(x))."""
        hi = x.to(self.t.dtype)
        self.valid().copy_(hi)
        if self.lo is not None:
            self.lo[:, self.H:self.H + self.T].copy_((x.float() - hi.float()).to(self.lo.dtype))

    def valid_f32(self):
        """The valid rows as fp32 (hi + lo for a split buffer)."""
        v = self.valid().float()
        return v if self.lo is None else v + self.lo[:, self.H:self.H + self.T].float()


class Mat:
    """Flat activation matrix [rows, C]: bf16, a split-bf16 (hi, lo) pair, or fp32 (lo = None)."""

    def __init__(self, rows, Cc, device, split=False, dtype=BF16):
        split = split and dtype == BF16
        self.planes = torch.zeros(2 if split else 1, rows, Cc, dtype=dtype, device=device)
        self.t = self.planes[0]
        self.lo = self.planes[1] if split else None

    def f32(self):
        return self.t.float() if self.lo is None else self.t.float() + self.lo.float()


def _planes(x):
    """(hi / only plane, lo plane or None) of a tensor, Mat or RowBuf."""
    if x is None or torch.is_tensor(x):
        return x, None
    return x.t, x.lo


def _ptr(t):
    return None if t is None else t.data_ptr()


class PlanCache:
    """bucket key -> plan with least-recently-used eviction.  A plan owns every activation buffer of its (batch, frames)
    bucket (gigabytes at large batches) and ragged tail windows create one bucket per length, so the caches of the
    wrappers / driver are bounded: beyond ``capacity`` buckets the one used longest ago is dropped (its buffers and
    CUDA graph are freed when the last reference goes)."""

    def __init__(self, capacity=8):
        from collections import OrderedDict
        self.capacity, self._d = capacity, OrderedDict()

    def get(self, key, build):
        d = self._d
        if key in d:
            d.move_to_end(key)
            return d[key]
        plan = d[key] = build()
        while len(d) > self.capacity:
            d.popitem(last=False)
        return plan

    def __len__(self):
        return len(self._d)

    def __contains__(self, key):
        return key in self._d


class Plan:
    """A recorded sequence of C-ABI launches over fixed buffers; ``run()`` replays it on the current stream."""

    def __init__(self, device):
        self.device = device
        self.steps = []
        self.keep = []          # keeps tensors referenced by raw pointers alive
        self.launches = 0

    # -- generic recorders
    def call(self, fn, *args, flops=0.0):
        self.steps.append((fn, args, flops))
        self.launches += 1

    def run(self):
        # the library configures and launches on the CURRENT device: make it the one that owns the plan's buffers
        with torch.cuda.device(self.device):
            st = _stream()
            for fn, args, _ in self.steps:
                rc = fn(*args, st)
                if rc != 0:
                    check(rc, fn.__name__)

    def run_timed(self, records):
        """Replay with a CUDA event pair around every launch; appends (name, start, end, flops, gemm_key) to records."""
        st = _stream()
        for fn, args, flops in self.steps:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args, st)
            if rc != 0:
                check(rc, fn.__name__)
            e1.record()
            key = None
            if fn.__name__ == "pt_gemm_bf16":       # (M, N, K, row_norm, has_residual): lets a caller match launches to profiles
                d = args[0]._obj
                key = (d.M, d.n_valid, d.taps * d.C, d.row_norm, int(bool(d.residual)))
            records.append((fn.__name__, e0, e1, flops, key))

    # -- GEMM recorders
    def gemm(self, a_t, a_rows, a_pitch, w: Weight, out_t, *, M, C_, s=1, base_row=0, rows_per_sample, t_valid,
             oss, ors, ocs=1, out_off=0, act=ACT_NONE, slope=0.2, residual=None, rss=0, rrs=0, res_off=0, a_off=0,
             n_valid=None, use_bias=True, row_norm=0, norm_a=None, norm_b=None, labels=None, label_rows=1,
             a_lo=None, out_lo=None, residual_lo=None, a_col_table=None):
        d = GemmDesc()
        d.a = a_t.data_ptr() + 2 * a_off
        if a_col_table is not None:
            # extra A column offset per 128-column output tile (grouped convolutions): a host array the call reads
            tab = (C.c_int * len(a_col_table))(*a_col_table)
            d.a_col_table = C.cast(tab, C.c_void_p)
            self.keep.append(tab)
        if w.split:
            # split-bf16 arithmetic: every bf16 operand / result is a (hi, lo) plane pair
            assert a_lo is not None and row_norm == 0, "a split weight needs a split A operand"
            assert out_t.dtype == F32 or out_lo is not None
            assert residual is None or residual.dtype == F32 or residual_lo is not None
            d.split = 1
            d.a_lo = a_lo.data_ptr() + 2 * a_off
            d.out_lo = out_lo.data_ptr() + 2 * out_off if (out_lo is not None and out_t.dtype != F32) else None
            d.residual_lo = residual_lo.data_ptr() + 2 * res_off if (residual_lo is not None and residual.dtype != F32) else None
        # (a bf16 weight reads only the hi plane of a split operand)
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
        d.row_norm, d.label_rows = row_norm, label_rows
        d.norm_a = norm_a.data_ptr() if norm_a is not None else None
        d.norm_b = norm_b.data_ptr() if norm_b is not None else None
        d.norm_classes = (norm_a.numel() // 512) if norm_a is not None else 0
        d.labels = labels.data_ptr() if labels is not None else None
        self.keep += [a_t, w, out_t, residual, d, norm_a, norm_b, labels, a_lo, out_lo, residual_lo]
        # algorithmic FLOPs: real output rows x real output channels x real input channels (no padding / halo credit)
        flops = 2.0 * M * (t_valid / float(rows_per_sample)) * d.n_valid * w.taps * w.cin
        self.call(_lib.lib().pt_gemm_bf16, C.byref(d), flops=flops)

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
            kw = dict(residual=residual.t, rss=residual.Tp * residual.C, rrs=residual.C, res_off=residual.H * residual.C,
                      residual_lo=residual.lo)
        self.gemm(x.t, x.B * x.Tp, x.C, w, out.t, M=M, C_=x.C, s=stride, base_row=base, rows_per_sample=rps,
                  t_valid=out.T, oss=out.Tp * out.C, ors=out.C, out_off=out.H * out.C + out_c_off, act=act, slope=slope,
                  a_lo=x.lo, out_lo=out.lo, **kw)

    def convT(self, x: RowBuf, phases, out: RowBuf, *, act=ACT_NONE, slope=0.2):
        """ConvTranspose1d k4 s2 p1 as two phase GEMMs (even / odd output rows)."""
        assert out.T == 2 * x.T and x.H >= 1
        for ph, w in enumerate(phases):
            assert w.Cp == x.C
            self.gemm(x.t, x.B * x.Tp, x.C, w, out.t, M=x.B * x.Tp, C_=x.C, s=1, base_row=x.H - 1 + ph,
                      rows_per_sample=x.Tp, t_valid=x.T, oss=out.Tp * out.C, ors=2 * out.C,
                      out_off=(out.H + ph) * out.C, act=act, slope=slope, a_lo=x.lo, out_lo=out.lo)

    def linear(self, a, M, w: Weight, out, *, out_cols=None, act=ACT_NONE, residual=None, rss=0, rrs=0,
               rows_per_sample=None, out_off=0, a_pitch=None, a_off=0, **norm):
        """Plain row-major GEMM: out[m, :] = act(a[m, :] @ W^T + b (+ residual)).
        a / out / residual: tensors, or Mat / RowBuf objects (whose lo planes ride along in split arithmetic)."""
        a_t, a_lo = _planes(a)
        out_t, out_lo = _planes(out)
        res_t, res_lo = _planes(residual)
        ors = out_cols if out_cols is not None else w.n_valid
        rps = rows_per_sample or max(M, 1)
        self.gemm(a_t, M, a_pitch or w.Cp, w, out_t, M=M, C_=w.Cp, rows_per_sample=rps, t_valid=rps,
                  oss=rps * ors, ors=ors, out_off=out_off, act=act, residual=res_t, rss=rss, rrs=rrs, a_off=a_off,
                  a_lo=a_lo if w.split else None, out_lo=out_lo, residual_lo=res_lo, **norm)


# ----------------------------------------------------------------------------------------------
# PQ-VAE  (reference: nets/inpainting/vqvae_1d_sc.py:197-324, nets/spg/vqvae_modules.py:107-245)
# ----------------------------------------------------------------------------------------------

PRECISIONS = ("bf16", "bf16x3")


def _check_precision(precision):
    if precision not in PRECISIONS:
        raise ValueError("precision must be one of %s (got %r)" % (PRECISIONS, precision))
    return precision == "bf16x3"


class VQVAEWeights:
    """Folded weights of VQVAE_SC(376, 512, 128, num_hiddens=1024, num_residual_layers=2).

    The ENCODER is always prepared for split-bf16 arithmetic (A_hi W_hi + A_hi W_lo + A_lo W_hi, fp32-grade): its
    output z feeds a nearest-code search, where a bf16-sized error flips indices, and it is not on the throughput path.
    The decoder follows ``precision`` ("bf16" or "bf16x3")."""

    def __init__(self, sd, device, groups=4, precision="bf16"):
        dev = device
        sd = {k: v.to(dev) for k, v in sd.items() if torch.is_tensor(v)}
        self.device = dev
        self.precision = precision
        self.precise = _check_precision(precision)

        def cnr(p, sample, split):
            bn = _bn_scale_shift(sd, p + ".norm")
            if sample == "none":
                return prep_conv(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn, split=split)
            ex = (sd[p + ".residual_layer.weight"], sd[p + ".residual_layer.bias"])
            if sample == "down":
                return prep_conv(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn, extra=ex, split=split)
            return prep_convT(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, bn=bn, extra=ex, split=split)

        def stack(p, split):        # last: conv + norm of the stack itself
            return [cnr(p + "._layers.0", "none", split), cnr(p + "._layers.1", "none", split), cnr(p, "none", split)]

        ds = self.precise
        self.enc_project = cnr("encoder.project", "none", True)
        self.enc_stacks = [stack("encoder._enc_%d" % i, True) for i in (1, 2, 3, 4)]
        self.enc_downs = [cnr("encoder._down_%d" % i, "down", True) for i in (1, 2, 3)]
        self.pre_vq = prep_linear(sd["encoder.pre_vq_conv.weight"], sd["encoder.pre_vq_conv.bias"], dev, split=True)
        self.aft_vq = prep_linear(sd["decoder.aft_vq_conv.weight"], sd["decoder.aft_vq_conv.bias"], dev, split=ds)
        self.dec_stacks = [stack("decoder._dec_%d" % i, ds) for i in (1, 2, 3, 4)]
        self.dec_ups = [cnr("decoder._up_%d" % i, "up", ds) for i in (2, 3, 4)]
        self.dec_project = prep_linear(sd["decoder.project.weight"], sd["decoder.project.bias"], dev, split=ds)
        self.codebooks = torch.stack([sd["vq_layer.embeddings_%d" % g].float() for g in range(groups)]).contiguous()
        self.groups = groups
        self.n_codes, self.code_dim = self.codebooks.shape[1], self.codebooks.shape[2]
        self.in_dim = sd["encoder.project.conv.weight"].shape[1]
        # gather fused into aft_vq_conv (vqvae_modules.py:557-564 -> vqvae_1d_sc.py:259-261): the 1x1 conv of a
        # concatenation of G codebook rows is a sum of G table rows, T_g = C_g . W[:, gD:(g+1)D]^T  (fp64 product,
        # stored fp32; G x K x 1024 = 2 MB)
        wa = sd["decoder.aft_vq_conv.weight"].double()
        wa = wa[:, :, 0] if wa.dim() == 3 else wa                                   # [N, G*D]
        D = self.code_dim
        self.aft_table = torch.stack([self.codebooks[g].double() @ wa[:, g * D:(g + 1) * D].t()
                                      for g in range(groups)]).float().contiguous()        # [G, K, N]
        self.aft_bias = sd["decoder.aft_vq_conv.bias"].float().contiguous()


def _stack_steps(plan, x, ws, B, device):
    """Res_CNR_Stack.forward (vqvae_modules.py:238-245): 2 x (conv+BN+LeakyReLU), conv+BN, + x, ReLU."""
    sp = x.lo is not None
    t1 = RowBuf(B, x.T, x.C, 1, device=device, split=sp)
    t2 = RowBuf(B, x.T, x.C, 1, device=device, split=sp)
    out = RowBuf(B, x.T, x.C, 1, device=device, split=sp)
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
        # the encoder runs in split-bf16 arithmetic whatever the decoder's precision (see VQVAEWeights)
        x0 = RowBuf(B, T, _ceil(W.in_dim, 64), 1, device=dev, split=True)
        p.call(L.pt_pack_rows, self.poses.data_ptr(), None, None, 0, x0.t.data_ptr(), x0.lo.data_ptr(), B, W.in_dim, T,
               x0.Tp, x0.H, x0.C, 0, 0)
        h = RowBuf(B, T, 256, 1, device=dev, split=True)
        p.conv(x0, W.enc_project, h, k=3, act=ACT_LEAKY)
        self.stages = [("project", h)]
        chans = [256, 512, 1024]
        for i in range(4):
            h = _stack_steps(p, h, W.enc_stacks[i], B, dev)
            self.stages.append(("_enc_%d" % (i + 1), h))
            if i < 3:
                d = RowBuf(B, (h.T - 2) // 2 + 1, chans[i], 1, device=dev, split=True)
                p.conv(h, W.enc_downs[i], d, k=4, stride=2, pad=1, act=ACT_LEAKY)
                h = d
                self.stages.append(("_down_%d" % (i + 1), h))
        self.Tq = h.T
        self.z = RowBuf(B, h.T, 512, 0, dtype=F32, device=dev)
        p.conv(h, W.pre_vq, self.z, k=1)
        self.codes = torch.zeros(B, self.Tq, W.groups, dtype=torch.int64, device=dev)
        zt = self.z.t
        p.call(L.pt_pq_search, zt.data_ptr(), self.z.Tp * 512, 1, 512, B, self.Tq, W.groups, W.n_codes, W.code_dim,
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

    def __init__(self, W: VQVAEWeights, B, Tq, codes=None, from_latents=False):
        dev = W.device
        self.W, self.B, self.Tq = W, B, Tq
        p = self.plan = Plan(dev)
        L = _lib.lib()
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)
        sp = W.precise
        h = RowBuf(B, Tq, W.aft_table.shape[2], 1, device=dev, split=sp)
        if from_latents:
            # VQ.decode(latents, None) on arbitrary [B, 512, T'] fp32 latents (scripts/demo.py:152-153)
            self.codes = None
            Cl = W.groups * W.code_dim
            e = RowBuf(B, Tq, Cl, 0, device=dev, split=sp)
            self.latents = torch.zeros(B, Cl, Tq, dtype=F32, device=dev)
            p.call(L.pt_pack_rows, self.latents.data_ptr(), None, None, 0, e.t.data_ptr(), _ptr(e.lo), B, Cl, Tq, e.Tp, 0,
                   Cl, 0, 0)
            p.conv(e, W.aft_vq, h, k=1)
            p.keep += [e]
        else:
            # codebook gather + dequant fused into the decoder's first conv: G table rows summed per output row
            self.codes = codes if codes is not None else torch.zeros(B, Tq, W.groups, dtype=torch.int64, device=dev)
            p.call(L.pt_pq_gather_conv1x1, self.codes.data_ptr(), int(self.codes.dtype == torch.int64),
                   W.aft_table.data_ptr(), W.aft_bias.data_ptr(), h.t.data_ptr() + 2 * h.H * h.C,
                   (h.lo.data_ptr() + 2 * h.H * h.C) if sp else None, h.Tp * h.C, h.C, B, Tq, W.groups, W.n_codes, h.C,
                   self.bad.data_ptr(), flops=2.0 * B * Tq * h.C * W.groups * W.code_dim)
        self.stages = [("aft_vq_conv", h)]
        chans = [512, 256, 256]
        for i in range(4):
            h = _stack_steps(p, h, W.dec_stacks[i], B, dev)
            self.stages.append(("_dec_%d" % (i + 1), h))
            if i < 3:
                u = RowBuf(B, 2 * h.T, chans[i], 1, device=dev, split=sp)
                p.convT(h, W.dec_ups[i], u, act=ACT_LEAKY)
                h = u
                self.stages.append(("_up_%d" % (i + 2), h))
        self.T_out = h.T
        self.out = torch.zeros(B, W.in_dim, self.T_out, dtype=F32, device=dev)
        w = W.dec_project
        p.gemm(h.t, B * h.Tp, h.C, w, self.out, M=B * h.Tp, C_=h.C, base_row=h.H, rows_per_sample=h.Tp, t_valid=h.T,
               oss=W.in_dim * self.T_out, ors=1, ocs=self.T_out, a_lo=h.lo)

    def run(self, codes=None, latents=None):
        if codes is not None:
            self.codes.copy_(codes)
        if latents is not None:
            self.latents.copy_(latents)
        self.plan.run()
        return self.out


# ----------------------------------------------------------------------------------------------
# transformer decoder stack shared by predictor and refiner
# (reference: DeepSupervisionLayer / StylizationBlock_V3, nets/inpainting/PTransformer.py:28-165, :284-372)
# ----------------------------------------------------------------------------------------------

D_MODEL, N_HEADS, HEAD_DIM = 512, 8, 64


class StyleWeights:
    """One StylizationBlock_V3 ('sb'): LN affine, per-class [scale | shift] table, output Linear."""

    def __init__(self, sd, p, class_embed, device, split=False):
        self.gamma = sd[p + ".norm_2.weight"].float().to(device).contiguous()
        self.beta = sd[p + ".norm_2.bias"].float().to(device).contiguous()
        w, b = sd[p + ".emb_layers_2.1.weight"].float().to(device), sd[p + ".emb_layers_2.1.bias"].float().to(device)
        # emb_layers_2 = Linear(SiLU(class_embed[label])) depends on the label only: tabulate it per class (fp32)
        self.mod = torch.nn.functional.linear(torch.nn.functional.silu(class_embed), w, b).contiguous()   # [n_classes, 1024]
        self.out = prep_linear(sd[p + ".out_layers_2.2.weight"], sd[p + ".out_layers_2.2.bias"], device, split=split)
        # fused epilogue form:  LN(h)*(1+scale)+shift = (h-mean)*rstd * A[c] + B[c]
        scale, shift = self.mod[:, :D_MODEL], self.mod[:, D_MODEL:]
        self.norm_a = (self.gamma[None] * (1 + scale)).contiguous()
        self.norm_b = (self.beta[None] * (1 + scale) + shift).contiguous()


class LayerWeights:
    def __init__(self, sd, p, class_embed, device, split=False):
        C_ = D_MODEL
        lin = lambda w, b: prep_linear(w, b, device, split=split)
        Wi, bi = sd[p + ".self_attn.in_proj_weight"], sd[p + ".self_attn.in_proj_bias"]
        self.qkv = lin(Wi, bi)
        self.sa_out = lin(sd[p + ".self_attn.out_proj.weight"], sd[p + ".self_attn.out_proj.bias"])
        Wc, bc = sd[p + ".multihead_attn.in_proj_weight"], sd[p + ".multihead_attn.in_proj_bias"]
        self.ca_q = lin(Wc[:C_], bc[:C_])
        self.ca_kv_w, self.ca_kv_b = Wc[C_:], bc[C_:]          # stacked across layers by the stack builder
        self.ca_out = lin(sd[p + ".multihead_attn.out_proj.weight"], sd[p + ".multihead_attn.out_proj.bias"])
        self.ff1 = lin(sd[p + ".linear1.weight"], sd[p + ".linear1.bias"])
        self.ff2 = lin(sd[p + ".linear2.weight"], sd[p + ".linear2.bias"])
        self.norms = [(sd[p + ".norm%d.weight" % i].float().to(device).contiguous(),
                       sd[p + ".norm%d.bias" % i].float().to(device).contiguous()) for i in (1, 2, 3)]
        self.styles = [StyleWeights(sd, p + ".style_gourp_%d" % i, class_embed, device, split) for i in (1, 2, 3)]


class StackWeights:
    def __init__(self, sd, prefix, class_embed, device, n_layers=6, split=False):
        self.split = split
        self.layers = [LayerWeights(sd, "%s.layers.%d" % (prefix, l), class_embed, device, split) for l in range(n_layers)]
        # cross-attention K/V projections of all layers as ONE GEMM over the (constant) memory
        w = torch.cat([lw.ca_kv_w for lw in self.layers], 0)
        b = torch.cat([lw.ca_kv_b for lw in self.layers], 0)
        self.kv_all = prep_linear(w, b, device, split=split)


class StackPlanner:
    """Records the launches of a 6-layer DeepSupervisionLayer stack over fixed buffers.

    xb   : the residual stream [B*N, 512] in bf16 (every sub-block ends in a LayerNorm, so rounding does not
           accumulate across layers); r: pre-norm sum x + style(...) in bf16; LayerNorm statistics are fp32
    memb : bf16 memory rows [B*Mem, 512]; labels int32 [B]."""

    def __init__(self, plan: Plan, W: StackWeights, B, N, Mem, labels, device, fuse_style=True, fuse_norm=True):
        self.p, self.W, self.B, self.N, self.Mem = plan, W, B, N, Mem
        # fuse_style: out-proj / FF2 GEMM + style LayerNorm-modulation-SiLU in one kernel (cluster row-norm epilogue)
        # fuse_norm : style GEMM + residual + LayerNorm in one kernel (107 us vs 54 + 59 us unfused at M = 90112)
        # split-bf16 arithmetic (W.split): GEMM operands are (hi, lo) plane pairs, everything the elementwise / attention
        # kernels read is fp32, and the row-norm epilogues (bf16 by construction) give way to the stand-alone kernels
        sp = self.split = W.split
        self.fuse_style, self.fuse_norm = fuse_style and not sp, fuse_norm and not sp
        M = B * N
        self.M = M
        m = lambda rows, cols, **k: Mat(rows, cols, device, **k)
        mid = F32 if sp else BF16                    # what attention / LayerNorm kernels consume
        self.xb = m(M, D_MODEL, split=sp)
        self.memb = m(B * Mem, D_MODEL, split=sp)
        self.kv = m(B * Mem, 2 * D_MODEL * len(W.layers), dtype=mid)
        self.qkv = m(M, 3 * D_MODEL, dtype=mid)
        self.q = m(M, D_MODEL, dtype=mid)
        self.att = m(M, D_MODEL, split=sp)
        self.proj = m(M, D_MODEL, dtype=mid)
        self.sil = m(M, D_MODEL, split=sp)
        self.r = m(M, D_MODEL, dtype=mid)
        self.ff = m(M, 4 * D_MODEL, split=sp)
        self.labels = labels

    def record_memory_kv(self):
        self.p.linear(self.memb, self.B * self.Mem, self.W.kv_all, self.kv)

    def _sublayer_tail(self, lw: LayerWeights, i, a_t, w_proj: Weight):
        """x = norm_i( x + style_i( proj(a) ) )   (PTransformer.py:329-331 with :134-140, :344-364) as two fused GEMMs:
             sil = SiLU( LN(a W_proj^T + b) (1+scale) + shift )         [row-norm epilogue, per-class tables]
             x   = LN_i( sil W_style^T + b + x )                         [row-norm epilogue, residual in place]"""
        st = lw.styles[i]
        p, M = self.p, self.M
        g, b = lw.norms[i]
        L = _lib.lib()
        if self.fuse_style:
            p.linear(a_t, M, w_proj, self.sil, row_norm=1, norm_a=st.norm_a, norm_b=st.norm_b, labels=self.labels,
                     label_rows=self.N)
        else:
            f32 = int(self.split)
            p.linear(a_t, M, w_proj, self.proj)
            p.call(L.pt_ln_mod_silu512, self.proj.t.data_ptr(), f32, st.gamma.data_ptr(), st.beta.data_ptr(),
                   st.mod.data_ptr(), self.labels.data_ptr(), self.N, self.sil.t.data_ptr(), _ptr(self.sil.lo), M)
        if self.fuse_norm:
            p.linear(self.sil, M, st.out, self.xb, residual=self.xb, rss=0, rrs=D_MODEL, row_norm=2, norm_a=g, norm_b=b)
            return
        p.linear(self.sil, M, st.out, self.r)
        # the residual add rides in the LayerNorm pass (a row-local read) instead of the GEMM epilogue
        p.call(L.pt_layernorm512, self.r.t.data_ptr(), int(self.split), self.xb.t.data_ptr(), _ptr(self.xb.lo), g.data_ptr(),
               b.data_ptr(), None, self.xb.t.data_ptr(), _ptr(self.xb.lo), M)

    def record_layers(self):
        L = _lib.lib()
        p, M, B, N, Mem = self.p, self.M, self.B, self.N, self.Mem
        scale = 1.0 / math.sqrt(HEAD_DIM)
        nl = len(self.W.layers)
        es = 4 if self.split else 2                      # bytes per q / k / v element

        def attention(qp, qs, kp, ks, vp, vs, Nk):
            if self.split:
                p.call(L.pt_attention_f32, qp, qs, kp, ks, vp, vs, self.att.t.data_ptr(), self.att.lo.data_ptr(), D_MODEL,
                       B, N_HEADS, HEAD_DIM, N, Nk, scale)
            else:
                p.call(L.pt_attention, qp, qs, kp, ks, vp, vs, self.att.t.data_ptr(), D_MODEL, B, N_HEADS, HEAD_DIM, N, Nk,
                       scale)

        for l, lw in enumerate(self.W.layers):
            # self attention
            p.linear(self.xb, M, lw.qkv, self.qkv)
            qkv = self.qkv.t.data_ptr()
            attention(qkv, 3 * D_MODEL, qkv + es * D_MODEL, 3 * D_MODEL, qkv + es * 2 * D_MODEL, 3 * D_MODEL, N)
            self._sublayer_tail(lw, 0, self.att, lw.sa_out)
            # cross attention over the memory
            p.linear(self.xb, M, lw.ca_q, self.q)
            kv_stride = 2 * D_MODEL * nl
            kptr = self.kv.t.data_ptr() + es * (l * 2 * D_MODEL)
            attention(self.q.t.data_ptr(), D_MODEL, kptr, kv_stride, kptr + es * D_MODEL, kv_stride, Mem)
            self._sublayer_tail(lw, 1, self.att, lw.ca_out)
            # feed forward
            p.linear(self.xb, M, lw.ff1, self.ff, act=ACT_RELU)
            self._sublayer_tail(lw, 2, self.ff, lw.ff2)


# ----------------------------------------------------------------------------------------------
# predictor  (reference: predictor.TrainWrapper.infer_on_batch / infer_to_code,
#             nets/inpainting/predictor.py:398-448, :597-647; PredictorNet.forward predictornet.py:76-112;
#             PTransformer.sample_good / tokens_to_logits, PTransformer.py:590-704)
# ----------------------------------------------------------------------------------------------

MASK_ID = 128
CHOICE_TEMPERATURE = 4.5


class PredictorWeights:
    def __init__(self, state, device, precision="bf16"):
        dev = device
        aud = {k: v.to(dev) for k, v in state["AudEnc"].items() if torch.is_tensor(v)}
        sd = {k: v.to(dev) for k, v in state["Predictor"].items() if torch.is_tensor(v)}
        self.device = dev
        self.precision = precision
        sp = self.precise = _check_precision(precision)
        # AudEnc (ConditionEncoder, vqvae_1d_sc.py:70-91)
        self.aud_map = prep_conv(aud["feature_map.0.weight"], aud["feature_map.0.bias"], dev,
                                 bn=_bn_scale_shift(aud, "feature_map.1"), split=sp)
        self.aud_downs = [prep_conv(aud["_down_%d.conv.weight" % i], aud["_down_%d.conv.bias" % i], dev,
                                    bn=_bn_scale_shift(aud, "_down_%d.norm" % i),
                                    extra=(aud["_down_%d.residual_layer.weight" % i], aud["_down_%d.residual_layer.bias" % i]),
                                    split=sp)
                          for i in (1, 2, 3)]
        # context encoder (predictornet.py:31-48, gated_pixelcnn_1d.py:39-83)
        self.mot_dim = sd["motion_embedding.weight"].shape[1] - 1
        self.embed = prep_conv(sd["motion_embedding.weight"], sd["motion_embedding.bias"], dev, split=sp)
        self.enc = []
        for i in range(10):
            p = "enc.%d" % i
            k = sd[p + ".vert_stack.weight"].shape[2]
            stride = 2 if (p + ".conv.weight") in sd else 1
            item = dict(k=k, stride=stride, pad=3 if k == 7 else 1,
                        cin=sd[p + ".vert_stack.weight"].shape[1], dim=sd[p + ".res.weight"].shape[0],
                        vert=prep_conv(sd[p + ".vert_stack.weight"], sd[p + ".vert_stack.bias"], dev, gate=True, split=sp),
                        res=prep_linear(sd[p + ".res.weight"], sd[p + ".res.bias"], dev, split=sp))
            if stride == 2:
                item["skip"] = prep_conv(sd[p + ".conv.weight"], sd[p + ".conv.bias"], dev, split=sp)
            self.enc.append(item)
        # PTransformer
        d = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
        self.tok_emb = d["tok_emb.weight"].float().contiguous()
        self.n_emb = self.tok_emb.shape[0]
        self.class_embed = d["class_embed.weight"].float()
        self.PE = d["PE.position_encoding.weight"].double()
        self.PPE = d["product_PE.position_encoding.weight"].double()
        self.groups = self.PPE.shape[0] - 1
        self.condi = prep_linear(d["condi_emb.weight"], d["condi_emb.bias"], dev, split=sp)
        self.stack = StackWeights(d, "transformer_decoder", self.class_embed, dev, split=sp)
        self.head0 = prep_linear(d["output_conv.0.weight"], d["output_conv.0.bias"], dev, split=sp)
        self.head1 = prep_linear(d["output_conv.2.weight"], d["output_conv.2.bias"], dev, split=sp)
        self.n_codes = d["output_conv.2.weight"].shape[0]


def maskgit_base_len(N, step, T):
    """floor(N * cos(pi/2 * (step+1)/T)) evaluated like the reference: int64 tensor * numpy float64 -> float32
    (PTransformer.py:610-621)."""
    ratio = 1.0 * (step + 1) / T
    mask_ratio = np.cos(ratio * np.pi / 2)
    return int(torch.floor(torch.tensor([N], dtype=torch.int64) * mask_ratio).item())


class PredictorPlan:
    """One (B, T) bucket of s2g_body_predictor inference on already-batched inputs:
         aud [B,768,T], gt_poses [B,376,T], mask [B,1,T], labels [B]  ->  poses [B,376,T], codes [B,T',G]."""

    def __init__(self, PW: PredictorWeights, VW: VQVAEWeights, B, T, maskgit_T=8, use_graph=True):
        dev = PW.device
        self.PW, self.VW, self.B, self.T, self.steps_T = PW, VW, B, T, maskgit_T
        L = _lib.lib()
        G = PW.groups
        f32 = lambda *s: torch.zeros(*s, dtype=F32, device=dev)
        self.aud, self.gt, self.mask = f32(B, 768, T), f32(B, PW.mot_dim, T), f32(B, 1, T)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        # ---------------- plan A: condition (AudEnc + context encoder + memory K/V) ----------------
        pa = self.plan_cond = Plan(dev)
        sp = PW.precise
        RB = lambda *a, **k: RowBuf(*a, device=dev, split=sp, **k)
        xa = RB(B, T, 768, 0)
        pa.call(L.pt_pack_rows, self.aud.data_ptr(), None, None, 0, xa.t.data_ptr(), _ptr(xa.lo), B, 768, T, xa.Tp, xa.H,
                xa.C, 0, 0)
        h = RB(B, T, 256, 1)
        pa.conv(xa, PW.aud_map, h, k=1, act=ACT_LEAKY, slope=0.1)
        Tq = T
        for i in range(3):
            Tq = (Tq - 2) // 2 + 1
        self.Tq = Tq
        self.N = Tq * G
        cond = RB(B, Tq, 768, 0)                                   # [mm (512) | audio (256)]  (predictornet.py:93-95)
        for i in range(3):
            To = (h.T - 2) // 2 + 1
            if i < 2:
                d = RB(B, To, 256, 1)
                pa.conv(h, PW.aud_downs[i], d, k=4, stride=2, pad=1, act=ACT_LEAKY)
                h = d
            else:
                pa.conv(h, PW.aud_downs[i], cond, k=4, stride=2, pad=1, act=ACT_LEAKY, out_c_off=512)
        xm = RB(B, T, _ceil(PW.mot_dim + 1, 64), 0)
        pa.call(L.pt_pack_rows, self.gt.data_ptr(), self.mask.data_ptr(), None, 0, xm.t.data_ptr(), _ptr(xm.lo), B,
                PW.mot_dim, T, xm.Tp, xm.H, xm.C, 0, 1)
        # (a 0 / 1 mask is exact in bf16: the lo plane of the mask channel stays zero)
        pa.call(L.pt_pack_mask, self.mask.data_ptr(), xm.t.data_ptr(), B, T, xm.Tp, xm.H, xm.C, PW.mot_dim, 0)
        mm = RB(B, T, 128, 3)
        pa.conv(xm, PW.embed, mm, k=1)
        self.ctx_stages = [("motion_embedding", mm)]
        for i, e in enumerate(PW.enc):
            To = mm.T if e["stride"] == 1 else (mm.T - 2) // 2 + 1
            gate = RB(B, To, e["dim"], 0)
            pa.conv(mm, e["vert"], gate, k=e["k"], stride=e["stride"], pad=e["pad"], act=ACT_GATE)
            if e["stride"] == 2:
                skip = RB(B, To, e["dim"], 0)
                pa.conv(mm, e["skip"], skip, k=4, stride=2, pad=1)
            else:
                skip = mm
            last = i == len(PW.enc) - 1
            if last:
                pa.conv(gate, e["res"], cond, k=1, residual=skip, out_c_off=0)
                nxt = cond
            else:
                nxt = RB(B, To, e["dim"], 1)
                pa.conv(gate, e["res"], nxt, k=1, residual=skip)
            mm = nxt
            self.ctx_stages.append(("enc.%d" % i, mm))
        assert mm.T == Tq
        self.cond = cond
        # positional tables (PTransformer.py:669-674): token (t, g) -> PE[t+1] + product_PE[g+1]; memory t -> PE[t+1]
        c_pe = PW.PE[1:Tq + 1]
        x_pe = c_pe.repeat_interleave(G, 0) + PW.PPE[1:G + 1].repeat(Tq, 1)
        self.x_pe64 = x_pe.contiguous().to(dev)
        self.c_pe32 = c_pe.float().contiguous().to(dev)
        self.stack = StackPlanner(pa, PW.stack, B, self.N, Tq, self.labels, dev)
        # condition embedding 1x1 (768 -> 512) + PE, rounded once to bf16 (only ever a GEMM operand)
        pa.gemm(cond.t, B * cond.Tp, 768, PW.condi, self.stack.memb.t, M=B * cond.Tp, C_=768, base_row=0,
                rows_per_sample=cond.Tp, t_valid=Tq, oss=Tq * D_MODEL, ors=D_MODEL, residual=self.c_pe32, rss=0,
                rrs=D_MODEL, a_lo=cond.lo, out_lo=self.stack.memb.lo)
        self.stack.record_memory_kv()
        # ---------------- plan B: one MaskGIT iteration (embed -> 6 layers -> logits) ----------------
        pb = self.plan_iter = Plan(dev)
        self.stack.p = pb
        M = B * self.N
        self.cur = torch.full((B, self.N), MASK_ID, dtype=torch.int32, device=dev)
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)
        pb.call(L.pt_embed_tokens512, self.cur.data_ptr(), PW.tok_emb.data_ptr(), PW.n_emb, self.x_pe64.data_ptr(), self.N,
                None, self.stack.xb.t.data_ptr(), _ptr(self.stack.xb.lo), M, self.bad.data_ptr())
        self.stack.record_layers()
        self.h1 = Mat(M, D_MODEL, dev, split=sp)
        self.logits = torch.zeros(M, PW.n_codes, dtype=F32, device=dev)
        pb.linear(self.stack.xb, M, PW.head0, self.h1, act=ACT_RELU)
        pb.linear(self.h1, M, PW.head1, self.logits)
        # ---------------- sampler steps (one tiny plan per MaskGIT iteration) ----------------
        self.u = f32(maskgit_T, B, self.N)
        self.gumbel = f32(maskgit_T, B, self.N)
        self.base_len = [maskgit_base_len(self.N, t, maskgit_T) for t in range(maskgit_T)]
        self.plan_sample = []
        for t in range(maskgit_T):
            ps = Plan(dev)
            ratio = 1.0 * (t + 1) / maskgit_T
            ps.call(L.pt_maskgit_step, self.logits.data_ptr(), self.cur.data_ptr(), self.u[t].data_ptr(),
                    self.gumbel[t].data_ptr(), self.cur.data_ptr(), B, self.N, PW.n_codes, MASK_ID, self.base_len[t],
                    CHOICE_TEMPERATURE * (1.0 - ratio))
            self.plan_sample.append(ps)
        # ---------------- plan C: decode (the token buffer IS the [B, T', G] int32 code tensor) ----------------
        self.codes = self.cur.view(B, Tq, G)
        self.dec = DecoderPlan(VW, B, Tq, codes=self.codes)
        self.out = f32(B, PW.mot_dim, T)
        self.dec.plan.call(L.pt_interp_blend, self.gt.data_ptr(), self.mask.data_ptr(), self.dec.out.data_ptr(), self.dec.T_out,
                           self.out.data_ptr(), B, PW.mot_dim, T, 2)
        self.launches = pa.launches + maskgit_T * (pb.launches + 1) + self.dec.plan.launches
        self.use_graph, self.graph, self._warm = use_graph, None, False

    def plans_in_order(self):
        seq = [self.plan_cond]
        for t in range(self.steps_T):
            seq += [self.plan_iter, self.plan_sample[t]]
        return seq + [self.dec.plan]

    def draw_noise(self, generator=None):
        """Sampler noise on the device: u ~ U(0,1) for the categorical draw, Gumbel(0,1) for the confidences."""
        self.u.uniform_(generator=generator)
        g = torch.empty_like(self.gumbel).uniform_(generator=generator).clamp_(1e-10, 1 - 1e-7)
        self.gumbel.copy_(-torch.log(-torch.log(g)))

    def run_condition(self):
        self.plan_cond.run()

    def run_iteration(self, t):
        """tokens_to_logits + one sample_good step (PTransformer.py:603-638) using noise slice t."""
        self.plan_iter.run()
        self.plan_sample[t].run()

    def run_decode(self):
        self.dec.plan.run()
        return self.out

    def run(self, aud, gt, mask, labels, noise=None, generator=None, force_codes=None):
        """Full infer_to_code for a batch.  noise = (u, gumbel) [T, B, N] to inject, else drawn on the device.
        force_codes [B, T', G]: teacher forcing -- skip the sampler and decode these codes (parity tests).
        Returns (poses [B,376,T] fp32, codes [B,T',G] int32) -- views of plan-owned buffers."""
        self.aud.copy_(aud); self.gt.copy_(gt); self.mask.copy_(mask); self.labels.copy_(labels)
        if force_codes is not None:
            self.cur.copy_(force_codes.reshape(self.B, -1))
            self.dec.plan.run()
            return self.out, self.codes
        if noise is not None:
            self.u.copy_(noise[0]); self.gumbel.copy_(noise[1])
        else:
            self.draw_noise(generator)
        self._replay()
        return self.out, self.codes

    def _launch_all(self):
        self.cur.fill_(MASK_ID)
        for pl in self.plans_in_order():
            pl.run()

    def _replay(self):
        """First call launches eagerly (also the warm-up that sets kernel attributes); from the second call on the
        whole window (~750 launches) is replayed as ONE CUDA graph, which removes the host launch cost that
        dominates small batches."""
        if not self.use_graph:
            return self._launch_all()
        if self.graph is None:
            if not self._warm:
                self._warm = True
                return self._launch_all()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_all()
            self.graph = g
        self.graph.replay()


# ----------------------------------------------------------------------------------------------
# refiner  (reference: refiner.TrainWrapper.infer_on_batch, nets/inpainting/refiner.py:512-558;
#           RTransformer.forward, nets/inpainting/RTransformer.py:209-230)
# ----------------------------------------------------------------------------------------------

class RefinerWeights:
    def __init__(self, state, device, precision="bf16"):
        dev = device
        sd = {k: v.to(dev) for k, v in state["HFTransformer"].items() if torch.is_tensor(v)}
        self.device = dev
        self.precision = precision
        sp = self.precise = _check_precision(precision)
        self.in_dim = sd["input_conv.weight"].shape[1]
        self.aud_dim = sd["audio_conv.weight"].shape[1] - 2
        self.class_embed = sd["class_embed.weight"].float()
        self.PE = sd["PE.position_encoding.weight"].double()
        self.input_conv = prep_linear(sd["input_conv.weight"], sd["input_conv.bias"], dev, split=sp)
        self.audio_conv = prep_linear(sd["audio_conv.weight"], sd["audio_conv.bias"], dev, split=sp)
        self.stack = StackWeights(sd, "transformer_decoder", self.class_embed, dev, split=sp)
        self.head0 = prep_linear(sd["output_conv.0.weight"], sd["output_conv.0.bias"], dev, split=sp)
        self.head1 = prep_linear(sd["output_conv.2.weight"], sd["output_conv.2.bias"], dev, split=sp)


class RefinerPlan:
    """One (B, T) bucket: aud [B,768,T], gt [B,376,T], mask [B,1,T], pred [B,376,Tpred], labels [B] -> [B,376,T]."""

    def __init__(self, RW: RefinerWeights, B, T, T_pred=None, use_graph=True):
        dev = RW.device
        self.RW, self.B, self.T = RW, B, T
        self.T_pred = T_pred or T
        L = _lib.lib()
        f32 = lambda *s: torch.zeros(*s, dtype=F32, device=dev)
        self.aud, self.gt, self.mask = f32(B, RW.aud_dim, T), f32(B, RW.in_dim, T), f32(B, 1, T)
        self.pred = f32(B, RW.in_dim, self.T_pred)
        self.labels = torch.zeros(B, dtype=torch.int32, device=dev)
        p = self.plan = Plan(dev)
        M = B * T
        # blank = gt*mask + interp(pred)*(1-mask)  -> rows [B*T, 384]
        sp = RW.precise
        xin = RowBuf(B, T, _ceil(RW.in_dim, 64), 0, device=dev, even=False, split=sp)      # any T (odd tail windows too)
        p.call(L.pt_pack_rows, self.gt.data_ptr(), self.mask.data_ptr(), self.pred.data_ptr(), self.T_pred, xin.t.data_ptr(),
               _ptr(xin.lo), B, RW.in_dim, T, xin.Tp, 0, xin.C, 0, 2)
        # condition = cat(audio, one_hot(mask))  -> rows [B*T, 832]
        xc = RowBuf(B, T, _ceil(RW.aud_dim + 2, 64), 0, device=dev, even=False, split=sp)
        p.call(L.pt_pack_rows, self.aud.data_ptr(), None, None, 0, xc.t.data_ptr(), _ptr(xc.lo), B, RW.aud_dim, T, xc.Tp, 0,
               xc.C, 0, 0)
        p.call(L.pt_pack_mask, self.mask.data_ptr(), xc.t.data_ptr(), B, T, xc.Tp, 0, xc.C, RW.aud_dim, 1)
        self.pe32 = RW.PE[1:T + 1].float().contiguous().to(dev)
        self.stack = StackPlanner(p, RW.stack, B, T, T, self.labels, dev)
        st = self.stack
        # x = float32(input_conv(blank) + PE)   (RTransformer.py:211-214)
        p.linear(xin, M, RW.input_conv, st.xb, residual=self.pe32, rss=0, rrs=D_MODEL, rows_per_sample=T)
        p.linear(xc, M, RW.audio_conv, st.memb, residual=self.pe32, rss=0, rrs=D_MODEL, rows_per_sample=T)
        st.record_memory_kv()
        st.record_layers()
        self.h1 = Mat(M, 1024, dev, split=sp)
        self.raw = f32(B, RW.in_dim, T)
        p.linear(st.xb, M, RW.head0, self.h1, act=ACT_RELU)
        w = RW.head1
        p.gemm(self.h1.t, M, 1024, w, self.raw, M=M, C_=1024, rows_per_sample=T, t_valid=T, oss=RW.in_dim * T, ors=1, ocs=T,
               a_lo=self.h1.lo)
        self.out = f32(B, RW.in_dim, T)
        p.call(L.pt_interp_blend, self.gt.data_ptr(), self.mask.data_ptr(), self.raw.data_ptr(), T, self.out.data_ptr(), B,
               RW.in_dim, T, 1)
        self.launches = p.launches
        self.use_graph, self.graph, self._warm = use_graph, None, False
        p.keep += [xin, xc]

    def run(self, aud, gt, mask, pred, labels):
        self.aud.copy_(aud); self.gt.copy_(gt); self.mask.copy_(mask); self.pred.copy_(pred); self.labels.copy_(labels)
        if not self.use_graph:
            self.plan.run()
        elif self.graph is None and not self._warm:
            self._warm = True
            self.plan.run()
        else:
            if self.graph is None:
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self.plan.run()
                self.graph = g
            self.graph.replay()
        return self.out


# ----------------------------------------------------------------------------------------------
# FGD feature extractor  (reference: EmbeddingNet.extract -> PoseEncoderConv.forward, nets/embedding_net.py:56-95,
#                         147-150; ConvNormRelu :29-53)  -- SURVEY.md 8(f)3
# ----------------------------------------------------------------------------------------------

class EmbeddingWeights:
    """Folded weights of PoseEncoderConv(length=90, dim): four VALID convolutions (k3, k3, k4 s2, k3; BatchNorm folded,
    LeakyReLU 0.2) and the MLP 1280 -> 512 -> 256 -> 128 -> fc_mu 128.  `nn.LeakyReLU(True)` in the MLP is
    negative_slope = 1.0, i.e. the identity (embedding_net.py:71-77), reproduced as such."""

    def __init__(self, sd, device, prefix="pose_encoder.", precision="bf16"):
        g = lambda k: sd[prefix + k]
        self.dim = g("net.0.0.weight").shape[1]
        self.precision = precision
        sp = self.precise = _check_precision(precision)
        self.c1 = prep_conv(g("net.0.0.weight"), g("net.0.0.bias"), device, bn=_bn_scale_shift(sd, prefix + "net.0.1"), split=sp)
        self.c2 = prep_conv(g("net.1.0.weight"), g("net.1.0.bias"), device, bn=_bn_scale_shift(sd, prefix + "net.1.1"), cin_pad=64, split=sp)
        self.c3 = prep_conv(g("net.2.0.weight"), g("net.2.0.bias"), device, bn=_bn_scale_shift(sd, prefix + "net.2.1"), split=sp)
        self.c4 = prep_conv(g("net.3.weight"), g("net.3.bias"), device, split=sp)
        # flatten(1) of [B, 32, 40] is channel-major; our rows are [B, 40, 64]: W'[n, t*64 + c] = W[n, c*40 + t]
        w = g("out_net.0.weight").float()                                   # [512, 1280]
        s0, h0 = _bn_scale_shift(sd, prefix + "out_net.1")
        wp = torch.zeros(512, 40, 64)
        wp[:, :, :32] = w.view(512, 32, 40).permute(0, 2, 1)
        self.l1 = prep_linear(wp.reshape(512, 2560) * s0[:, None], g("out_net.0.bias").float() * s0 + h0, device, split=sp)
        s1, h1 = _bn_scale_shift(sd, prefix + "out_net.4")
        self.l2 = prep_linear(g("out_net.3.weight").float() * s1[:, None], g("out_net.3.bias").float() * s1 + h1, device, split=sp)
        self.l3 = prep_linear(g("out_net.6.weight"), g("out_net.6.bias"), device, split=sp)
        self.mu = prep_linear(g("fc_mu.weight"), g("fc_mu.bias"), device, split=sp)


class EmbeddingPlan:
    """poses [B, dim, 90] fp32 (channel-first, as pushed by test_holisticbody.py:279) -> features [B, 128] fp32."""

    def __init__(self, W: EmbeddingWeights, B, dev="cuda"):
        self.W, self.B = W, B
        p = self.plan = Plan(dev)
        L = _lib.lib()
        T = 90
        self.poses = torch.zeros(B, W.dim, T, dtype=F32, device=dev)
        sp = W.precise
        RB = lambda *a: RowBuf(*a, device=dev, split=sp)
        x0 = RB(B, T, _ceil(W.dim, 64), 0)
        p.call(L.pt_pack_rows, self.poses.data_ptr(), None, None, 0, x0.t.data_ptr(), _ptr(x0.lo), B, W.dim, T, x0.Tp, 0,
               x0.C, 0, 0)
        h1, h2 = RB(B, 88, 64, 0), RB(B, 86, 64, 0)
        h3, h4 = RB(B, 42, 64, 0), RB(B, 40, 64, 0)

        def valid_conv(x, w, out, act):                  # k3, stride 1, no padding: out[t] = sum_j x[t + j]
            p.gemm(x.t, x.B * x.Tp, x.C, w, out.t, M=x.B * x.Tp, C_=x.C, s=1, base_row=0, rows_per_sample=x.Tp,
                   t_valid=out.T, oss=out.Tp * out.C, ors=out.C, act=act, slope=0.2, a_lo=x.lo, out_lo=out.lo)

        valid_conv(x0, W.c1, h1, ACT_LEAKY)
        valid_conv(h1, W.c2, h2, ACT_LEAKY)
        # k4, stride 2, no padding: rows viewed as [Tp / 2, 2 C]; tap j = row t + j / 2, column half j % 2
        p.gemm(h2.t, B * h2.Tp, h2.C, W.c3, h3.t, M=B * h2.Tp // 2, C_=h2.C, s=2, base_row=0, rows_per_sample=h2.Tp // 2,
               t_valid=h3.T, oss=h3.Tp * h3.C, ors=h3.C, act=ACT_LEAKY, slope=0.2, a_lo=h2.lo, out_lo=h3.lo)
        valid_conv(h3, W.c4, h4, ACT_NONE)
        f1, f2, f3 = Mat(B, 512, dev, split=sp), Mat(B, 256, dev, split=sp), Mat(B, 128, dev, split=sp)
        self.out = torch.zeros(B, 128, dtype=F32, device=dev)
        p.linear(h4, B, W.l1, f1)                        # A = the [B, 40 * 64] view of the conv output
        p.linear(f1, B, W.l2, f2)
        p.linear(f2, B, W.l3, f3)
        p.linear(f3, B, W.mu, self.out)
        self.keep = [x0, h1, h2, h3, h4, f1, f2, f3]

    def extract(self, poses):
        """poses [n <= B, dim, 90] -> [n, 128] (a view of the plan's output buffer; clone to keep it)."""
        n = poses.shape[0]
        assert n <= self.B and tuple(poses.shape[1:]) == (self.W.dim, 90)
        self.poses[:n].copy_(poses)
        self.plan.run()
        return self.out[:n]
