"""Thin torch-tensor wrappers over the C ABI (``include/probtalk_b200.h``).

torch is used for device memory and streams only; all arithmetic happens in the CUDA library.
Every wrapper requires CUDA tensors and raises if the library is missing (no CPU fallback).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import GemmDesc, check

ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_GATE, ACT_GELU = 0, 1, 2, 3, 4


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.PtError("probtalk_b200 ops need CUDA tensors (no CPU fallback)")


def pq_search(z: torch.Tensor, codebooks: torch.Tensor, want_gap: bool = False, score_dump: bool = False,
              count_rescored: bool = False):
    """Nearest-code search.  z [B, G*D, W] fp32 (any strides), codebooks [G, K, D] fp32.
    Returns idx int64 [B, W, G] (+ gap fp32 [B, W, G], + scores [B, W, G, K], + n_rescored tensor).
    Mirrors ProductQuantization.get_code_indices (reference nets/spg/vqvae_modules.py:546-555)."""
    _need_cuda(z, codebooks)
    assert z.dtype == torch.float32 and codebooks.dtype == torch.float32 and z.dim() == 3
    G, K, D = codebooks.shape
    B, Cc, W = z.shape
    assert Cc == G * D, "z has %d channels, codebooks cover %d" % (Cc, G * D)
    codebooks = codebooks.contiguous()
    idx = torch.empty((B, W, G), dtype=torch.int64, device=z.device)
    gap = torch.empty((B, W, G), dtype=torch.float32, device=z.device) if want_gap else None
    dump = torch.empty((B, W, G, K), dtype=torch.float32, device=z.device) if score_dump else None
    nres = torch.zeros((1,), dtype=torch.int32, device=z.device) if count_rescored else None
    sb, sc, sw = z.stride()
    check(_lib.lib().pt_pq_search(z.data_ptr(), sb, sc, sw, B, W, G, K, D, codebooks.data_ptr(), idx.data_ptr(),
                                  _ptr(gap), _ptr(dump), _ptr(nres), _stream()), "pt_pq_search")
    out = [idx]
    if want_gap:
        out.append(gap)
    if score_dump:
        out.append(dump)
    if count_rescored:
        out.append(nres)
    return out[0] if len(out) == 1 else tuple(out)


def pq_gather(idx: torch.Tensor, codebooks: torch.Tensor, out: torch.Tensor = None, channels_last: bool = False,
              dtype=torch.float32):
    """Codebook gather: idx [B, W, G] -> quantized [B, G*D, W] (reference layout, default) or
    [B, W, G*D] (channels_last).  Mirrors ProductQuantization.quantize_all (vqvae_modules.py:557-564)
    followed by the transpose of predictor.py:440."""
    _need_cuda(idx, codebooks)
    G, K, D = codebooks.shape
    B, W, G2 = idx.shape
    assert G2 == G and idx.dtype in (torch.int64, torch.int32)
    idx = idx.contiguous()
    codebooks = codebooks.contiguous()
    if out is None:
        shape = (B, W, G * D) if channels_last else (B, G * D, W)
        out = torch.empty(shape, dtype=dtype, device=idx.device)
    if channels_last:
        os_b, os_w, os_c = out.stride()
    else:
        os_b, os_c, os_w = out.stride()
    bad = torch.zeros((1,), dtype=torch.int32, device=idx.device)
    check(_lib.lib().pt_pq_gather(idx.data_ptr(), int(idx.dtype == torch.int64), codebooks.data_ptr(), out.data_ptr(),
                                  int(out.dtype == torch.bfloat16), os_b, os_c, os_w, B, W, G, K, D, bad.data_ptr(),
                                  _stream()), "pt_pq_gather")
    return out, bad


def gemm(a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, *, M: int, C_: int, taps: int = 1, s: int = 1,
         base_row: int = 0, rows_per_sample: int = None, t_valid: int = None, n_valid: int = None,
         oss: int = 0, ors: int = None, ocs: int = 1, out_offset: int = 0, bias: torch.Tensor = None,
         residual: torch.Tensor = None, rss: int = 0, rrs: int = 0, res_offset: int = 0, act: int = ACT_NONE,
         slope: float = 0.2, a_rows: int = None, a_pitch: int = None, a_offset: int = 0, row_norm: int = 0,
         norm_a: torch.Tensor = None, norm_b: torch.Tensor = None, labels: torch.Tensor = None, label_rows: int = 1,
         a_lo: torch.Tensor = None, out_lo: torch.Tensor = None, residual_lo: torch.Tensor = None, a_col_table=None):
    """Launch the tensor-core contraction engine (see pt_gemm_bf16 in include/probtalk_b200.h).
    ``a`` / ``out`` / ``residual`` are flat device buffers; offsets are in elements.
    Split-bf16 arithmetic: pass ``a_lo`` (and ``out_lo`` / ``residual_lo`` for bf16 outputs / residuals) and a weight
    of shape [N, 2 * taps * C] = [hi | lo] (``split_weight``)."""
    _need_cuda(a, w, out, bias, residual, a_lo, out_lo, residual_lo)
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and w.dim() == 2
    N = w.shape[0]
    split = a_lo is not None
    assert w.shape[1] == (2 if split else 1) * taps * C_ and w.is_contiguous()
    d = GemmDesc()
    tab = None
    if a_col_table is not None:                 # per-128-column-tile A column offsets (host ints), see pt_gemm_desc
        tab = (C.c_int * len(a_col_table))(*a_col_table)
        d.a_col_table = C.cast(tab, C.c_void_p)
    if split:
        d.split = 1
        d.a_lo = a_lo.data_ptr() + 2 * a_offset
        d.out_lo = out_lo.data_ptr() + 2 * out_offset if out_lo is not None else None
        d.residual_lo = residual_lo.data_ptr() + 2 * res_offset if residual_lo is not None else None
    d.a = a.data_ptr() + 2 * a_offset
    d.a_pitch = a_pitch if a_pitch is not None else C_
    d.a_rows = a_rows if a_rows is not None else (a.numel() - a_offset) // d.a_pitch
    d.w = w.data_ptr()
    d.M, d.N, d.C, d.taps, d.s = M, N, C_, taps, s
    d.base_row = base_row
    d.rows_per_sample = rows_per_sample if rows_per_sample is not None else max(M, 1)
    d.t_valid = t_valid if t_valid is not None else d.rows_per_sample
    d.n_valid = n_valid if n_valid is not None else N
    d.out = out.data_ptr() + out.element_size() * out_offset
    d.oss, d.ors, d.ocs = oss, (ors if ors is not None else d.n_valid), ocs
    d.out_f32 = int(out.dtype == torch.float32)
    assert out.dtype in (torch.float32, torch.bfloat16)
    d.bias = _ptr(bias)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() >= N
    if residual is not None:
        assert residual.dtype in (torch.float32, torch.bfloat16)
        d.residual = residual.data_ptr() + residual.element_size() * res_offset
        d.res_f32 = int(residual.dtype == torch.float32)
    else:
        d.residual = None
        d.res_f32 = 0
    d.rss, d.rrs = rss, rrs
    d.act, d.slope = act, slope
    d.row_norm, d.norm_a, d.norm_b, d.labels, d.label_rows = row_norm, _ptr(norm_a), _ptr(norm_b), _ptr(labels), label_rows
    d.norm_classes = (norm_a.numel() // 512) if norm_a is not None else 0
    check(_lib.lib().pt_gemm_bf16(C.byref(d), _stream()), "pt_gemm_bf16")
    return out


def split_bf16(x: torch.Tensor):
    """fp32 tensor -> (hi, lo) bf16 pair with hi = bf16(x), lo = bf16(x - hi): the operand format of the split-bf16
    ("bf16x3") arithmetic mode (hi + lo carries 16 mantissa bits of x)."""
    hi = x.to(torch.bfloat16)
    return hi, (x.float() - hi.float()).to(torch.bfloat16)


def split_weight(w: torch.Tensor):
    """fp32 weight matrix [N, K] -> bf16 [N, 2K] = [hi | lo], the layout pt_gemm_desc.split expects."""
    hi, lo = split_bf16(w)
    return torch.cat([hi, lo], dim=1).contiguous()


def pq_gather_conv1x1(idx: torch.Tensor, table: torch.Tensor, bias: torch.Tensor, split: bool = False):
    """Fused codebook gather + decoder entry conv (pt_pq_gather_conv1x1): idx [B, W, G], table fp32 [G, K, N], bias [N]
    -> bf16 [B, W, N] (plus the lo plane when ``split``) and the out-of-range index counter."""
    _need_cuda(idx, table, bias)
    B, W, G = idx.shape
    G2, K, N = table.shape
    assert G2 == G and idx.dtype in (torch.int64, torch.int32) and table.dtype == torch.float32
    idx, table, bias = idx.contiguous(), table.contiguous(), bias.contiguous()
    out = torch.empty(B, W, N, dtype=torch.bfloat16, device=idx.device)
    lo = torch.empty_like(out) if split else None
    bad = torch.zeros((1,), dtype=torch.int32, device=idx.device)
    check(_lib.lib().pt_pq_gather_conv1x1(idx.data_ptr(), int(idx.dtype == torch.int64), table.data_ptr(), bias.data_ptr(),
                                          out.data_ptr(), _ptr(lo), W * N, N, B, W, G, K, N, bad.data_ptr(), _stream()),
          "pt_pq_gather_conv1x1")
    return out, lo, bad


def attention_f32(q, k, v, B, Nq, Nk, heads=8, head_dim=64, scale=None, split=True):
    """fp32 attention core (pt_attention_f32): q [B*Nq, heads*head_dim], k / v [B*Nk, ...] fp32 (any row stride)
    -> (hi, lo) bf16 [B*Nq, heads*head_dim]."""
    _need_cuda(q, k, v)
    assert q.dtype == k.dtype == v.dtype == torch.float32
    out = torch.empty(B * Nq, heads * head_dim, dtype=torch.bfloat16, device=q.device)
    lo = torch.empty_like(out) if split else None
    scale = scale if scale is not None else head_dim ** -0.5
    check(_lib.lib().pt_attention_f32(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
                                      out.data_ptr(), _ptr(lo), heads * head_dim, B, heads, head_dim, Nq, Nk, scale,
                                      _stream()), "pt_attention_f32")
    return out, lo


def pq_ema_update(z, idx, codebooks, ema_dw, ema_cluster_size, decay, epsilon, step):
    """EMA codebook update + dead-code reset of ProductQuantization's training forward (pt_pq_ema_update;
    vqvae_modules.py:495-534).  z [B, G*D, W] fp32 (any strides), idx [B, W, G] int64; codebooks [G, K, D], ema_dw [G, K, D]
    and ema_cluster_size [G, K] are updated IN PLACE.  Returns (mse fp64 [1], re-seeded codes per group int32 [G])."""
    _need_cuda(z, idx, codebooks, ema_dw, ema_cluster_size)
    G, K, D = codebooks.shape
    B, Cc, W = z.shape
    assert Cc == G * D and idx.shape == (B, W, G) and idx.dtype == torch.int64 and idx.is_contiguous()
    for t in (codebooks, ema_dw, ema_cluster_size):
        assert t.dtype == torch.float32 and t.is_contiguous()
    L = _lib.lib()
    nbytes = L.pt_pq_ema_workspace_bytes(G, K, D)
    ws = torch.empty((nbytes + 7) // 8, dtype=torch.float64, device=z.device)
    mse = torch.zeros(1, dtype=torch.float64, device=z.device)
    dead = torch.zeros(G, dtype=torch.int32, device=z.device)
    sb, sc, sw = z.stride()
    check(L.pt_pq_ema_update(z.data_ptr(), sb, sc, sw, B, W, G, K, D, idx.data_ptr(), codebooks.data_ptr(), ema_dw.data_ptr(),
                             ema_cluster_size.data_ptr(), float(decay), float(epsilon), int(step), mse.data_ptr(),
                             dead.data_ptr(), ws.data_ptr(), nbytes, _stream()), "pt_pq_ema_update")
    return mse, dead


def attention(q, k, v, B, Nq, Nk, heads=8, head_dim=64, scale=None):
    """bf16 attention core (pt_attention): q [B*Nq, heads*head_dim], k / v [B*Nk, ...] bf16 (any row stride that is a
    multiple of 8) -> bf16 [B*Nq, heads*head_dim].  Nk <= 256: tcgen05; beyond what fits on chip the keys are streamed."""
    _need_cuda(q, k, v)
    assert q.dtype == k.dtype == v.dtype == torch.bfloat16
    out = torch.empty(B * Nq, heads * head_dim, dtype=torch.bfloat16, device=q.device)
    scale = scale if scale is not None else head_dim ** -0.5
    check(_lib.lib().pt_attention(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
                                  out.data_ptr(), heads * head_dim, B, heads, head_dim, Nq, Nk, scale, _stream()),
          "pt_attention")
    return out


def layernorm_rows(x, residual, gamma, beta, eps=1e-5, out=None):
    """LayerNorm over bf16 rows [M, C] (+ bf16 residual), C a multiple of 128 up to 1024 (pt_layernorm_rows)."""
    _need_cuda(x, residual, gamma, beta)
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.is_contiguous()
    assert gamma.dtype == beta.dtype == torch.float32
    M, Cc = x.shape
    out = torch.empty_like(x) if out is None else out
    check(_lib.lib().pt_layernorm_rows(x.data_ptr(), _ptr(residual), gamma.data_ptr(), beta.data_ptr(), float(eps),
                                       out.data_ptr(), M, Cc, _stream()), "pt_layernorm_rows")
    return out


def w2v_frontend(wave, conv0_w, gn_gamma, gn_beta):
    """wav2vec2 processor normalisation + conv0 + GroupNorm + GELU (pt_w2v_frontend): wave [B, L] fp32 ->
    bf16 rows [B, T0, 512], T0 = (L - 10) // 5 + 1."""
    _need_cuda(wave, conv0_w, gn_gamma, gn_beta)
    assert wave.dtype == torch.float32 and wave.dim() == 2 and wave.is_contiguous()
    B, L = wave.shape
    T0 = (L - 10) // 5 + 1
    lib = _lib.lib()
    nbytes = lib.pt_w2v_frontend_workspace_bytes(B)
    ws = torch.empty((nbytes + 7) // 8, dtype=torch.float64, device=wave.device)
    out = torch.empty(B, max(T0, 0), 512, dtype=torch.bfloat16, device=wave.device)
    check(lib.pt_w2v_frontend(wave.data_ptr(), B, L, conv0_w.data_ptr(), gn_gamma.data_ptr(), gn_beta.data_ptr(),
                              out.data_ptr(), max(T0, 0), ws.data_ptr(), nbytes, _stream()), "pt_w2v_frontend")
    return out


def w2v_interp(rows, t_out):
    """rows [B, T, C] (bf16 / fp32) -> F.interpolate(rows.transpose(1, 2), size=t_out, mode='linear') as fp32 [B, C, t_out]."""
    _need_cuda(rows)
    assert rows.dim() == 3 and rows.is_contiguous() and rows.dtype in (torch.bfloat16, torch.float32)
    B, T, Cc = rows.shape
    out = torch.empty(B, Cc, t_out, dtype=torch.float32, device=rows.device)
    check(_lib.lib().pt_w2v_interp(rows.data_ptr(), int(rows.dtype == torch.float32), T, B, T, Cc, out.data_ptr(), t_out,
                                   _stream()), "pt_w2v_interp")
    return out
