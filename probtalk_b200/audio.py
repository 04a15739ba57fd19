"""wav2vec2 audio features on the B200: the conditioning input of the predictor / refiner.

Replaces, for the reference's demo / data-preparation path (scripts/demo.py:76-77, :105-111),

    inputs   = aud_p(audio, sampling_rate=16000, return_tensors="pt")      # Wav2Vec2Processor: zero mean / unit variance
    audio_ft = aud_m(**inputs).last_hidden_state                           # transformers.Wav2Vec2Model (wav2vec2-base)
    aud      = linear_interpolation(audio_ft.transpose(1, 2), int(len / 16000 * 30))     # data_utils/utils.py:14-16

with one recorded plan of C-ABI launches (``Wav2VecPlan``) over fixed device buffers:

    pt_w2v_frontend      normalisation + conv0 (k10 s5) + GroupNorm + GELU, one pass over the waveform
    pt_gemm_bf16 x 6     conv1-4 (k3 s2), conv5-6 (k2 s2), GELU epilogue, valid convolution over token-major rows
    pt_layernorm_rows    feature_projection.layer_norm (512)
    pt_gemm_bf16         feature_projection.projection 512 -> 768 (twice: flat rows and the halo buffer of pos_conv)
    pt_gemm_bf16         pos_conv_embed: Conv1d(768, 768, k128, pad 64, groups 16) + GELU in one launch: six 128-column output
                         tiles, each a 128-tap contraction over the 192 input channels its (2-4) groups read (a_col_table)
    pt_layernorm_rows    encoder.layer_norm(h + pos)
    12 x [ QKV GEMM (N = 2304), pt_attention (12 heads x 64, keys streamed), out-proj GEMM, LN(h + .),
           FF 768 -> 3072 GELU -> 768, LN(h + .) ]                       post-norm Wav2Vec2EncoderLayer
    pt_w2v_interp        50 fps rows -> [B, 768, t] fp32 at 30 fps

Arithmetic: bf16 tensor-core operands, fp32 accumulation / statistics / softmax (the reference runs fp32 torch); the
observed deviation from the fp32 model is stated in tests/test_gpu_audio.py.  There is no CPU fallback.
"""
from __future__ import annotations

import math

import torch

from . import _lib
from .engine import BF16, F32, Mat, Plan, RowBuf, Weight, _ceil, _finish, prep_conv, prep_linear
from .ops import ACT_GELU, ACT_NONE

W2V_CONV = ((512, 10, 5), (512, 3, 2), (512, 3, 2), (512, 3, 2), (512, 3, 2), (512, 2, 2), (512, 2, 2))
W2V_HIDDEN, W2V_HEADS, W2V_HEAD_DIM, W2V_FF, W2V_LAYERS = 768, 12, 64, 3072, 12
W2V_POS_K, W2V_POS_GROUPS = 128, 16
# first input channel of the 192-wide slab each 128-column output tile of pos_conv reads (multiples of 8: TMA alignment)
W2V_POS_OFFSETS = (0, 96, 240, 384, 480, 576)


def feat_lengths(L):
    """Frames after each feature-encoder layer (valid convolutions): transformers' _get_feat_extract_output_lengths."""
    out = []
    for _, k, s in W2V_CONV:
        L = (L - k) // s + 1
        out.append(L)
    return out


def _strip(sd):
    """Accept Wav2Vec2Model / Wav2Vec2ForCTC state dicts (the latter prefixes everything with 'wav2vec2.')."""
    if any(k.startswith("wav2vec2.") for k in sd):
        sd = {k[len("wav2vec2."):]: v for k, v in sd.items() if k.startswith("wav2vec2.")}
    return sd


def _pos_conv_weight(sd):
    """Effective weight of the weight-normalised positional convolution: w = g * v / |v| with the norm over (out, in)
    per tap (torch weight_norm(dim=2)); old (weight_g / weight_v) and parametrised (original0 / original1) key names."""
    p = "encoder.pos_conv_embed.conv."
    if p + "weight" in sd:
        return sd[p + "weight"].float()
    if p + "weight_g" in sd:
        g, v = sd[p + "weight_g"].float(), sd[p + "weight_v"].float()
    else:
        g, v = sd[p + "parametrizations.weight.original0"].float(), sd[p + "parametrizations.weight.original1"].float()
    return v * (g / v.norm(dim=(0, 1), keepdim=True))


class Wav2VecWeights:
    """Device-resident, GEMM-ready weights of a wav2vec2-base ``Wav2Vec2Model`` (feat_extract_norm = 'group',
    conv_bias = False, do_stable_layer_norm = False -- the configuration of facebook/wav2vec2-base-960h)."""

    def __init__(self, state_dict, device="cuda"):
        sd = _strip(state_dict)
        dev = torch.device(device)
        f = lambda k: sd[k].detach().float()
        fe = "feature_extractor.conv_layers."
        if fe + "1.layer_norm.weight" in sd or fe + "0.conv.bias" in sd:
            raise ValueError("only the 'group'-norm, bias-free feature encoder of wav2vec2-base is built")
        w0 = f(fe + "0.conv.weight")
        if tuple(w0.shape) != (512, 1, 10):
            raise ValueError("unexpected conv0 geometry %s" % (tuple(w0.shape),))
        self.conv0_w = w0.reshape(512, 10).contiguous().to(dev)
        self.gn_g = f(fe + "0.layer_norm.weight").to(dev)
        self.gn_b = f(fe + "0.layer_norm.bias").to(dev)
        self.convs = []
        for i in range(1, 7):
            w = f(fe + "%d.conv.weight" % i)
            assert tuple(w.shape) == (512, 512, W2V_CONV[i][1]), w.shape
            self.convs.append(prep_conv(w, torch.zeros(512), dev))
        vec = lambda k: f(k).contiguous().to(dev)
        self.fp_ln = (vec("feature_projection.layer_norm.weight"), vec("feature_projection.layer_norm.bias"))
        self.fp = prep_linear(f("feature_projection.projection.weight"), f("feature_projection.projection.bias"), dev)
        # positional convolution: block-diagonal (16 groups of 48 -> 48 channels), one zero-padded [128, 128 taps, 192]
        # matrix per 128-column output tile
        wp, bp = _pos_conv_weight(sd), f("encoder.pos_conv_embed.conv.bias")
        cg = W2V_HIDDEN // W2V_POS_GROUPS
        assert tuple(wp.shape) == (W2V_HIDDEN, cg, W2V_POS_K), wp.shape
        tiles = []
        for j, off in enumerate(W2V_POS_OFFSETS):
            wm = torch.zeros(128, W2V_POS_K, 192)
            for n in range(128):
                g = (128 * j + n) // cg
                lo = cg * g - off
                assert 0 <= lo and lo + cg <= 192
                wm[n, :, lo:lo + cg] = wp[128 * j + n].t()
            tiles.append(wm)
        # ONE weight matrix [768, 128 taps * 192]: the launch covers all six output tiles, tile j reading its 192-channel slab
        # at column W2V_POS_OFFSETS[j] of the activations (pt_gemm_desc.a_col_table)
        self.pos = _finish(torch.cat(tiles, 0), bp, W2V_POS_K, 192, dev)
        self.pos.cin = cg                                  # algorithmic FLOPs: 48 input channels per output channel
        self.enc_ln = (vec("encoder.layer_norm.weight"), vec("encoder.layer_norm.bias"))
        self.layers = []
        for l in range(W2V_LAYERS):
            p = "encoder.layers.%d." % l
            qkv_w = torch.cat([f(p + "attention.%s_proj.weight" % n) for n in "qkv"], 0)
            qkv_b = torch.cat([f(p + "attention.%s_proj.bias" % n) for n in "qkv"], 0)
            self.layers.append(dict(
                qkv=prep_linear(qkv_w, qkv_b, dev),
                out=prep_linear(f(p + "attention.out_proj.weight"), f(p + "attention.out_proj.bias"), dev),
                ln1=(vec(p + "layer_norm.weight"), vec(p + "layer_norm.bias")),
                ff1=prep_linear(f(p + "feed_forward.intermediate_dense.weight"), f(p + "feed_forward.intermediate_dense.bias"), dev),
                ff2=prep_linear(f(p + "feed_forward.output_dense.weight"), f(p + "feed_forward.output_dense.bias"), dev),
                ln2=(vec(p + "final_layer_norm.weight"), vec(p + "final_layer_norm.bias"))))
        self.device = dev


class Wav2VecPlan:
    """The launch plan of one (utterances, samples) bucket: ``run(wave [B, L] fp32) -> aud [B, 768, t_out] fp32``."""

    def __init__(self, W: Wav2VecWeights, B, L, t_out=None, sampling_rate=16000, fps=30, use_graph=False):
        dev = W.device
        if L < 400:
            raise ValueError("wav2vec2's feature encoder needs at least 400 samples (got %d)" % L)
        self.W, self.B, self.L = W, B, L
        Ts = feat_lengths(L)
        self.T = T = Ts[-1]
        # scripts/demo.py:111: int(speech_array.shape[0] / sampling_rate * 30)
        self.t_out = int(L / sampling_rate * fps) if t_out is None else t_out
        lib = _lib.lib()
        p = self.plan = Plan(dev)
        # test / bring-up hook: name -> (number of recorded launches after which `tensor` holds that quantity, tensor, rows, T)
        self.marks = {}
        mark = lambda name, t, rows_per_sample, t_valid: self.marks.__setitem__(name, (len(p.steps), t, rows_per_sample, t_valid))
        self.wave = torch.zeros(B, L, dtype=F32, device=dev)
        ws_bytes = lib.pt_w2v_frontend_workspace_bytes(B)
        self.ws = torch.zeros((ws_bytes + 7) // 8, dtype=torch.float64, device=dev)
        # feature encoder: token-major rows [B, P_i, 512], P_i even so that a stride-2 layer reads [P_i / 2, 1024] rows
        P = [_ceil(t, 2) for t in Ts]
        self.feat = [torch.zeros(B, P[i], 512, dtype=BF16, device=dev) for i in range(6)]
        self.c = Mat(B * T, 512, dev)                       # conv6 output, flat rows
        p.call(lib.pt_w2v_frontend, self.wave.data_ptr(), B, L, W.conv0_w.data_ptr(), W.gn_g.data_ptr(), W.gn_b.data_ptr(),
               self.feat[0].data_ptr(), P[0], self.ws.data_ptr(), ws_bytes)
        mark("conv0", self.feat[0], P[0], Ts[0])
        for i in range(1, 7):
            src = self.feat[i - 1]
            last = i == 6
            dst = self.c.t if last else self.feat[i]
            p.gemm(src, B * P[i - 1], 512, W.convs[i - 1], dst, M=B * P[i - 1] // 2, C_=512, s=2, base_row=0,
                   rows_per_sample=P[i - 1] // 2, t_valid=Ts[i], oss=(T if last else P[i]) * 512, ors=512, act=ACT_GELU,
                   use_bias=False)
            mark("conv%d" % i, dst, T if last else P[i], Ts[i])
        # feature projection
        H = W2V_HIDDEN
        M = self.M = B * T
        p.call(lib.pt_layernorm_rows, self.c.t.data_ptr(), None, W.fp_ln[0].data_ptr(), W.fp_ln[1].data_ptr(), 1e-5,
               self.c.t.data_ptr(), M, 512)
        self.h = Mat(M, H, dev)
        self.halo = RowBuf(B, T, H, H=W2V_POS_K // 2, device=dev)
        p.linear(self.c, M, W.fp, self.h)
        mark("projected", self.h.t, T, T)
        hb = self.halo
        p.gemm(self.c.t, M, 512, W.fp, hb.t, M=M, C_=512, rows_per_sample=T, t_valid=T, oss=hb.Tp * H, ors=H,
               out_off=hb.H * H)
        # positional convolution + GELU (the k = 128 'same' convolution drops its last output frame)
        self.pos = Mat(M, H, dev)
        p.gemm(hb.t, B * hb.Tp, H, W.pos, self.pos.t, M=B * hb.Tp, C_=192, base_row=0, rows_per_sample=hb.Tp, t_valid=T,
               oss=T * H, ors=H, act=ACT_GELU, a_col_table=list(W2V_POS_OFFSETS))
        ln = lambda x, res, gb, out: p.call(lib.pt_layernorm_rows, x.data_ptr(), None if res is None else res.data_ptr(),
                                            gb[0].data_ptr(), gb[1].data_ptr(), 1e-5, out.data_ptr(), M, H)
        mark("pos_conv", self.pos.t, T, T)
        ln(self.pos.t, self.h.t, W.enc_ln, self.h.t)
        mark("hidden0", self.h.t, T, T)
        # encoder layers (post-norm)
        self.qkv = Mat(M, 3 * H, dev)
        self.att = Mat(M, H, dev)
        self.tmp = Mat(M, H, dev)
        self.ff = Mat(M, W2V_FF, dev)
        scale = 1.0 / math.sqrt(W2V_HEAD_DIM)
        for l, lw in enumerate(W.layers):
            p.linear(self.h, M, lw["qkv"], self.qkv)
            q = self.qkv.t.data_ptr()
            p.call(lib.pt_attention, q, 3 * H, q + 2 * H, 3 * H, q + 4 * H, 3 * H, self.att.t.data_ptr(), H, B, W2V_HEADS,
                   W2V_HEAD_DIM, T, T, scale, flops=4.0 * B * W2V_HEADS * T * T * W2V_HEAD_DIM)
            p.linear(self.att, M, lw["out"], self.tmp)
            ln(self.tmp.t, self.h.t, lw["ln1"], self.h.t)
            p.linear(self.h, M, lw["ff1"], self.ff, act=ACT_GELU)
            p.linear(self.ff, M, lw["ff2"], self.tmp)
            ln(self.tmp.t, self.h.t, lw["ln2"], self.h.t)
            mark("hidden%d" % (l + 1), self.h.t, T, T)
        # 50 fps rows -> 30 fps channel-first fp32
        self.aud = torch.zeros(B, H, self.t_out, dtype=F32, device=dev)
        p.call(lib.pt_w2v_interp, self.h.t.data_ptr(), 0, T, B, T, H, self.aud.data_ptr(), self.t_out)
        self.launches = p.launches
        self.flops = sum(s[2] for s in p.steps)

    def run(self, wave):
        """wave: [B, L] (or [L] when B = 1) fp32 samples at 16 kHz, host or device."""
        w = torch.as_tensor(wave)
        if w.dim() == 1:
            w = w[None]
        if tuple(w.shape) != (self.B, self.L):
            raise ValueError("this plan was built for waveforms of shape %s, got %s" % ((self.B, self.L), tuple(w.shape)))
        self.wave.copy_(w.to(torch.float32), non_blocking=True)
        self.plan.run()
        return self.aud

    def run_until(self, wave, name):
        """Bring-up / test hook: replay the plan up to the mark `name`, return that stage as fp32 [B, T_stage, C]."""
        n, t, rps, tv = self.marks[name]
        self.wave.copy_(torch.as_tensor(wave).to(torch.float32).reshape(self.B, self.L))
        full, self.plan.steps = self.plan.steps, self.plan.steps[:n]
        try:
            self.plan.run()
        finally:
            self.plan.steps = full
        return t.reshape(self.B, rps, -1)[:, :tv].float()

    def last_hidden_state(self):
        """[B, T, 768] fp32 copy of the encoder output of the last run (Wav2Vec2Model(...).last_hidden_state)."""
        return self.h.t.view(self.B, self.T, W2V_HIDDEN).float()


class Wav2Vec2Features:
    """The reference-facing callable: ``Wav2Vec2Features.from_hf(aud_m)(audio)`` = scripts/demo.py:105-111."""

    def __init__(self, weights: Wav2VecWeights, sampling_rate=16000, fps=30, cache=4):
        from .engine import PlanCache
        self.W, self.sr, self.fps = weights, sampling_rate, fps
        self._plans = PlanCache(cache)

    @classmethod
    def from_hf(cls, model, device="cuda", **kw):
        """model: a transformers ``Wav2Vec2Model`` (or ``Wav2Vec2ForCTC``), or its state dict."""
        sd = model if isinstance(model, dict) else model.state_dict()
        return cls(Wav2VecWeights(sd, device), **kw)

    def __call__(self, audio, sampling_rate=None, frames=None):
        if sampling_rate is not None and sampling_rate != self.sr:
            raise ValueError("wav2vec2 expects %d Hz audio (got %s)" % (self.sr, sampling_rate))
        w = torch.as_tensor(audio)
        if w.dim() == 1:
            w = w[None]
        B, L = w.shape
        plan = self._plans.get((B, L, frames), lambda: Wav2VecPlan(self.W, B, L, frames, self.sr, self.fps))
        return plan.run(w).clone()          # the plan's output buffer is reused by the next call
