"""Deterministic random-init checkpoints in the reference's layout (no network, no dataset).

The reference's constructor init is unusable for parity work: ``zero_module`` zeroes every style
block's output Linear, so each transformer layer collapses to ``x = LN(x)``, and fresh BatchNorm
statistics are the identity (caveat 1).  The generator below draws EVERY tensor from a
seeded CPU generator with fan-in scaled weights, non-trivial BatchNorm statistics and non-zero style
outputs, so attention / FFN / style bugs are visible.  The same seed gives the same values on any
machine with the same torch build (CPU generator).
"""
from __future__ import annotations

import math

import torch

from .nets import modules as M

N_CLASSES = 4
POSE_DIM = 376
AUDIO_DIM = 768


def _randomize(module: torch.nn.Module, seed: int, gain: float = 1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd = module.state_dict()
    out = {}
    for k, v in sd.items():
        shape = tuple(v.shape)
        if k.endswith("num_batches_tracked"):
            out[k] = torch.tensor(100, dtype=torch.int64)
        elif "position_encoding" in k:
            out[k] = v.clone()                                        # sinusoid tables stay as constructed (fp64)
        elif k.endswith("running_mean"):
            out[k] = torch.randn(shape, generator=g) * 0.1
        elif k.endswith("running_var"):
            out[k] = torch.rand(shape, generator=g) + 0.5
        elif ".norm" in k and k.endswith(".weight") or k.endswith("feature_map.1.weight"):
            out[k] = torch.rand(shape, generator=g) + 0.5             # BN / LN gains in [0.5, 1.5)
        elif k.endswith(".bias") or k.endswith("in_proj_bias") or k.endswith("feature_map.1.bias"):
            out[k] = torch.randn(shape, generator=g) * 0.05
        elif "embeddings_" in k:
            bound = math.sqrt(6.0 / (shape[0] + shape[1]))            # xavier_uniform, the reference's codebook init
            out[k] = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif "ema_" in k:
            out[k] = torch.zeros(shape)
        elif k.endswith("tok_emb.weight") or k.endswith("class_embed.weight"):
            out[k] = torch.randn(shape, generator=g)
        elif "pad_" in k:
            out[k] = torch.randn(shape, generator=g) * 0.02
        elif len(shape) >= 2:
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            if len(shape) == 3 and ("_up_" in k):
                fan_in = shape[0] * shape[2] / 2.0                    # ConvTranspose1d [Cin, Cout, k], stride 2
            out[k] = torch.randn(shape, generator=g) * (gain / math.sqrt(fan_in))
        else:
            out[k] = torch.randn(shape, generator=g) * 0.05
        out[k] = out[k].to(v.dtype) if out[k].dtype != v.dtype and not k.endswith("num_batches_tracked") else out[k]
    return out


def make_vq_state(seed=0):
    """state_dict of VQVAE_SC(376, 512, 128, 1024, 2, groups=4)  (vq_teacher.py:64-67)."""
    return _randomize(M.VQVAE_SC(POSE_DIM, 512, 128, 1024, 2, 4), seed)


def make_predictor_state(seed=1):
    """{'AudEnc', 'Predictor', 'TextEnc'} like predictor.TrainWrapper.state_dict (predictor.py:128-143)."""
    return {
        "AudEnc": _randomize(M.ConditionEncoder(AUDIO_DIM, 256), seed),
        "Predictor": _randomize(M.PredictorNet(POSE_DIM, 128, 512, N_CLASSES, 4, 8), seed + 1),
        "TextEnc": _randomize(M.ConditionEncoder(AUDIO_DIM, 256), seed + 2),
    }


def make_refiner_state(seed=5):
    """{'HFTransformer', 'AudEnc'} like refiner.TrainWrapper.state_dict (refiner.py:141-156)."""
    return {"HFTransformer": _randomize(M.RTransformer(POSE_DIM, AUDIO_DIM, 512, 6, N_CLASSES), seed), "AudEnc": {}}


def synthetic_poses(B, T, seed):
    """Normalised 6-D pose clips ~ N(0.6, 1.45^2) clipped to +-6 (statistics of the real demo clips)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(B, POSE_DIM, T, generator=g) * 1.45 + 0.6).clamp_(-6, 6)


def synthetic_audio(B, T, seed):
    """wav2vec2-shaped per-frame features [B, 768, T] ~ N(0, 1)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(B, AUDIO_DIM, T, generator=g)


def make_wav2vec2(seed=11):
    """A transformers ``Wav2Vec2Model`` of wav2vec2-base geometry (the architecture scripts/demo.py:76-77 downloads as
    facebook/wav2vec2-base-960h; no network here) with every tensor drawn from a seeded CPU generator: fan-in scaled
    weights (He gain for the GELU convolutions), LayerNorm / GroupNorm gains in [0.5, 1.5), small biases.  The library's
    own init (std 0.02 Linears) would leave every encoder layer close to x = LN(x) and hide attention / FF bugs."""
    from transformers import Wav2Vec2Config, Wav2Vec2Model
    model = Wav2Vec2Model(Wav2Vec2Config()).eval()
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd = model.state_dict()
    new = {}
    for k, v in sd.items():
        shape = tuple(v.shape)
        if k.endswith("parametrizations.weight.original0"):
            continue                                                  # set from original1 below
        if "layer_norm.weight" in k:
            new[k] = torch.rand(shape, generator=g) + 0.5
        elif k.endswith(".bias"):
            new[k] = torch.randn(shape, generator=g) * 0.05
        elif k.endswith("parametrizations.weight.original1"):
            vv = torch.randn(shape, generator=g) * (2.0 / math.sqrt(shape[1] * shape[2]))
            new[k] = vv
            gain = torch.rand(1, 1, shape[2], generator=g) + 0.5
            new[k.replace("original1", "original0")] = vv.norm(dim=(0, 1), keepdim=True) * gain
        elif len(shape) == 3:
            new[k] = torch.randn(shape, generator=g) * math.sqrt(2.0 / (shape[1] * shape[2]))
        elif len(shape) == 2:
            new[k] = torch.randn(shape, generator=g) / math.sqrt(shape[1])
        else:
            new[k] = torch.randn(shape, generator=g) * 0.05
    model.load_state_dict(new, strict=True)
    return model


def synthetic_waveform(B, L, seed, sr=16000):
    """[B, L] fp32 'speech-like' audio: a few drifting harmonics under a slow envelope plus a noise floor, peak ~0.3."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    t = torch.arange(L, dtype=torch.float64) / sr
    out = torch.zeros(B, L, dtype=torch.float64)
    for b in range(B):
        f0 = 90.0 + 120.0 * torch.rand(1, generator=g).item()
        env = 0.55 + 0.45 * torch.sin(2 * math.pi * (1.5 + torch.rand(1, generator=g).item() * 3.0) * t)
        for h in range(1, 7):
            ph = 2 * math.pi * torch.rand(1, generator=g).item()
            out[b] += env * torch.sin(2 * math.pi * f0 * h * t * (1.0 + 0.01 * torch.sin(2 * math.pi * 0.7 * t)) + ph) / (h * 6.0)
        out[b] += 0.02 * torch.randn(L, generator=g, dtype=torch.float64)
    return out.float()
