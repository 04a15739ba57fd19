"""CPU restatement of the reference's two sliding-window loops (TEST INFRASTRUCTURE, like the rest of oracle/).

  * ``windows_eval``  -- scripts/test_holisticbody.py:239-262: windows [0,180), then [150+180(i-1), 150+180i), each one
                         All_In_One_Model('infer_on_batch') = predictor -> refiner (nets/ai1.py:44-60), the previous
                         output fed back as known context through (input_gt, input_mask);
  * ``windows_demo``  -- scripts/demo.py:116-181: stage 1 predictor windows (180, stride 150) collecting codes, one PQ-VAE
                         decode of the concatenated codes, linear interpolation to t, stage 2 refiner windows (176,
                         stride 146).  ``demo.py:151`` hard-codes two windows; the concatenation is generalised to
                         ``[c0] + [c[:, 3:] for c in rest]`` as SURVEY.md 3.2 prescribes.

Single-window arithmetic is ``oracle/probtalk_oracle.py``.  The two RNG draws of the sampler are inputs (``noise`` = one
(u, gumbel) pair per predictor window), or the codes of every window can be teacher-forced (``codes``).
Pinned against the reference's own wrappers driven by a literal transcription of the script loops:
``tests/test_oracle_vs_reference.py`` (needs /root/reference) and the committed golden vectors
``tests/golden/window_vectors.pt`` (``oracle/make_golden_windows.py``).
Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may import this module.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from . import probtalk_oracle as O


def eval_slices(t):
    """scripts/test_holisticbody.py:240-246 (slices that start beyond the clip are dropped: the reference would crash)."""
    n = 1 + math.ceil((t - 180) / 150)
    out = []
    for i in range(max(n, 1)):
        s = 0 if i == 0 else 150 + 180 * (i - 1)
        e = 180 if i == 0 else 150 + 180 * i
        if s < t:
            out.append((s, min(e, t)))
    return out


def demo_slices(t, length, cover=30):
    """scripts/demo.py:118, 133-135 / :158-162."""
    n = math.ceil((t - cover) / (length - cover))
    return [((length - cover) * i, min((length - cover) * i + length, t)) for i in range(n)]


def _predict(aud, gt, mask, ids, pred_state, vq_sd, noise, codes):
    """One predictor window: sampled with injected noise, or decoded from teacher-forced codes."""
    if codes is not None:
        return O.predictor_decode(codes.long(), gt * mask, mask, vq_sd, aud.shape[2]), codes
    u, gum = noise
    return O.predictor_infer_on_batch(aud, gt, mask, ids, pred_state, vq_sd, u, gum)


def windows_eval(aud, gt_poses, mask, ids, pred_state, vq_sd, ref_sd, noise=None, codes=None):
    """aud [B,768,t], gt_poses [B,376,t], mask [B,1,t], ids [B] -> (poses [B,376,t], [codes per window]).
    noise / codes: one entry per window (see module docstring)."""
    input_gt, input_mask = gt_poses.clone(), mask.clone()
    pred, used = None, []
    for w, (s, e) in enumerate(eval_slices(aud.shape[-1])):
        a, g, m = aud[..., s:e], input_gt[..., s:e], input_mask[..., s:e]
        body, cd = _predict(a, g, m, ids, pred_state, vq_sd, None if noise is None else noise[w],
                            None if codes is None else codes[w])
        pred_0 = O.refiner_infer_on_batch(a, g, m, ids, body, ref_sd)          # ai1.py:52-56: pred_poses = predictor output
        used.append(cd)
        pred = pred_0 if pred is None else torch.cat([pred, pred_0[..., 30:]], -1)
        input_gt[..., s:e] = pred_0
        input_mask[..., s:e] = 1
    return gt_poses * mask + pred * (1 - mask), used


def windows_demo(aud, ids, pred_state, vq_sd, ref_sd, noise=None, codes=None, groups=4):
    """aud [B,768,t], ids [B] -> (poses [B,376,t], codes [B, T'_total, G], [codes per window])."""
    B, _, t = aud.shape
    input_gt = torch.zeros(B, 376, t)
    input_mask = torch.zeros(B, 1, t)
    code_list = []
    for w, (s, e) in enumerate(demo_slices(t, 180)):
        pred_0, cd = _predict(aud[..., s:e], input_gt[..., s:e], input_mask[..., s:e], ids, pred_state, vq_sd,
                              None if noise is None else noise[w], None if codes is None else codes[w])
        code_list.append(cd)
        input_gt[..., s:e] = pred_0
        input_mask[..., s:e] = 1
    all_codes = torch.cat([code_list[0]] + [c[:, 3:] for c in code_list[1:]], dim=1)
    latents = O.pq_gather(all_codes.long(), O.pq_codebooks(vq_sd, groups)).transpose(1, 2).contiguous()
    prelim = F.interpolate(O.decoder_sc(latents, vq_sd), size=t, align_corners=False, mode="linear")
    input_mask = torch.zeros(B, 1, t)
    pred_all = None
    for s, e in demo_slices(t, 176):
        pred_0 = O.refiner_infer_on_batch(aud[..., s:e], input_gt[..., s:e], input_mask[..., s:e], ids, prelim[..., s:e], ref_sd)
        pred_all = pred_0 if pred_all is None else torch.cat([pred_all, pred_0[..., 30:]], -1)
        input_gt[..., s:e] = pred_0
        input_mask[..., s:e] = 1
    return pred_all, all_codes, code_list
