"""Batched inference driver: sliding windows over utterances, sharded across GPUs.

Re-implements the two window loops of the reference on top of the CUDA plans:
  * ``windows_eval``  -- scripts/test_holisticbody.py:239-262: windows [0,180), then [150+180(i-1), 150+180i),
                         each window = predictor -> refiner, 30-frame context fed back through (gt, mask);
  * ``windows_demo``  -- scripts/demo.py:116-181: stage 1 predictor windows (180, stride 150) collecting codes,
                         one PQ-VAE decode of all codes, linear interpolation to t, stage 2 refiner windows
                         (176, stride 146).  The reference hard-codes two windows (``demo.py:151``); the
                         concatenation ``[c0] + [c[:, 3:] ...]`` is generalised to any number of windows.
Windows of one utterance are sequential (feedback); parallelism is across utterances x samples, and across
GPUs by sharding utterances (one process per GPU, no collective until the final gather).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from . import engine as E


def shard_range(n: int, rank: int, world: int):
    """Contiguous block partition of n independent units; sizes differ by at most one."""
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def gather_to_rank0(local: torch.Tensor, counts, group=None):
    """The path's only collective: gather per-rank result blocks (possibly unequal) on rank 0.
    Works on any backend (NCCL on GPUs, gloo in the CPU tests).  Returns the concatenation on rank 0, else None."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local
    mx = max(counts)
    pad = local
    if local.shape[0] < mx:
        pad = torch.cat([local, local.new_zeros((mx - local.shape[0],) + tuple(local.shape[1:]))])
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad.contiguous(), bufs, dst=0, group=group)
    if rank != 0:
        return None
    return torch.cat([b[:c] for b, c in zip(bufs, counts)])


def eval_window_bounds(t: int):
    """scripts/test_holisticbody.py:240-246."""
    n = 1 + math.ceil(max(t - 180, 0) / 150)
    out = []
    for i in range(n):
        s = 0 if i == 0 else 150 + 180 * (i - 1)
        e = 180 if i == 0 else 150 + 180 * i
        if s >= t:
            break
        out.append((s, min(e, t)))
    return out


def demo_window_bounds(t: int, length: int, cover: int = 30):
    """scripts/demo.py:118,133-135 (and :158-162 for the refiner with length 176)."""
    n = math.ceil((t - cover) / (length - cover))
    return [((length - cover) * i, min((length - cover) * i + length, t)) for i in range(n)]


class HolisticPipeline:
    """predictor + refiner + PQ-VAE over fixed weights; plans are cached per (batch, frames) bucket."""

    def __init__(self, vq_state, pred_state, ref_state, device="cuda", precision="bf16"):
        self.device = torch.device(device)
        self.precision = precision
        self.VW = E.VQVAEWeights(vq_state, self.device, precision=precision)
        self.PW = E.PredictorWeights(pred_state, self.device, precision)
        self.RW = E.RefinerWeights(ref_state, self.device, precision)
        self._pred, self._ref, self._dec = E.PlanCache(), E.PlanCache(), E.PlanCache(4)
        self.generator = None          # torch.Generator for the sampler noise (None = global CUDA RNG)

    @classmethod
    def from_weights(cls, VW, PW, RW):
        """A pipeline over already-prepared weights (engine.VQVAEWeights / PredictorWeights / RefinerWeights)."""
        self = cls.__new__(cls)
        self.device, self.precision = torch.device(PW.device), PW.precision
        self.VW, self.PW, self.RW = VW, PW, RW
        self._pred, self._ref, self._dec = E.PlanCache(), E.PlanCache(), E.PlanCache(4)
        self.generator = None
        return self

    def predictor(self, B, T):
        return self._pred.get((B, T), lambda: E.PredictorPlan(self.PW, self.VW, B, T))

    def refiner(self, B, T, Tp):
        return self._ref.get((B, T, Tp), lambda: E.RefinerPlan(self.RW, B, T, Tp))

    def window(self, aud, gt, mask, labels, noise=None, force_codes=None):
        """One All_In_One_Model('infer_on_batch') call (ai1.py:44-60) on already-batched inputs."""
        B, _, T = aud.shape
        poses, codes = self.predictor(B, T).run(aud, gt, mask, labels, noise=noise, generator=self.generator,
                                                force_codes=force_codes)
        out = self.refiner(B, T, T).run(aud, gt, mask, poses, labels)
        return out, codes

    def windows_eval(self, aud, gt, mask, labels, noise=None, force_codes=None, return_codes=False):
        """aud [B,768,t], gt [B,376,t], mask [B,1,t] -> poses [B,376,t]  (test_holisticbody.py:239-262).
        noise / force_codes: optional lists with one entry per window ((u, gumbel) to inject / codes to teacher-force)."""
        input_gt, input_mask = gt.clone(), mask.clone()
        pred, used = None, []
        for w, (s, e) in enumerate(eval_window_bounds(aud.shape[-1])):
            p0, cd = self.window(aud[..., s:e].contiguous(), input_gt[..., s:e].contiguous(),
                                 input_mask[..., s:e].contiguous(), labels, noise=None if noise is None else noise[w],
                                 force_codes=None if force_codes is None else force_codes[w])
            used.append(cd.clone())
            pred = p0.clone() if pred is None else torch.cat([pred, p0[..., 30:]], -1)
            input_gt[..., s:e] = p0
            input_mask[..., s:e] = 1
        out = gt * mask + pred * (1 - mask)
        return (out, used) if return_codes else out

    def windows_demo(self, aud, labels, noise=None, force_codes=None):
        """aud [B,768,t] -> poses [B,376,t]  (scripts/demo.py:116-181, generalised to any t)."""
        B, _, t = aud.shape
        dev = self.device
        input_gt = torch.zeros(B, self.PW.mot_dim, t, device=dev)
        input_mask = torch.zeros(B, 1, t, device=dev)
        code_list = []
        for w, (s, e) in enumerate(demo_window_bounds(t, 180)):
            plan = self.predictor(B, e - s)
            p0, codes = plan.run(aud[..., s:e].contiguous(), input_gt[..., s:e].contiguous(),
                                 input_mask[..., s:e].contiguous(), labels, generator=self.generator,
                                 noise=None if noise is None else noise[w],
                                 force_codes=None if force_codes is None else force_codes[w])
            code_list.append(codes.clone())
            input_gt[..., s:e] = p0
            input_mask[..., s:e] = 1
        codes = torch.cat([code_list[0]] + [c[:, 3:] for c in code_list[1:]], dim=1)
        dec = self._dec.get((B, codes.shape[1]),
                            lambda: E.DecoderPlan(self.VW, B, codes.shape[1], codes=torch.zeros_like(codes)))
        prelim = F.interpolate(dec.run(codes=codes), size=t, align_corners=False, mode="linear")
        input_mask.zero_()
        pred = None
        for s, e in demo_window_bounds(t, 176):
            p0 = self.refiner(B, e - s, e - s).run(aud[..., s:e].contiguous(), input_gt[..., s:e].contiguous(),
                                                   input_mask[..., s:e].contiguous(), prelim[..., s:e].contiguous(), labels)
            pred = p0.clone() if pred is None else torch.cat([pred, p0[..., 30:]], -1)
            input_gt[..., s:e] = p0
            input_mask[..., s:e] = 1
        return pred, codes


    def infer_from_audio(self, features, waveform, labels, sampling_rate=16000, norm_stats=None, stand=False, noise=None):
        """The whole of scripts/demo.py::infer between `librosa.load` and `get_vertices` (:104-205), on the device:
        waveform [L] / [B, L] (16 kHz) -> wav2vec2 features at 30 fps (`features` = audio.Wav2Vec2Features, f1) ->
        predictor windows, PQ-VAE decode, refiner windows (`windows_demo`) -> SMPL-X parameter rows (f2:
        denormalize -> 6-D -> axis-angle -> part2full).  Returns (smplx [B, t, 265] , poses [B, 376, t], codes).
        ``norm_stats``: the reference's (mean, std) pair of data_utils/norm_stats.npy, or None (no de-normalisation)."""
        from . import postprocess as P
        aud = features(waveform, sampling_rate=sampling_rate)
        B = aud.shape[0]
        labels = torch.as_tensor(labels, dtype=torch.int32, device=self.device).reshape(-1)
        if labels.numel() == 1 and B > 1:
            labels = labels.repeat(B)
        poses, codes = self.windows_demo(aud, labels, noise=noise)
        if norm_stats is not None:
            smplx = P.poses_to_smplx(poses, norm_stats[0], norm_stats[1], P.c_index_6d(), stand=stand)
        else:
            smplx = P.poses_to_smplx(poses, stand=stand)
        return smplx, poses, codes


def run_sharded(n_units: int, compute, group=None):
    """Shard ``n_units`` independent clips/utterances over the ranks of ``group``; ``compute(start, end)`` returns
    this rank's result block [end-start, ...]; rank 0 gets the concatenation in unit order."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    s, e = shard_range(n_units, rank, world)
    local = compute(s, e)
    if world == 1:
        return local
    counts = [shard_range(n_units, r, world)[1] - shard_range(n_units, r, world)[0] for r in range(world)]
    return gather_to_rank0(local, counts, group)
