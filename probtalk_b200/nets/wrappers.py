"""Drop-in model wrappers with the reference's constructor / checkpoint / inference API.

    getattr(probtalk_b200.nets, name)(args, config)        name in {s2g_body_vqt, s2g_body_predictor, s2g_body_refiner}

mirrors ``getattr(nets, model_name)(args, config)`` of the reference (nets/init_model.py:8,
nets/ai1.py:11); SURVEY.md 8(b) is the contract:
  * ``state_dict()`` / ``load_state_dict()`` use the reference's dictionary layout
    ({'VQ'}, {'AudEnc','Predictor','TextEnc'}, {'HFTransformer','AudEnc'}), 'module.' prefixes stripped;
  * predictor / refiner fetch the PQ-VAE from ``torch.load(config.Model.vq_path)['generator']['VQ']`` in their
    constructor (predictor.py:81-83, refiner.py:88-90) when that file exists;
  * ``infer_on_batch`` / ``infer_to_code`` / ``infer_on_audio`` keep the reference signatures and return
    ``(tensor, seconds)`` (seconds measured with CUDA events instead of time.time()).
All arithmetic runs in the CUDA library through ``probtalk_b200.engine``; there is no CPU path.
"""
from __future__ import annotations

import os
from collections import OrderedDict

import torch

from .. import engine as E
from .. import ops
from . import modules as M


class QuantizationOutput:
    """Same 2-field record as the reference (vqvae_modules.py:24-29); in eval mode the code indices travel
    in ``.loss`` (vqvae_modules.py:489-492)."""

    def __init__(self, quantized, loss, q_loss=0, lr_loss=0):
        self.quantized, self.loss, self.q_loss, self.lr_loss = quantized, loss, q_loss, lr_loss


def _strip(sd):
    return OrderedDict((k.replace("module.", ""), v) for k, v in sd.items())


def _n_classes(config):
    root = getattr(getattr(config, "Data", None), "data_root", None)
    if root and os.path.isdir(root):
        return len(os.listdir(root))            # data_utils/consts.py:13-18 (model shape depends on a directory listing)
    return 4


def _pose_dims(config):
    """nets/base.py:69-100 with convert_to_6d: jaw 6 + body 90 + hands 180 + expression."""
    scale = 2 if config.Data.pose.convert_to_6d else 1
    exp_dim = int(os.environ.get("expression_dim", "100"))
    each = [3 * scale, 45 * scale, 90 * scale, exp_dim]
    return sum(each), each


def _device(args):
    gpu = getattr(args, "gpu", 0)
    if isinstance(gpu, str) and not gpu.startswith("cuda"):
        raise E._lib.PtError("probtalk_b200 needs a CUDA device (args.gpu=%r); there is no CPU path" % (gpu,))
    return torch.device("cuda", gpu) if isinstance(gpu, int) else torch.device(gpu)


class _Timer:
    """The reference brackets every infer_* call with torch.cuda.synchronize() + time.time() and returns the seconds
    (predictor.py:423-443, refiner.py:541-554).  ``blocking=True`` keeps that contract (CUDA events instead of
    wall-clock); ``blocking=False`` (wrapper attribute ``sync_timing = False``) returns 0.0 seconds and never
    synchronises, so a caller can overlap copies of the next batch with this batch's kernels."""

    def __init__(self, blocking=True):
        self.blocking, self.seconds = blocking, 0.0

    def __enter__(self):
        if self.blocking:
            self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            self.e0.record()
        return self

    def __exit__(self, *a):
        if self.blocking:
            self.e1.record()
            torch.cuda.synchronize()
            self.seconds = self.e0.elapsed_time(self.e1) * 1e-3


class PQLayer(M.ProductQuantizationParams):
    """ProductQuantization buffers + the two calls the scripts make on ``VQ.vq_layer``."""

    def _weights(self):
        return self.__dict__["_owner"]().weights()

    def quantize_all(self, encoding_indices):
        """[..., G] indices -> [..., G*D] (vqvae_modules.py:557-564)."""
        W = self._weights()
        idx = encoding_indices.to(W.device)
        if idx.dtype not in (torch.int32, torch.int64):
            idx = idx.long()
        lead = idx.shape[:-1]
        flat = idx.reshape(1, -1, idx.shape[-1]).contiguous()
        q, bad = ops.pq_gather(flat, W.codebooks, channels_last=True)
        if int(bad) != 0:
            raise IndexError("code index out of range in quantize_all")
        return q.reshape(*lead, -1)

    def train_forward(self, z, commitment_cost=0.25, decay=0.99, epsilon=1e-5):
        """ProductQuantization.forward in TRAINING mode (vqvae_modules.py:470-545, share_code = False): nearest-code search,
        EMA codebook update + dead-code reset on this module's own buffers (pt_pq_ema_update), commitment loss.  Returns
        QuantizationOutput(quantized [B, G*D, W] -- the straight-through VALUE, no autograd graph --, loss)."""
        dev = self.__dict__["_owner"]()._device
        z = z.to(dev, torch.float32)
        G = self.num_chunks
        cb = torch.stack([getattr(self, "embeddings_%d" % g).to(dev, torch.float32) for g in range(G)]).contiguous()
        dw = torch.stack([self.ema_dw[g].hidden.to(dev, torch.float32) for g in range(G)]).contiguous()
        cs = torch.stack([self.ema_cluster_size[g].hidden.to(dev, torch.float32) for g in range(G)]).contiguous()
        idx = ops.pq_search(z, cb)
        q, _ = ops.pq_gather(idx, cb)                                   # quantized with the codebooks BEFORE the update
        self.__dict__["_ema_step"] = self.__dict__.get("_ema_step", 0) + 1
        mse, _ = ops.pq_ema_update(z, idx, cb, dw, cs, decay, epsilon, self.__dict__["_ema_step"])
        with torch.no_grad():
            for g in range(G):
                getattr(self, "embeddings_%d" % g).copy_(cb[g])
                self.ema_dw[g].hidden.copy_(dw[g])
                self.ema_cluster_size[g].hidden.copy_(cs[g])
        owner = self.__dict__["_owner"]()
        owner._weights, owner._plans = None, E.PlanCache()              # prepared weights / plans hold the old codebooks
        return QuantizationOutput(q, commitment_cost * mse.to(torch.float32)[0])

    def forward(self, z, r=0.0, **kw):
        """ProductQuantization.forward in eval mode (vqvae_modules.py:470-492)."""
        W = self._weights()
        z = z.to(W.device, torch.float32)
        idx = ops.pq_search(z, W.codebooks)
        q, _ = ops.pq_gather(idx, W.codebooks)
        return QuantizationOutput(q, idx)


class VectorQuantizerEMA(torch.nn.Module):
    """Inference surface of the single-codebook quantizer of the TalkSHOW-style VQ-VAEs (vqvae_modules.py:279-377, built
    with embedding_dim = 64, num_embeddings = config.Model.code_num by nets/smplx_body_vq.py:48-53): the ``embeddings``
    buffer with the reference's name / shape, ``get_code_indices`` (:350-358), ``quantize`` (:360-362) and the eval
    ``forward`` (indices -> quantized, [B, D, T] in and out).  The search is pt_pq_search with G = 1."""

    def __init__(self, embedding_dim, num_embeddings, commitment_cost=0.25, decay=0.99, epsilon=1e-5):
        super().__init__()
        self.embedding_dim, self.num_embeddings = embedding_dim, num_embeddings
        self.commitment_cost, self.decay, self.epsilon = commitment_cost, decay, epsilon
        self.register_buffer("embeddings", torch.zeros(num_embeddings, embedding_dim))
        self.register_buffer("ema_dw", torch.zeros(num_embeddings, embedding_dim))
        self.register_buffer("ema_cluster_size", torch.zeros(num_embeddings))

    def get_code_indices(self, flat_x):
        """flat_x [N, D] (CUDA fp32) -> indices [N] int64."""
        cb = self.embeddings.to(flat_x.device, torch.float32).unsqueeze(0).contiguous()
        z = flat_x.to(torch.float32).contiguous().t().unsqueeze(0)            # [1, D, N] view, channels-last strides
        return ops.pq_search(z, cb).reshape(-1)

    def quantize(self, encoding_indices):
        idx = encoding_indices.reshape(1, -1, 1)
        q, bad = ops.pq_gather(idx.to(self.embeddings.device), self.embeddings.float().unsqueeze(0).contiguous(),
                               channels_last=True)
        if int(bad) != 0:
            raise IndexError("code index out of range in quantize")
        return q.reshape(*encoding_indices.shape, self.embedding_dim)

    def forward(self, x):
        """Eval path of vqvae_modules.py:305-343: x [B, D, T] -> (quantized [B, D, T], loss = 0)."""
        B, D, T = x.shape
        flat = x.permute(0, 2, 1).reshape(-1, D)
        idx = self.get_code_indices(flat)
        q = self.quantize(idx).view(B, T, D).permute(0, 2, 1).contiguous()
        return q, torch.zeros((), device=x.device)


class VQVAE(M.VQVAE_SC):
    """VQVAE_SC with the reference's call surface (vqvae_1d_sc.py:277-324) on top of the CUDA plans."""

    def __init__(self, *a, device="cuda", precision="bf16", **k):
        super().__init__(*a, **k)
        self._precision = precision
        import weakref
        pq = PQLayer(self.embedding_dim, self.num_embeddings, self.groups)
        pq.__dict__["_owner"] = weakref.ref(self)
        self.vq_layer = pq
        self._device = torch.device(device)
        self._weights = None
        self._plans = E.PlanCache()

    def enable_facade(self):
        pass

    def load_state_dict(self, sd, strict=True):
        out = super().load_state_dict(_strip(sd), strict=strict)
        self._weights, self._plans = None, E.PlanCache()
        return out

    def weights(self):
        if self._weights is None:
            self._weights = E.VQVAEWeights(self.state_dict(), self._device, self.groups, self._precision)
        return self._weights

    def _plan(self, kind, *key):
        def build():
            W = self.weights()
            if kind == "enc":
                return E.EncoderPlan(W, *key)
            if kind == "dec":
                return E.DecoderPlan(W, *key)
            return E.DecoderPlan(W, *key, from_latents=True)
        return self._plans.get((kind,) + key, build)

    def encoder(self, x):
        B, _, T = x.shape
        plan = self._plan("enc", B, T)
        plan.run(x.to(self._device, torch.float32))
        return plan.z_channel_first().contiguous(), {}

    def encode(self, gt_poses, r=0.0, **kw):
        B, _, T = gt_poses.shape
        plan = self._plan("enc", B, T)
        codes = plan.run(gt_poses.to(self._device, torch.float32)).clone()
        z = plan.z_channel_first().contiguous()
        q, _ = ops.pq_gather(codes, self.weights().codebooks)
        return z, QuantizationOutput(q, codes), {}

    def decode(self, e, enc_feats=None):
        B, _, Tq = e.shape
        plan = self._plan("dec_lat", B, Tq)
        return plan.run(latents=e.to(self._device, torch.float32)).clone()

    def decode_codes(self, codes):
        B, Tq, _ = codes.shape
        plan = self._plan("dec", B, Tq)
        return plan.run(codes=codes.to(self._device)).clone()

    def forward(self, gt_poses, result_form="pose", r=0.0, **kw):
        z, qo, _ = self.encode(gt_poses)
        pred = self.decode_codes(qo.loss)
        if result_form == "full":
            return z, qo, pred
        if result_form == "part":
            return qo.loss, pred
        return pred

    __call__ = forward


class _WrapperBase:
    def __init__(self, args, config):
        self.args, self.config = args, config
        self.device = _device(args)
        # arithmetic of the dense ops: "bf16" (throughput) or "bf16x3" (split-bf16, fp32-grade).
        # Not part of the reference's interface: an optional args attribute, else the PROBTALK_B200_PRECISION variable.
        self.precision = getattr(args, "precision", None) or os.environ.get("PROBTALK_B200_PRECISION", "bf16")
        E._check_precision(self.precision)
        self.full_dim, self.each_dim = _pose_dims(config)
        self.num_classes = _n_classes(config)
        self.global_step = 0
        self.generator_optimizer = None
        self.sync_timing = True       # False: infer_* calls do not synchronise and report 0.0 seconds

    def _make_vq(self, in_dim):
        m = self.config.Model
        vq = VQVAE(in_dim, m.groups * 128, m.code_num, 1024, 2, m.groups, device=self.device, precision=self.precision)
        vq.enable_facade()
        path = getattr(m, "vq_path", None)
        if path and os.path.exists(path):
            ckpt = torch.load(path, map_location="cpu")
            vq.load_state_dict(ckpt["generator"]["VQ"])
        return vq

    def _labels(self, id, B):
        t = torch.as_tensor(id).reshape(-1).to(torch.int32)
        if t.numel() == 1:
            t = t.repeat(B)
        return t.to(self.device)

    def to_parallel(self):
        pass


class s2g_body_vqt(_WrapperBase):
    """vq_teacher.TrainWrapper (nets/inpainting/vq_teacher.py:23-285), inference surface."""

    def __init__(self, args, config):
        super().__init__(args, config)
        self.type = config.Model.vq_type
        assert self.type == "fbhe", "only the holistic 'fbhe' PQ-VAE of config/vq.json is built"
        self.VQ = VQVAE(self.full_dim, config.Model.groups * 128, config.Model.code_num, 1024, 2, config.Model.groups,
                        device=self.device, precision=self.precision)
        self.VQ.enable_facade()
        self.generator = self.VQ

    def state_dict(self):
        return {"VQ": self.VQ.state_dict(), "generator_optim": None}

    def load_state_dict(self, state_dict):
        if "VQ" in state_dict:
            self.VQ.load_state_dict(state_dict["VQ"])

    def infer_on_audio(self, gt_poses, aud=None, text=None, id=None, B=1, **kw):
        """encode -> product-quantise -> decode (vq_teacher.py:255-275)."""
        assert self.args.infer, "train mode"
        with _Timer(self.sync_timing) as t:
            pred = self.VQ(gt_poses=gt_poses)
        return pred, t.seconds

    def get_codes(self, gt_poses):
        return self.VQ.encode(gt_poses)[1].loss


class s2g_body_predictor(_WrapperBase):
    """predictor.TrainWrapper (nets/inpainting/predictor.py:27-685), inference surface."""

    def __init__(self, args, config):
        super().__init__(args, config)
        m = config.Model
        self.type = m.vq_type
        assert self.type == "fbhe" and m.maskgit and m.transformer and m.identity and not m.knn, \
            "only the configuration of config/transformer.json is built"
        self.audio, self.text = m.p_audio, m.p_text
        assert self.audio is not None and self.text is None
        self.AudEnc = M.ConditionEncoder(768, 256)
        self.TextEnc = M.ConditionEncoder(768, 256)
        self.Predictor = M.PredictorNet(self.full_dim, m.code_num, 512, self.num_classes, m.groups, m.maskgit_T)
        self.VQ = self._make_vq(self.full_dim)
        self.in_dim = self.full_dim
        self._weights, self._plans = None, E.PlanCache()
        self.noise_generator = None

    def state_dict(self):
        return {"AudEnc": self.AudEnc.state_dict(), "Predictor": self.Predictor.state_dict(),
                "TextEnc": self.TextEnc.state_dict(), "generator_optim": None}

    def load_state_dict(self, state_dict):
        sd = {k: _strip(v) for k, v in state_dict.items() if v is not None and k != "generator_optim"}
        for name in ("AudEnc", "Predictor", "TextEnc"):
            if name in sd:
                getattr(self, name).load_state_dict(sd[name])
        self._weights, self._plans = None, E.PlanCache()

    def _plan(self, B, T):
        if self._weights is None:
            self._weights = E.PredictorWeights({"AudEnc": self.AudEnc.state_dict(), "Predictor": self.Predictor.state_dict()},
                                               self.device, self.precision)
        return self._plans.get((B, T), lambda: E.PredictorPlan(self._weights, self.VQ.weights(), B, T, self.Predictor.T))

    def _infer(self, aud, id, B, gt_poses, mask, noise=None):
        assert self.args.infer, "train mode"
        dev = self.device
        aud = aud.to(dev, torch.float32)
        if aud.shape[0] != B:
            aud = aud.repeat(B, 1, 1)                         # predictor.py:415
        T = aud.shape[2]
        plan = self._plan(B, T)
        with _Timer(self.sync_timing) as t:
            poses, codes = plan.run(aud, gt_poses.to(dev, torch.float32), mask.to(dev, torch.float32),
                                    self._labels(id, B), noise=noise, generator=self.noise_generator)
        return poses.clone(), codes.to(torch.int32).clone(), t.seconds

    def infer_on_batch(self, aud, text=None, id=0, B=1, gt_poses=None, mask=None, noise=None, **kw):
        poses, _, sec = self._infer(aud, id, B, gt_poses, mask, noise)
        return poses, sec

    def infer_to_code(self, aud, text=None, id=0, B=1, gt_poses=None, mask=None, noise=None, **kw):
        return self._infer(aud, id, B, gt_poses, mask, noise)


class s2g_body_refiner(_WrapperBase):
    """refiner.TrainWrapper (nets/inpainting/refiner.py:32-663), inference surface."""

    def __init__(self, args, config):
        super().__init__(args, config)
        m = config.Model
        self.type, self.vqtype = m.hf_type, m.hf_vq_type
        assert self.type == "fbhe" and m.two_stage and m.motion_context and m.identity, \
            "only the configuration of config/transformer.json is built"
        self.two_stage = True
        self.RTransformer = M.RTransformer(self.full_dim, 768, 512, 6, self.num_classes)
        self.AudEnc = torch.nn.Identity()
        self.VQ = self._make_vq(self.full_dim)      # kept for checkpoint/API parity; inference never uses it (refiner.py:84-90)
        self.in_dim = self.vq_dim = self.full_dim
        self._weights, self._plans = None, E.PlanCache()

    def state_dict(self):
        return {"HFTransformer": self.RTransformer.state_dict(), "AudEnc": self.AudEnc.state_dict(), "generator_optim": None}

    def load_state_dict(self, state_dict):
        if state_dict.get("HFTransformer") is not None:
            self.RTransformer.load_state_dict(_strip(state_dict["HFTransformer"]))
        self._weights, self._plans = None, E.PlanCache()

    def _plan(self, B, T, Tp):
        if self._weights is None:
            self._weights = E.RefinerWeights({"HFTransformer": self.RTransformer.state_dict()}, self.device, self.precision)
        return self._plans.get((B, T, Tp), lambda: E.RefinerPlan(self._weights, B, T, Tp))

    def infer_on_batch(self, aud, B=1, id=0, gt_poses=None, mask=None, pred_poses=None, **kw):
        assert self.args.infer, "train mode"
        dev = self.device
        aud = aud.to(dev, torch.float32)
        if aud.shape[0] != B:
            aud = aud.repeat(B, 1, 1)
        T = aud.shape[2]
        plan = self._plan(B, T, pred_poses.shape[2])
        with _Timer(self.sync_timing) as t:
            out = plan.run(aud, gt_poses.to(dev, torch.float32), mask.to(dev, torch.float32),
                           pred_poses.to(dev, torch.float32), self._labels(id, B))
        return out.clone(), t.seconds
