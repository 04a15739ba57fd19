"""Orchestrator with the reference's All_In_One_Model call surface (nets/ai1.py:19-89)."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def init_model(model_name, model_path, args, config):
    """nets/ai1.py:10-16: build by registry name, load ``ckpt['generator']``."""
    from . import wrappers
    generator = getattr(wrappers, model_name)(args, config)
    if model_path:
        ckpt = torch.load(model_path, map_location="cpu") if isinstance(model_path, str) else model_path
        generator.load_state_dict(ckpt["generator"])
    return generator


class All_In_One_Model:
    """Chains [predictor, refiner]: each model gets the previous output as ``pred_poses`` (ai1.py:44-60)."""

    def __init__(self, face_name, face_path, body_name, body_path, device, args, config):
        assert face_name is None, "the face model is outside the token-level hot path"
        self.args, self.config = args, config
        if isinstance(body_name, (list, tuple)):
            self.body_model = [init_model(n, p, args, config) for n, p in zip(body_name, body_path)]
        else:
            self.body_model = init_model(body_name, body_path, args, config)
        self.face_model = None

    def __call__(self, forward_type="infer_on_audio", target_frame=0, result_format="one", **kwargs):
        cost = 0.0
        if isinstance(self.body_model, list):
            body, body_list = None, []
            for model in self.body_model:
                body, t0 = getattr(model, forward_type)(**kwargs, target_frame=target_frame, pred_poses=body)
                body_list.append(body)
                cost += t0
        else:
            body, t0 = getattr(self.body_model, forward_type)(**kwargs)
            body_list = [body]
            cost += t0
        holistic = body
        if body.shape[2] < target_frame:
            holistic = F.interpolate(holistic, size=target_frame, align_corners=False, mode="linear")
        if result_format == "one":
            return holistic, cost
        if body_list[0].shape[2] < target_frame:
            body_list[0] = F.interpolate(body_list[0], size=target_frame, align_corners=False, mode="linear")
        return body_list, cost
