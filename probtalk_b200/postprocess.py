"""Pose post-processing: normalised 6-D network output -> SMPL-X axis-angle parameter rows.

Mirrors the tail of the reference's demo / evaluation scripts (scripts/demo.py:183-205, scripts/test_holisticbody.py):
``denormalize(pred, mean, std, 'fbhe', c_index)`` -> ``matrix_to_axis_angle(rotation_6d_to_matrix(.))`` per joint ->
``cat(expression)`` -> ``part2full``.  All arithmetic happens in one CUDA kernel (``pt_pose_postprocess``); this module
only prepares the small constant tables.  CUDA tensors only, no CPU fallback.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

EXP_DIM = 100          # data_utils/consts.py: smplx_hyperparams['expression_dim']

# data_utils/lower_body.py:4-9 (`lower_pose`): the fixed lower-body joints part2full splices in (sitting pose)
LOWER_POSE = (0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 3.0747, -0.0158, -0.0152, -1.1826512813568115, 0.23866955935955048,
              0.15146760642528534, -1.2604516744613647, -0.3160211145877838, -0.1603458970785141, 1.1654603481292725,
              0.0, 0.0, 1.2521806955337524, 0.041598282754421234, -0.06312154978513718, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0)
# part2full(stand=True) (lower_body.py:67-71): zeros except the global orientation
LOWER_POSE_STAND = tuple([0.0] * 6 + [3.0747, -0.0158, -0.0152] + [0.0] * 18)


def c_index_6d() -> np.ndarray:
    """Channels of the 330-wide 6-D pose vector the models are trained on (lower_body.py:40-62): every 3-D index
    except the fixed ones, two 6-D channels per kept 3-D index."""
    fixed = set(list(range(0, 18)) + list(range(21, 27)) + list(range(30, 36)))
    keep = [i for i in range(165) if i not in fixed]
    return np.asarray([c for i in keep for c in (2 * i, 2 * i + 1)])


def select_fbhe(stat: torch.Tensor, c_index) -> torch.Tensor:
    """``tofbhe`` of nets/utils.py:13-25 applied to a per-channel statistic: jaw channels, ``c_index``, expression."""
    stat = stat.reshape(-1)
    c_index = torch.as_tensor(np.asarray(c_index), dtype=torch.long, device=stat.device)
    jaw_dim = 3 if c_index.numel() == 129 else 6
    return torch.cat([stat[:jaw_dim], stat[c_index], stat[-EXP_DIM:]])


_LOWER_CACHE = {}


def _lower_table(device, stand):
    key = (str(device), bool(stand))
    if key not in _LOWER_CACHE:
        _LOWER_CACHE[key] = torch.tensor(LOWER_POSE_STAND if stand else LOWER_POSE, dtype=torch.float32, device=device)
    return _LOWER_CACHE[key]


def poses_to_smplx(pred: torch.Tensor, mean: torch.Tensor = None, std: torch.Tensor = None, c_index=None,
                   stand: bool = False, n_expr: int = EXP_DIM) -> torch.Tensor:
    """pred [B, 6J + n_expr, T] fp32 CUDA (the models' output) -> [B, T, 3J + n_expr + 27] SMPL-X parameter rows.

    ``mean`` / ``std``: the reference's ``norm_stats`` halves (any shape, 430 channels) with ``c_index``
    (``c_index_6d()``), or already-selected per-channel vectors when ``c_index`` is None; both None = the
    ``Data.pose.normalization = False`` path."""
    if not pred.is_cuda:
        raise _lib.PtError("probtalk_b200.postprocess needs CUDA tensors (no CPU fallback)")
    assert pred.dtype == torch.float32 and pred.dim() == 3
    pred = pred.contiguous()
    B, C_, T = pred.shape
    assert (C_ - n_expr) % 6 == 0 and C_ > n_expr, "channels must be 6 * joints + expression"
    J = (C_ - n_expr) // 6
    m = s = None
    if mean is not None:
        m = (select_fbhe(mean, c_index) if c_index is not None else mean.reshape(-1)).to(pred.device, torch.float32).contiguous()
        s = (select_fbhe(std, c_index) if c_index is not None else std.reshape(-1)).to(pred.device, torch.float32).contiguous()
        assert m.numel() == C_ and s.numel() == C_
    lower = _lower_table(pred.device, stand)
    out = torch.empty(B, T, 3 * J + n_expr + 27, dtype=torch.float32, device=pred.device)
    _lib.check(_lib.lib().pt_pose_postprocess(pred.data_ptr(), None if m is None else m.data_ptr(),
                                              None if s is None else s.data_ptr(), lower.data_ptr(), out.data_ptr(), B, T, J,
                                              n_expr, torch.cuda.current_stream().cuda_stream), "pt_pose_postprocess")
    return out
