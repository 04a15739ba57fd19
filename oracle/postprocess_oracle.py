"""CPU restatement of the reference's pose post-processing (TEST INFRASTRUCTURE -- only tests/ may import this).

Follows scripts/demo.py:183-205 step by step in fp32 torch:
  denormalize            nets/utils.py:27-43 ('fbhe' selection :13-25)
  rotation_6d_to_matrix  data_utils/rotation_conversion.py:512-534
  matrix_to_quaternion   :97-119   (_sqrt_positive_part :86-94, _copysign :68-83)
  quaternion_to_axis_angle :481-509
  part2full              data_utils/lower_body.py:66-83
Pinned against the unmodified reference functions by tests/test_oracle_vs_reference.py and by the golden vectors of
oracle/make_golden_post.py (tests/golden/postprocess_vectors.pt).
"""
import torch
import torch.nn.functional as F


def rotation_6d_to_matrix(d6):
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = F.normalize(a2 - (b1 * a2).sum(-1, keepdim=True) * b1, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def _sqrt_pos(x):
    return torch.where(x > 0, torch.sqrt(torch.clamp(x, min=0)), torch.zeros_like(x))


def _copysign(a, b):
    return torch.where((a < 0) != (b < 0), -a, a)


def matrix_to_axis_angle(m):
    m00, m11, m22 = m[..., 0, 0], m[..., 1, 1], m[..., 2, 2]
    o0 = 0.5 * _sqrt_pos(1 + m00 + m11 + m22)
    x = _copysign(0.5 * _sqrt_pos(1 + m00 - m11 - m22), m[..., 2, 1] - m[..., 1, 2])
    y = _copysign(0.5 * _sqrt_pos(1 - m00 + m11 - m22), m[..., 0, 2] - m[..., 2, 0])
    z = _copysign(0.5 * _sqrt_pos(1 - m00 - m11 + m22), m[..., 1, 0] - m[..., 0, 1])
    v = torch.stack((x, y, z), -1)
    norms = torch.norm(v, p=2, dim=-1, keepdim=True)
    half = torch.atan2(norms, o0.unsqueeze(-1))
    ang = 2 * half
    s = torch.where(ang.abs() < 1e-6, 0.5 - ang * ang / 48, torch.sin(half) / ang)
    return v / s


def part2full(x, lower):
    lp = lower.unsqueeze(0).repeat(x.shape[0], 1)
    return torch.cat([x[:, :3], lp[:, :15], x[:, 3:6], lp[:, 15:21], x[:, 6:9], lp[:, 21:27], x[:, 9:]], dim=1)


def poses_to_smplx(pred, mean, std, lower, n_expr=100):
    """pred [B, 6J + n_expr, T]; mean / std per channel (already in 'fbhe' order) or None -> [B, T, 3J + n_expr + 27]."""
    B, C, T = pred.shape
    if mean is not None:
        pred = pred * std.reshape(1, -1, 1) + mean.reshape(1, -1, 1)
    rows = []
    for b in range(B):
        p = pred[b].transpose(0, 1)                       # [T, C]
        face, body = p[:, C - n_expr:], p[:, :C - n_expr]
        aa = matrix_to_axis_angle(rotation_6d_to_matrix(body.reshape(T, -1, 6))).reshape(T, -1)
        rows.append(part2full(torch.cat([aa, face], dim=-1), lower))
    return torch.stack(rows)
