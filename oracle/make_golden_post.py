"""Generate tests/golden/postprocess_vectors.pt by running the UNMODIFIED reference post-processing on CPU.

TEST INFRASTRUCTURE; run in the authoring container only (needs /root/reference):
    python -m oracle.make_golden_post
Stores the inputs (small) and the reference's outputs for scripts/demo.py:183-205's sequence
(denormalize 'fbhe' -> rotation_6d_to_matrix -> matrix_to_axis_angle -> cat(expression) -> part2full), sitting and
standing variants, plus c_index_6d and the `lower_pose` row so that the restated constants are pinned too.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH   # noqa: E402  (installs the import shims for /root/reference)


def main():
    RH.load_reference()
    from data_utils import lower_body as LB
    from data_utils.rotation_conversion import matrix_to_axis_angle, rotation_6d_to_matrix
    from nets.utils import denormalize
    g = torch.Generator().manual_seed(2024)
    B, T = 3, 37
    pred = torch.randn(B, 376, T, generator=g)
    # a few hard rows: tiny rotations (small-angle branch), near-pi rotations, a degenerate 6-D vector
    pred[0, 0:6, 0] = torch.tensor([1.0, 0.0, 0.0, 0.0, 1.0, 0.0])
    pred[0, 6:12, 0] = torch.tensor([1.0, 1e-7, 0.0, -1e-7, 1.0, 0.0])
    pred[0, 12:18, 0] = torch.tensor([-1.0, 0.0, 0.0, 0.0, -1.0, 1e-3])
    pred[0, 18:24, 0] = torch.tensor([0.0, 0.0, 0.0, 1.0, 0.0, 0.0])
    mean430 = torch.randn(1, 1, 430, generator=g) * 0.3
    std430 = torch.rand(1, 1, 430, generator=g) + 0.5
    c_index = LB.c_index_6d
    out = {"torch_version": str(torch.__version__), "pred": pred, "mean430": mean430, "std430": std430,
           "c_index_6d": torch.from_numpy(np.asarray(c_index)).long(), "lower_pose": LB.lower_pose.clone()}
    for name, norm in (("denorm", True), ("raw", False)):
        for stand in (False, True):
            rows = []
            for b in range(B):
                p = pred[b:b + 1]
                if norm:
                    p = denormalize(p, mean430, std430, "fbhe", c_index)
                p = p.squeeze(0).transpose(0, 1)
                face, body = p[:, -100:], p[:, :-100]
                body = matrix_to_axis_angle(rotation_6d_to_matrix(body.reshape(body.shape[0], -1, 6))).reshape(body.shape[0], -1)
                rows.append(LB.part2full(torch.cat([body, face], dim=-1), stand))
            out["%s_%s" % (name, "stand" if stand else "sit")] = torch.stack(rows)
    path = os.path.join(ROOT, "tests", "golden", "postprocess_vectors.pt")
    torch.save(out, path)
    print("wrote", path, {k: tuple(v.shape) for k, v in out.items() if torch.is_tensor(v)})


if __name__ == "__main__":
    main()
