"""Generate tests/golden/reference_vectors.pt by running the UNMODIFIED reference on CPU.

TEST INFRASTRUCTURE; run in the authoring container only (needs /root/reference):
    python -m oracle.make_golden
Inputs and weights are re-creatable from seeds (probtalk_b200.synthetic), so only the reference's
OUTPUTS are stored.  tests/test_golden.py replays the oracle against them on any machine; the GPU
tests compare the CUDA path with the same vectors.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness as RH                      # noqa: E402
from probtalk_b200 import synthetic as S                  # noqa: E402
from tests.test_oracle_vs_reference import _noise, _patched_sampler   # noqa: E402

SEEDS = dict(vq=0, pred=1, ref=5)


def main():
    vq, pst, rst = S.make_vq_state(SEEDS["vq"]), S.make_predictor_state(SEEDS["pred"]), S.make_refiner_state(SEEDS["ref"])
    out = {"torch_version": str(torch.__version__), "threads": torch.get_num_threads(), "seeds": SEEDS}
    ref_vq = RH.build_vqt(vq)
    ref_pred = RH.build_predictor(vq, pst)
    ref_ref = RH.build_refiner(vq, rst)
    with torch.no_grad():
        # config 1: one 300-frame clip through encode -> PQ -> decode (vq_teacher.infer_on_audio)
        x = S.synthetic_poses(1, 300, 1001)
        z, _ = ref_vq.VQ.encoder(x)
        qo = ref_vq.VQ.vq_layer(z)
        rec = ref_vq.VQ.decode(qo.quantized, None)
        out["cfg1"] = dict(poses_seed=1001, z=z.clone(), codes=qo.loss.to(torch.int16), recon=rec.clone())
        # config 2 (reduced batch): PQ search layer alone, two input scales
        g = torch.Generator().manual_seed(1002)
        cases = []
        for scale in (1.0, 0.1):
            zz = torch.randn(32, 512, 22, generator=g) * scale
            cases.append(dict(scale=scale, codes=ref_vq.VQ.vq_layer(zz).loss.to(torch.int16)))
        out["cfg2"] = dict(seed=1002, B=32, W=22, cases=cases)
        # config 3: one tokens_to_logits pass
        g = torch.Generator().manual_seed(1005)
        tokens = torch.randint(0, 129, (3, 22, 4), generator=g, dtype=torch.int32)
        cond = torch.randn(3, 768, 22, generator=g)
        label = torch.tensor([0, 3, 1])
        out["cfg3"] = dict(seed=1005, label=label, logits=ref_pred.Predictor.decoder.tokens_to_logits(tokens, label, cond).clone())
        # config 4: predictor + refiner on one 180-frame window, 30 known frames, injected sampler noise
        B, T = 2, 180
        aud1 = S.synthetic_audio(1, T, 1008)
        gt = S.synthetic_poses(B, T, 1009)
        mask = torch.zeros(B, 1, T)
        mask[:, :, :30] = 1
        u, gum = _noise(8, B, 88, 1010)
        restore = _patched_sampler(u, gum)
        try:
            poses, codes, _ = ref_pred.infer_to_code(aud=aud1, text=aud1, id=1, B=B, gt_poses=gt, mask=mask)
        finally:
            restore()
        refined, _ = ref_ref.infer_on_batch(aud=aud1, B=B, id=torch.tensor([1]), gt_poses=gt, mask=mask, pred_poses=poses)
        out["cfg4"] = dict(aud_seed=1008, gt_seed=1009, noise_seed=1010, B=B, T=T, id=1, codes=codes.to(torch.int16),
                           poses=poses.clone(), refined=refined.clone())
    path = os.path.join(ROOT, "tests", "golden", "reference_vectors.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
