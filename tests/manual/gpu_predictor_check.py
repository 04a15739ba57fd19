"""Developer diagnostic: stage-by-stage error of the CUDA predictor / refiner vs the oracle."""
import os
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import probtalk_oracle as O      # noqa: E402
from probtalk_b200 import engine as E, synthetic as S   # noqa: E402


def err(name, got, ref):
    got, ref = got.float().cpu(), ref.float()
    print("%-28s ref rms %.4f  max err %.5f  rms err %.6f" % (name, ref.pow(2).mean().sqrt(), (got - ref).abs().max(), (got - ref).pow(2).mean().sqrt()))


def noise(T, B, N, seed):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(T, B, N, generator=g)
    gum = -torch.log(-torch.log(torch.rand(T, B, N, generator=g).clamp_(1e-10, 1 - 1e-7)))
    return u, gum


def main():
    vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
    B, T = 4, 180
    aud = S.synthetic_audio(B, T, 1008)
    gt = S.synthetic_poses(B, T, 1009)
    mask = torch.zeros(B, 1, T); mask[:, :, :30] = 1
    labels = torch.tensor([1, 0, 3, 2])
    VW = E.VQVAEWeights(vq, "cuda")
    PW = E.PredictorWeights(pst, "cuda")
    plan = E.PredictorPlan(PW, VW, B, T)
    plan.aud.copy_(aud); plan.gt.copy_(gt); plan.mask.copy_(mask); plan.labels.copy_(labels.int())
    plan.run_condition()
    torch.cuda.synchronize()
    input_poses, cond_ref = O.predictor_condition(aud, gt, mask, pst)
    err("condition [mm|audio]", plan.cond.valid().transpose(1, 2), cond_ref)
    err("  audio part", plan.cond.valid()[..., 512:].transpose(1, 2), cond_ref[:, 512:])
    # teacher-forced logits
    dec = O._sub(pst["Predictor"], "decoder")
    g = torch.Generator().manual_seed(3)
    tokens = torch.randint(0, 129, (B, plan.Tq, 4), generator=g, dtype=torch.int32)
    plan.cur.copy_(tokens.view(B, -1))
    plan.plan_iter.run()
    torch.cuda.synchronize()
    logits_ref = O.tokens_to_logits(tokens, labels, cond_ref, dec)
    err("logits (teacher-forced)", plan.logits.view(B, plan.Tq, 4, 128), logits_ref)
    # maskgit step on the oracle logits
    u, gum = noise(8, B, plan.N, 1010)
    L = E._lib.lib()
    cur = torch.full((B, plan.N), 128, dtype=torch.int32)
    lg = logits_ref.reshape(B, plan.N, 128).contiguous()
    for t in (0, 3, 7):
        cur_t = cur.clone()
        if t > 0:
            cur_t[:, ::3] = torch.randint(0, 128, cur_t[:, ::3].shape, dtype=torch.int32)
        ref = O.maskgit_step(lg, cur_t, t, 8, u[t], gum[t])
        out = torch.empty_like(cur_t).cuda()
        E.check(L.pt_maskgit_step(lg.cuda().data_ptr(), cur_t.cuda().data_ptr(), u[t].cuda().data_ptr(), gum[t].cuda().data_ptr(),
                                  out.data_ptr(), B, plan.N, 128, 128, E.maskgit_base_len(plan.N, t, 8),
                                  4.5 * (1 - (t + 1) / 8), 0))
        torch.cuda.synchronize()
        print("maskgit step %d: %d / %d tokens differ; masked ref %d got %d" % (t, int((out.cpu() != ref).sum()), ref.numel(), int((ref == 128).sum()), int((out.cpu() == 128).sum())))
    # full run with injected noise
    poses, codes = plan.run(aud.cuda(), gt.cuda(), mask.cuda(), labels.int().cuda(), noise=(u.cuda(), gum.cuda()))
    torch.cuda.synchronize()
    poses_ref, codes_ref = O.predictor_infer_on_batch(aud, gt, mask, labels, pst, vq, u, gum)
    print("codes: %d / %d differ (free-running sampler, bf16 logits)" % (int((codes.cpu() != codes_ref).sum()), codes_ref.numel()))
    err("poses (free-running)", poses, poses_ref)
    poses_tf = O.predictor_decode(codes.cpu().long(), input_poses, mask, vq, T)
    err("poses (given CUDA codes)", poses, poses_tf)
    # refiner
    RW = E.RefinerWeights(rst, "cuda")
    rp = E.RefinerPlan(RW, B, T, T)
    out = rp.run(aud.cuda(), gt.cuda(), mask.cuda(), poses_tf.cuda(), labels.int().cuda())
    torch.cuda.synchronize()
    out_ref = torch.cat([O.refiner_infer_on_batch(aud[i:i + 1], gt[i:i + 1], mask[i:i + 1], labels[i:i + 1], poses_tf[i:i + 1], rst["HFTransformer"]) for i in range(B)])
    err("refiner output", out, out_ref)
    err("  unknown frames only", out[:, :, 30:], out_ref[:, :, 30:])
    # timing
    for Bt in (16, 64, 256):
        pl = E.PredictorPlan(PW, VW, Bt, 180)
        rf = E.RefinerPlan(RW, Bt, 180, 180)
        a = S.synthetic_audio(Bt, 180, 1).cuda(); gg = S.synthetic_poses(Bt, 180, 2).cuda(); mm = torch.zeros(Bt, 1, 180).cuda()
        lab = torch.zeros(Bt, dtype=torch.int32).cuda()
        for _ in range(2):
            p_, _c = pl.run(a, gg, mm, lab); rf.run(a, gg, mm, p_, lab)
        torch.cuda.synchronize()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        n = 3
        t0 = time.time()
        e0.record()
        for _ in range(n):
            p_, _c = pl.run(a, gg, mm, lab)
        e1.record()
        for _ in range(n):
            rf.run(a, gg, mm, p_, lab)
        e2.record()
        torch.cuda.synchronize()
        tp, tr = e0.elapsed_time(e1) / n, e1.elapsed_time(e2) / n
        print("B=%d: predictor %.2f ms (%.0f TFLOP/s), refiner %.2f ms (%.0f TFLOP/s) -> %.0f clips/s; launches %d + %d" % (
            Bt, tp, Bt * 41.71 / tp, tr, Bt * 11.32 / tr, Bt / (tp + tr) * 1e3, pl.launches, rf.launches))


if __name__ == "__main__":
    main()
