"""Soak run (developer tool): many repetitions of the step's kernels and of the wav2vec2 path; any dead-lock ends in the
kernels' spin-limit trap ('unspecified launch failure').  [SOAK=10] python tools/soak.py"""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from probtalk_b200 import ops, engine as E, synthetic as S, audio as A
t0 = time.time()
R = int(os.environ.get("SOAK", "1"))            # repetition multiplier
for (B, Nq, Nk, H, n) in ((1024, 88, 88, 8, 600), (1024, 88, 22, 8, 600), (1024, 180, 180, 8, 200), (8, 2999, 2999, 12, 60), (3, 700, 700, 12, 200)):
    q = torch.randn(B * Nq, H * 64, device="cuda").bfloat16(); kv = torch.randn(B * Nk, 2 * H * 64, device="cuda").bfloat16()
    for _ in range(n * R):
        ops.attention(q, kv[:, :H * 64], kv[:, H * 64:], B, Nq, Nk, heads=H)
    torch.cuda.synchronize(); print("attention", B, Nq, Nk, "x", n, "ok %.1fs" % (time.time() - t0), flush=True)
vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
VW, PW, RW = E.VQVAEWeights(vq, "cuda"), E.PredictorWeights(pst, "cuda"), E.RefinerWeights(rst, "cuda")
for B, n in ((1024, 60), (64, 300), (7, 300)):
    pred = E.PredictorPlan(PW, VW, B, 180); ref = E.RefinerPlan(RW, B, 180, 180)
    aud = torch.randn(B, 768, 180, device="cuda"); gt = torch.zeros(B, 376, 180, device="cuda"); mask = torch.zeros(B, 1, 180, device="cuda")
    ids = (torch.arange(B, device="cuda") % 4).int()
    for _ in range(n * R):
        poses, _ = pred.run(aud, gt, mask, ids); ref.run(aud, gt, mask, poses, ids)
    torch.cuda.synchronize(); print("window B =", B, "x", n, "ok %.1fs" % (time.time() - t0), flush=True)
    del pred, ref; torch.cuda.empty_cache()
W = A.Wav2VecWeights(S.make_wav2vec2(11).state_dict(), "cuda")
for B, L, n in ((8, 960000, 40), (1, 960000, 200), (2, 100003, 200)):
    plan = A.Wav2VecPlan(W, B, L); wave = S.synthetic_waveform(1, L, 5).cuda().repeat(B, 1).contiguous()
    for _ in range(n * R):
        plan.run(wave)
    torch.cuda.synchronize(); print("wav2vec2 B =", B, "L =", L, "x", n, "ok %.1fs" % (time.time() - t0), flush=True)
    del plan; torch.cuda.empty_cache()
print("soak ok")
