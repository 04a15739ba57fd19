"""Per-launch breakdown of one bench step (event-bracketed replay): GEMM shapes with achieved TFLOP/s."""
import collections
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import engine as E, synthetic as S   # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
VW, PW, RW = E.VQVAEWeights(vq, "cuda"), E.PredictorWeights(pst, "cuda"), E.RefinerWeights(rst, "cuda")
pred, ref = E.PredictorPlan(PW, VW, B, 180), E.RefinerPlan(RW, B, 180, 180)
aud = S.synthetic_audio(B, 180, 1).cuda(); gt = torch.zeros(B, 376, 180).cuda(); mask = torch.zeros(B, 1, 180).cuda()
lab = torch.zeros(B, dtype=torch.int32).cuda()
for _ in range(2):
    p_, _c = pred.run(aud, gt, mask, lab); ref.run(aud, gt, mask, p_, lab)
torch.cuda.synchronize()
recs = []
plans = pred.plans_in_order() + [ref.plan]
for pl in plans:
    st = torch.cuda.current_stream().cuda_stream
    for fn, args, flops in pl.steps:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); rc = fn(*args, st); e1.record()
        assert rc == 0
        key = fn.__name__
        if key == "pt_gemm_bf16":
            d = args[0]._obj
            key = "gemm M=%d N=%d K=%d taps=%d s=%d out=%s res=%s act=%d ocs1=%d rn=%d" % (d.M, d.n_valid, d.taps * d.C, d.taps, d.s, "f32" if d.out_f32 else "bf16",
                                                                            ("f32" if d.res_f32 else "bf16") if d.residual else "-", d.act, int(d.ocs == 1), d.row_norm)
        elif key == "pt_attention":
            key = "attention Nq=%d Nk=%d" % (args[11], args[12])
        recs.append((key, e0, e1, flops))
torch.cuda.synchronize()
agg = collections.OrderedDict()
for k, e0, e1, fl in recs:
    a = agg.setdefault(k, [0, 0.0, 0.0])
    a[0] += 1; a[1] += e0.elapsed_time(e1); a[2] += fl
tot = sum(a[1] for a in agg.values())
print("total %.2f ms, %d launches" % (tot, len(recs)))
for k, (n, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%6.2f ms %5.1f%%  n=%4d  avg %7.1f us  %7.1f TFLOP/s  %s" % (ms, 100 * ms / tot, n, ms / n * 1e3, fl / ms / 1e9 if ms > 0 else 0, k))
