"""BASELINE config 4 batch sweep (64-1024 clips on one GPU): clips/s with and without CUDA-graph replay."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import engine as E, synthetic as S   # noqa: E402

vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
VW, PW, RW = E.VQVAEWeights(vq, "cuda"), E.PredictorWeights(pst, "cuda"), E.RefinerWeights(rst, "cuda")
rows = []
for B in [int(x) for x in (sys.argv[1:] or [1, 16, 64, 128, 256, 512, 1024])]:
    aud = S.synthetic_audio(B, 180, 1).cuda(); gt = torch.zeros(B, 376, 180).cuda(); mask = torch.zeros(B, 1, 180).cuda()
    lab = torch.zeros(B, dtype=torch.int32).cuda()
    res = {"clips": B}
    for graph in (False, True):
        pred, ref = E.PredictorPlan(PW, VW, B, 180, use_graph=graph), E.RefinerPlan(RW, B, 180, 180, use_graph=graph)
        for _ in range(3):
            p_, _c = pred.run(aud, gt, mask, lab); ref.run(aud, gt, mask, p_, lab)
        torch.cuda.synchronize()
        n = 10 if B <= 256 else 4
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            p_, _c = pred.run(aud, gt, mask, lab); ref.run(aud, gt, mask, p_, lab)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        res["graph" if graph else "eager"] = {"ms_per_window": round(ms, 3), "clips_per_s": round(B / ms * 1e3, 1),
                                               "tflops": round(B * 53.03 / ms, 1)}
        del pred, ref
    rows.append(res)
    print(json.dumps(res), flush=True)
