"""BASELINE config 5 on one GPU: U utterances x 1800 frames (60 s) through the long-form driver
(12 predictor + 13 refiner windows each, scripts/demo.py generalised).  Prints utterances/s and frames/s."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from probtalk_b200 import driver as D, synthetic as S   # noqa: E402

U = int(sys.argv[1]) if len(sys.argv) > 1 else 64
t = int(sys.argv[2]) if len(sys.argv) > 2 else 1800
pipe = D.HolisticPipeline(S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5), "cuda")
aud = S.synthetic_audio(U, t, 1005).cuda()
labels = (torch.arange(U) % 4).int().cuda()
for _ in range(2):
    out, codes = pipe.windows_demo(aud, labels)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
out, codes = pipe.windows_demo(aud, labels)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("config 5: %d utterances x %d frames: %.1f ms -> %.1f utterances/s, %.0f frames/s (%.1f TFLOP/s at 0.65 TFLOP/utterance)"
      % (U, t, ms, U / ms * 1e3, U * t / ms * 1e3, U * 0.65 / ms * 1e3))
