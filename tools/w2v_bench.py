"""Timing of the wav2vec2 feature path on one B200: per-kernel breakdown of a 60 s utterance, device
time for B = 1 / 8, the call through the facade with a HOST waveform, and the transformers fp32 model on the host cores
for a bounded sample.  Prints one JSON object.

    python tools/w2v_bench.py [--seconds 60] [--cpu-seconds 10]
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60.0)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--batches", default="1,8")
    args = ap.parse_args()
    from probtalk_b200 import audio as A, synthetic as S
    model = S.make_wav2vec2(11)
    W = A.Wav2VecWeights(model.state_dict(), "cuda")
    L = int(args.seconds * 16000)
    out = {"seconds_per_utterance": args.seconds, "samples": L}
    for B in [int(b) for b in args.batches.split(",")]:
        plan = A.Wav2VecPlan(W, B, L)
        wave = S.synthetic_waveform(B, L, 5).cuda()
        for _ in range(3):
            plan.run(wave)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            plan.run(wave)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        rec = {"ms": ms, "audio_seconds_per_s": B * args.seconds / (ms * 1e-3), "frames_50fps": plan.T, "frames_30fps": plan.t_out,
               "launches": plan.launches, "algorithmic_tflop": plan.flops / 1e12, "tflops": plan.flops / (ms * 1e-3) / 1e12}
        # per-kernel breakdown (event pair around every launch: includes launch gaps, so the sum exceeds `ms`)
        records = []
        plan.plan.run_timed(records)
        torch.cuda.synchronize()
        agg = {}
        names = sorted(plan.marks.items(), key=lambda kv: kv[1][0])
        bounds = [(nm, v[0]) for nm, v in names]
        for i, (name, a, b, flops, key) in enumerate(records):
            stage = next((nm for nm, end in bounds if i < end), "interp")
            if stage.startswith("hidden") and stage != "hidden0":
                stage = "encoder_layers/" + name
            elif stage.startswith("conv") and stage != "conv0":
                stage = "conv1-6"
            elif stage == "pos_conv":
                stage = "pos_conv"
            d = agg.setdefault(stage, [0.0, 0.0, 0])
            d[0] += a.elapsed_time(b)
            d[1] += flops
            d[2] += 1
        rec["stages"] = {k: {"ms": v[0], "launches": v[2], "tflops": (v[1] / (v[0] * 1e-3) / 1e12 if v[1] else None)} for k, v in agg.items()}
        # through the facade with a host waveform (pinned), result read back to the host
        if B == 1:
            feats = A.Wav2Vec2Features(W)
            hw = wave[0].cpu().pin_memory()
            feats(hw)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                r = feats(hw).cpu()
            rec["e2e_ms_host_to_host"] = (time.perf_counter() - t0) / 5 * 1e3
        out["B%d" % B] = rec
        del plan
        torch.cuda.empty_cache()
    # the reference's implementation (transformers, fp32) on the host cores, bounded sample
    if args.cpu_seconds > 0:
        Lc = int(args.cpu_seconds * 16000)
        x = S.synthetic_waveform(1, Lc, 6)
        x = (x - x.mean()) / torch.sqrt(x.var(unbiased=False) + 1e-7)
        with torch.no_grad():
            model(x)
            t0 = time.perf_counter()
            model(x)
            dt = time.perf_counter() - t0
        out["cpu_reference"] = {"kind": "transformers Wav2Vec2Model fp32", "threads": torch.get_num_threads(),
                                "sample_audio_seconds": args.cpu_seconds, "s": dt, "audio_seconds_per_s": args.cpu_seconds / dt}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
