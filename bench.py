#!/usr/bin/env python
"""Headline benchmark: motion clips/s through predict + refine + PQ-VAE decode (BASELINE.json config 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]

One "step" = one 180-frame window for B independent clips per GPU (distinct synthetic audio, known-frame
mask all zero = first window of scripts/test_holisticbody.py:239-262):
    s2g_body_predictor.infer_on_batch (AudEnc, context encoder, 8 MaskGIT iterations, codebook gather,
    PQ-VAE decode, interpolate + blend)  ->  s2g_body_refiner.infer_on_batch.
`value`  : clips/s with the inputs already resident in HBM (device timed, max over ranks);
`e2e`    : the same through the public wrapper API (probtalk_b200.nets.All_In_One_Model) from pinned HOST
           buffers, host->device and device->host copies inside the timed region;
`roofline`: the tensor-core GEMM kernel (dominant kernel) against the measured bf16 peak;
`cpu_baseline` / `--impl reference`: the reference algorithm's CPU restatement (oracle/, fp32 PyTorch, all host
           threads) on a bounded sample of the same workload -- the unmodified reference is Python and needs
           /root/reference, which does not exist on the GPU box.
Multi-GPU (torchrun): clips are sharded, weights replicated, the only collective is the final NCCL gather of
the poses to rank 0 (weak scaling: per-GPU batch fixed).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

T_FRAMES = 180
GFLOP_PER_CLIP = 53.03          # predictor 41.71 + refiner 11.32 (algorithmic, fp32 reference)
METRIC = "motion clips/sec (encode+predict+refine+decode)"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sust=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    source="MEASURED_PEAKS.json")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def synthetic_batch(B, seed, device="cpu", pin=False):
    """Config-4 inputs: distinct audio per clip, gt zeros, mask zeros, speaker ids cycling 0..3."""
    g = torch.Generator().manual_seed(seed)
    aud = torch.randn(B, 768, T_FRAMES, generator=g)
    gt = torch.zeros(B, 376, T_FRAMES)
    mask = torch.zeros(B, 1, T_FRAMES)
    ids = (torch.arange(B) % 4).to(torch.int32)
    ts = [aud, gt, mask, ids]
    if pin:
        ts = [t.pin_memory() for t in ts]
    return [t.to(device) for t in ts] if device != "cpu" else ts


# --------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port on the host cores
# --------------------------------------------------------------------------------------------------

def cpu_step(B, states, seed):
    from oracle import probtalk_oracle as O
    vq, pst, rst = states
    aud, gt, mask, ids = synthetic_batch(B, seed)
    g = torch.Generator().manual_seed(seed + 1)
    u = torch.rand(8, B, 88, generator=g)
    gum = -torch.log(-torch.log(torch.rand(8, B, 88, generator=g).clamp_(1e-10, 1 - 1e-7)))
    t0 = time.perf_counter()
    with torch.no_grad():
        poses, _ = O.predictor_infer_on_batch(aud, gt, mask, ids.long(), pst, vq, u, gum)
        out = O.refiner_infer_on_batch(aud, gt, mask, ids.long(), poses, rst["HFTransformer"])
    return time.perf_counter() - t0, out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from probtalk_b200 import synthetic as S
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    states = (S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5))
    Bs = args.ref_batch
    for i in range(args.warmup):
        cpu_step(Bs, states, 100 + i)
    t = 0.0
    for i in range(args.steps):
        dt, _ = cpu_step(Bs, states, 200 + i)
        t += dt
    ms = t / max(args.steps, 1) * 1e3
    v = Bs / (ms * 1e-3)
    sample = "%d of the %d clips per step, oracle port (fp32 PyTorch CPU), %d threads" % (Bs, args.batch, cores)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "clips/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config 4: predictor+refiner+PQ-VAE decode, one 180-frame window", "clips_per_step": Bs,
                       "frames": T_FRAMES, "full_batch_per_gpu": args.batch},
            "cpu_baseline": {"value": v, "unit": "clips/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------

def run_ours(args):
    import torch.distributed as dist
    from probtalk_b200 import engine as E, synthetic as S, _lib
    from probtalk_b200.nets import ai1
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if "PT_BENCH_NCCL_DEBUG" not in os.environ:
            os.environ["NCCL_DEBUG"] = "WARN"          # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        else:
            os.environ["NCCL_DEBUG"] = os.environ["PT_BENCH_NCCL_DEBUG"]
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    B = args.batch
    vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
    VW, PW, RW = E.VQVAEWeights(vq, dev), E.PredictorWeights(pst, dev), E.RefinerWeights(rst, dev)
    pred = E.PredictorPlan(PW, VW, B, T_FRAMES, use_graph=not args.no_graph)
    ref = E.RefinerPlan(RW, B, T_FRAMES, T_FRAMES, use_graph=not args.no_graph)
    aud, gt, mask, ids = synthetic_batch(B, 1004 + rank, dev)
    gather_buf = torch.empty(world * B, 376, T_FRAMES, device=dev) if (world > 1 and rank == 0) else None
    launches_per_step = pred.launches + ref.launches

    def step():
        poses, _ = pred.run(aud, gt, mask, ids)
        out = ref.run(aud, gt, mask, poses, ids)
        if world > 1:      # the path's only collective: gather the final poses on rank 0
            dist.gather(out, list(gather_buf.chunk(world)) if rank == 0 else None, dst=0)
        return out

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    value = world * B / (ms * 1e-3)

    # ---- roofline of the dominant kernel (tensor-core GEMM): per-launch device times from an event-bracketed replay
    gemm_ms, gemm_flops, n_gemm, kinds, gemm_keys = profile_kernels(pred, ref)
    traffic, traffic_hits = gemm_traffic(gemm_keys)

    # ---- end to end through the public wrapper API with host buffers
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, ai1, (vq, pst, rst), B, dev, world, rank)

    cpu = None
    if rank == 0 and not args.no_cpu:
        if world == 1:
            cpu = cpu_baseline(args, (vq, pst, rst))
        else:   # the CPU baseline is a rank-0, N = 1 measurement (torchrun also pins OMP to one thread per rank)
            cpu = {"value": None, "unit": "clips/s", "cores": 0, "kind": "port", "sample": "measured at --gpus 1 only"}
    if rank == 0:
        pk = _peaks()
        ach = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
        line = {"metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": "config 4: s2g_body_predictor + s2g_body_refiner + PQ-VAE decode, one 180-frame window",
                           "clips_per_gpu": B, "global_clips": world * B, "frames": T_FRAMES, "maskgit_T": 8,
                           "parallelism": "dp%d (clips sharded, weights replicated, NCCL gather of poses)" % world,
                           "launch": "eager" if args.no_graph else "one CUDA graph per model per window",
                           "l2": "inputs + activations per step (>1 GB) exceed the 126 MB L2; no flush needed"},
                "clocks": clocks, "gpu_launches": launches_per_step * args.steps,
                "roofline": {"bound": "tensor", "kernel": "gemm_kernel (tcgen05 bf16)", "achieved": ach, "peak": pk["tf_sust"],
                             "unit": "TFLOP/s", "frac": ach / pk["tf_sust"], "traffic": traffic,
                             "traffic_source": ("DRAM read+write bytes per GEMM launch, mean over the %d of %d launches of this "
                                                "step whose shape is in profiles/r01_ncu_gemm_suite.csv (ncu --set full)"
                                                % (traffic_hits, n_gemm)) if traffic else None,
                             "peak_source": pk["source"] + " bf16 sustained (kernel timed inside a long step)",
                             "launches_per_step": n_gemm, "gemm_ms_per_step": gemm_ms,
                             "kernel_share_of_step": kinds},
                "whole_step_tflops": world * B * GFLOP_PER_CLIP / ms,
                "cpu_baseline": cpu, "e2e": e2e}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def profile_kernels(pred, ref):
    """Replay one step with a CUDA event pair around every launch (same stream); returns the GEMM totals
    (ms, algorithmic FLOPs, launches) and each kernel family's share of the step's device time."""
    records = []
    pred.cur.fill_(128)
    for pl in pred.plans_in_order():
        pl.run_timed(records)
    ref.plan.run_timed(records)
    torch.cuda.synchronize()
    totals, counts, flops, keys = {}, {}, 0.0, []
    for name, e0, e1, fl, key in records:
        totals[name] = totals.get(name, 0.0) + e0.elapsed_time(e1)
        counts[name] = counts.get(name, 0) + 1
        if name == "pt_gemm_bf16":
            flops += fl
            keys.append(key)
    tot = sum(totals.values())
    share = {k: round(v / tot, 4) for k, v in sorted(totals.items(), key=lambda kv: -kv[1])}
    return totals.get("pt_gemm_bf16", 0.0), flops, counts.get("pt_gemm_bf16", 0), share, keys


def gemm_traffic(keys):
    """DRAM bytes per GEMM launch (read + write), from the committed `ncu --set full` capture of the six transformer GEMM
    shapes (profiles/r01_ncu_gemm_suite.csv, M = 90112; bytes scale with M for the refiner's M = 184320 launches).
    Average over the step's launches whose shape was captured; returns (bytes_per_launch, covered_launches)."""
    import csv
    path = os.path.join(ROOT, "profiles", "r01_ncu_gemm_suite.csv")
    try:
        rows = [r for r in csv.reader(open(path)) if r and not r[0].startswith("#")]
        hdr = rows[0]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        cap = {}
        for r in rows[2:]:
            name = r[0]                              # e.g. "rn2 N512 K512 (+res,LN)", "relu N2048 K512"
            n = int(name.split(" N")[1].split()[0]); k = int(name.split(" K")[1].split()[0])
            rn = 2 if name.startswith("rn2") else 1 if name.startswith("rn1") else 0
            cap[(n, k, rn)] = (float(r[ir]) + float(r[iw])) * 1e6      # ncu reports Mbyte
    except Exception:
        return None, 0
    total, hit = 0.0, 0
    for key in keys:
        M, N, K, rn, _res = key
        c = cap.get((N, K, rn))
        if c is not None and M >= 90112:
            total += c * (M / 90112.0)
            hit += 1
    return (total / hit, hit) if hit else (None, 0)


def run_e2e(args, ai1, states, B, dev, world, rank):
    """Public API, host buffers: All_In_One_Model('infer_on_batch') like scripts/test_holisticbody.py:249-255."""
    import types
    import torch.distributed as dist
    vq, pst, rst = states
    cfg = _config()
    a = types.SimpleNamespace(gpu=dev.index, infer=True)
    model = ai1.All_In_One_Model(None, None, ["s2g_body_predictor", "s2g_body_refiner"], [None, None], dev, a, cfg)
    model.body_model[0].VQ.load_state_dict(vq)
    model.body_model[0].load_state_dict(pst)
    model.body_model[1].load_state_dict(rst)
    for m in model.body_model:
        m.sync_timing = False                      # no device-wide sync inside the API call: copies may overlap compute
    aud, gt, mask, ids = synthetic_batch(B, 2004 + rank, "cpu", pin=True)
    out_host = [torch.empty(B, 376, T_FRAMES).pin_memory() for _ in range(2)]
    h2d = sum(t.numel() * t.element_size() for t in (aud, gt, mask, ids))
    d2h = out_host[0].numel() * 4
    # two-deep software pipeline, as a serving loop would run it: a copy stream uploads step i+1's inputs (and
    # drains step i-1's poses) while step i computes on the main stream.  Every step still pays its own H2D + D2H.
    main = torch.cuda.current_stream()
    copy = torch.cuda.Stream()
    dev_in = [[torch.empty_like(t, device=dev) for t in (aud, gt, mask, ids)] for _ in range(2)]
    dev_out = [torch.empty(B, 376, T_FRAMES, device=dev) for _ in range(2)]
    up_done = [torch.cuda.Event() for _ in range(2)]
    comp_done = [torch.cuda.Event() for _ in range(2)]
    down_done = [torch.cuda.Event() for _ in range(2)]

    def upload(i):
        s = i & 1
        with torch.cuda.stream(copy):
            copy.wait_event(comp_done[s])          # the compute that last read these device buffers has finished
            for d, h in zip(dev_in[s], (aud, gt, mask, ids)):
                d.copy_(h, non_blocking=True)
            up_done[s].record(copy)

    def compute(i):
        s = i & 1
        main.wait_event(up_done[s])
        main.wait_event(down_done[s])              # dev_out[s] has been drained
        a_d, g_d, m_d, i_d = dev_in[s]
        out, _ = model("infer_on_batch", aud=a_d, text=None, gt_poses=g_d, mask=m_d, id=i_d, B=B)
        dev_out[s].copy_(out)
        comp_done[s].record(main)

    def download(i):
        s = i & 1
        with torch.cuda.stream(copy):
            copy.wait_event(comp_done[s])
            out_host[s].copy_(dev_out[s], non_blocking=True)
            down_done[s].record(copy)

    def run(n):
        upload(0)
        for i in range(n):
            if i + 1 < n:
                upload(i + 1)
            compute(i)
            download(i)
        torch.cuda.synchronize()

    for e in comp_done + down_done:
        e.record(main)
    run(3)
    if world > 1:
        dist.barrier()
    n = max(2, min(args.steps, 6))
    t0 = time.perf_counter()
    run(n)
    dt = torch.tensor([(time.perf_counter() - t0) / n], device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return {"value": world * B / float(dt), "unit": "clips/s", "h2d_bytes_per_step": h2d * world,
            "d2h_bytes_per_step": d2h * world, "api": "probtalk_b200.nets.All_In_One_Model('infer_on_batch')",
            "steps": n, "pipeline": "copy stream overlaps step i+1 upload / step i-1 download with step i compute"}


def _config():
    """Attribute tree with the fields of config/transformer.json the wrappers read."""
    import types
    ns = types.SimpleNamespace
    return ns(Data=ns(data_root=None, pose=ns(convert_to_6d=True, expression=True, generate_length=180)),
              Model=ns(vq_type="fbhe", hf_type="fbhe", hf_vq_type="fbhe", q_type="pro", groups=4, code_num=128, code_dim=128,
                       share_code=False, knn=False, p_audio="wav2vec", p_text=None, two_stage=True, motion_context=True,
                       transformer=True, identity=True, maskgit=True, maskgit_T=8, vq_path=None, encoder_choice="wav2vec"))


def cpu_baseline(args, states):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Bs, reps = args.ref_batch, 2
    cpu_step(Bs, states, 300)
    t = sum(cpu_step(Bs, states, 301 + i)[0] for i in range(reps))
    return {"value": Bs * reps / t, "unit": "clips/s", "cores": cores, "kind": "port",
            "sample": "%d clips x %d repetitions of the same window workload, oracle port (fp32 PyTorch CPU)" % (Bs, reps)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1024, help="clips per GPU per step (BASELINE config 4 sweeps 64-1024)")
    ap.add_argument("--ref-batch", type=int, default=16, help="clips per CPU-reference step (bounded sample)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
