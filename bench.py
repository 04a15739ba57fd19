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
`cpu_baseline` / `--impl reference`: the UNMODIFIED reference (PyTorch CPU fp32, all host threads) from the snapshot
           oracle/_ref (oracle/build_ref.py) on a bounded sample of the same workload; the oracle port only if the
           snapshot is missing (`kind` says which).
Extra keys (`--no-extras` skips them): `bf16x3` (the tolerance-tight arithmetic mode on the same workload), `cfg1`, `cfg2`,
`cfg3`, `cfg4_sweep`, `cfg4_window2`, `cfg5`, `strong` -- the other BASELINE configs, each a short device-timed run.
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
# reference arm / cpu baseline: the reference's OWN code on the host cores (oracle/_ref), else the oracle port
# --------------------------------------------------------------------------------------------------

def _cpu_info():
    """CPU model, logical / physical cores, torch version (SURVEY.md 8d asks for them next to the CPU number)."""
    model, phys = None, set()
    try:
        pkg = core = None
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and model is None:
                model = line.split(":", 1)[1].strip()
            elif line.startswith("physical id"):
                pkg = line.split(":", 1)[1].strip()
            elif line.startswith("core id"):
                core = line.split(":", 1)[1].strip()
                phys.add((pkg, core))
    except Exception:
        pass
    return {"cpu_model": model, "logical_cores": os.cpu_count() or 1, "physical_cores": len(phys) or None,
            "torch": str(torch.__version__)}


class ReferenceCpu:
    """One window workload on the host: the UNMODIFIED reference (All_In_One_Model.__call__ -> s2g_body_predictor /
    s2g_body_refiner .infer_on_batch, nets/ai1.py:44-60) from oracle/_ref through oracle/ref_harness.py's shims when the
    snapshot is there (kind "reference"), else the oracle port (kind "port")."""

    def __init__(self, states):
        self.states = states
        self.kind, self.gen = "port", None
        try:
            from oracle import ref_harness as RH
            if RH.reference_available():
                from oracle.make_golden_windows import reference_generator
                vq, pst, rst = states
                self.gen = reference_generator(RH.build_predictor(vq, pst), RH.build_refiner(vq, rst))
                self.kind = "reference"
        except Exception as e:          # the port is the documented fall-back, say why it was taken
            sys.stderr.write("reference arm: falling back to the oracle port (%s: %s)\n" % (type(e).__name__, e))
            self.gen = None

    def step(self, B, seed):
        """B clips of one 180-frame window; returns seconds.  The reference API takes ONE audio clip and repeats it B
        times (predictor.py:415), recomputing everything per sample: the same work per clip as B distinct clips."""
        aud, gt, mask, ids = synthetic_batch(B, seed)
        with torch.no_grad():
            if self.gen is not None:
                t0 = time.perf_counter()
                self.gen(forward_type="infer_on_batch", aud=aud[:1], text=aud[:1], gt_poses=gt, mask=mask,
                         id=ids[:1].long(), B=B)
                return time.perf_counter() - t0
            from oracle import probtalk_oracle as O
            vq, pst, rst = self.states
            g = torch.Generator().manual_seed(seed + 1)
            u = torch.rand(8, B, 88, generator=g)
            gum = -torch.log(-torch.log(torch.rand(8, B, 88, generator=g).clamp_(1e-10, 1 - 1e-7)))
            t0 = time.perf_counter()
            poses, _ = O.predictor_infer_on_batch(aud, gt, mask, ids.long(), pst, vq, u, gum)
            O.refiner_infer_on_batch(aud, gt, mask, ids.long(), poses, rst["HFTransformer"])
            return time.perf_counter() - t0


def reference_cfg1(vq_state, reps=10):
    """BASELINE config 1 on the host: one 300-frame clip through the UNMODIFIED reference PQ-VAE
    (s2g_body_vqt: VQ.encoder -> vq_layer -> VQ.decode, vq_teacher.py:255-275), PyTorch CPU fp32; ms per clip or None."""
    try:
        from oracle import ref_harness as RH
        from probtalk_b200 import synthetic as S
        if not RH.reference_available():
            return None
        w = RH.build_vqt(vq_state)
        x = S.synthetic_poses(1, 300, 1001)
        with torch.no_grad():
            def one():
                z, _ = w.VQ.encoder(x)
                qo = w.VQ.vq_layer(z, audio=None, label=None, load_pretrain=False)
                return w.VQ.decode(qo.quantized, None)
            for _ in range(3):
                one()
            t0 = time.perf_counter()
            for _ in range(reps):
                one()
            return (time.perf_counter() - t0) / reps * 1e3
    except Exception as e:
        sys.stderr.write("reference arm: config 1 not timed (%s: %s)\n" % (type(e).__name__, e))
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from probtalk_b200 import synthetic as S
    info = _cpu_info()
    cores = info["logical_cores"]
    torch.set_num_threads(cores)
    states = (S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5))
    arm = ReferenceCpu(states)
    Bs = args.ref_batch
    for i in range(args.warmup):
        arm.step(Bs, 100 + i)
    t = sum(arm.step(Bs, 200 + i) for i in range(args.steps))
    ms = t / max(args.steps, 1) * 1e3
    v = Bs / (ms * 1e-3)
    what = ("the unmodified reference (oracle/_ref: nets.ai1.All_In_One_Model -> infer_on_batch, PyTorch CPU fp32)"
            if arm.kind == "reference" else "oracle port (fp32 PyTorch CPU)")
    sample = "%d of the %d clips per step, %s, %d threads" % (Bs, args.batch, what, cores)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "clips/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config 4: predictor+refiner+PQ-VAE decode, one 180-frame window", "clips_per_step": Bs,
                       "frames": T_FRAMES, "full_batch_per_gpu": args.batch},
            "cpu_baseline": dict({"value": v, "unit": "clips/s", "cores": cores, "kind": arm.kind, "sample": sample,
                                  "cfg1_ms_per_clip": reference_cfg1(states[0])}, **info),
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
        # NCCL's INIT log (communicator size, transport, NVLS) goes to STDERR, so that it can be inspected while stdout
        # stays the single JSON line of rank 0
        # NCCL's INIT log (communicator size, transport, NVLS) must be visible, but stdout is ONE JSON line: the log goes
        # to a per-process file that every rank copies to stderr before it exits, and while NCCL initialises (it prints its
        # version banner straight to stdout) file descriptor 1 points at stderr
        os.environ["NCCL_DEBUG"] = os.environ.get("PT_BENCH_NCCL_DEBUG", "INFO")
        os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
        nccl_log = os.path.join(tempfile.gettempdir(), "pt_bench_nccl_%d.log" % os.getpid())
        os.environ["NCCL_DEBUG_FILE"] = nccl_log
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()                                     # communicator set-up happens here, banner included
        torch.cuda.synchronize()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    _lib.lib()
    B = args.batch
    vq, pst, rst = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
    VW, PW, RW = E.VQVAEWeights(vq, dev), E.PredictorWeights(pst, dev), E.RefinerWeights(rst, dev)
    pred = E.PredictorPlan(PW, VW, B, T_FRAMES, use_graph=not args.no_graph)
    ref = E.RefinerPlan(RW, B, T_FRAMES, T_FRAMES, use_graph=not args.no_graph)
    aud, gt, mask, ids = synthetic_batch(B, 1004 + rank, dev)
    launches_per_step = pred.launches + ref.launches
    # The path's only collective: the NCCL gather of the final poses on rank 0.  It runs on a side stream from a
    # double-buffered staging copy, so step i's gather overlaps step i+1's compute (rank 0 receives 7 x 277 MB per step
    # at N = 8); a staging buffer is reused only after its previous gather has completed.
    side = torch.cuda.Stream() if world > 1 else None
    stage = [torch.empty(B, 376, T_FRAMES, device=dev) for _ in range(2)] if world > 1 else None
    gather_buf = [torch.empty(world * B, 376, T_FRAMES, device=dev) for _ in range(2)] if (world > 1 and rank == 0) else None
    pending = [None, None]
    staged = torch.cuda.Event() if world > 1 else None
    counter = [0]

    def step():
        poses, _ = pred.run(aud, gt, mask, ids)
        out = ref.run(aud, gt, mask, poses, ids)
        if world > 1:
            s_ = counter[0] & 1
            counter[0] += 1
            if pending[s_] is not None:
                pending[s_].wait()                    # the gather that last read stage[s_] (two steps ago) is done
            stage[s_].copy_(out)
            staged.record()
            with torch.cuda.stream(side):
                side.wait_event(staged)
                pending[s_] = dist.gather(stage[s_], list(gather_buf[s_].chunk(world)) if rank == 0 else None, dst=0,
                                          async_op=True)
        return out

    def sync():
        for w in pending:
            if w is not None:
                w.wait()
        if world > 1:
            torch.cuda.current_stream().wait_stream(side)
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
    for w in pending:                              # the last gathers are inside the timed region
        if w is not None:
            w.wait()
    if world > 1:
        torch.cuda.current_stream().wait_stream(side)
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

    extras = {}
    if not args.no_extras:
        del pred, ref                                  # free the headline plans' buffers before the other configs
        torch.cuda.empty_cache()
        extras = run_extras(args, E, S, (vq, pst, rst), (VW, PW, RW), dev, world, rank)

    cpu = None
    if rank == 0 and not args.no_cpu:
        if world == 1:
            cpu = cpu_baseline(args, (vq, pst, rst))
        else:   # the CPU baseline is a rank-0, N = 1 measurement (torchrun also pins OMP to one thread per rank)
            cpu = {"value": None, "unit": "clips/s", "cores": 0, "kind": "reference", "sample": "measured at --gpus 1 only"}
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
                                                "step whose shape is in profiles/r02_ncu_gemm_suite.csv (ncu --set full)"
                                                % (traffic_hits, n_gemm)) if traffic else None,
                             "peak_source": pk["source"] + " bf16 sustained (kernel timed inside a long step)",
                             "launches_per_step": n_gemm, "gemm_ms_per_step": gemm_ms,
                             "kernel_share_of_step": kinds},
                "whole_step_tflops": world * B * GFLOP_PER_CLIP / ms,
                "cpu_baseline": cpu, "e2e": e2e}
        line.update(extras)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        try:
            with open(nccl_log) as f:
                sys.stderr.write(f.read())
            os.unlink(nccl_log)
        except OSError:
            pass


def _time_steps(fn, steps, warm=2):
    """Device time per call of fn (CUDA events on the current stream), after `warm` untimed calls."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def _max_over_ranks(ms, dev, world):
    import torch.distributed as dist
    if world == 1:
        return ms
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def run_extras(args, E, S, states, weights, dev, world, rank):
    """The other BASELINE configs and the tolerance-tight mode, each a short device-timed run (inputs resident in HBM).
    cfg1 / cfg2 / cfg3 / cfg4_sweep / cfg4_window2 / bf16x3 describe ONE GPU (identical on every rank; rank 0 reports);
    `strong` and `cfg5` split a fixed global workload over the N ranks (max over ranks)."""
    from probtalk_b200 import driver as D, ops
    vq, pst, rst = states
    VW, PW, RW = weights
    out = {}
    T = T_FRAMES

    def window_rate(PWx, VWx, RWx, B, frames, known=0, steps=3):
        pred = E.PredictorPlan(PWx, VWx, B, frames)
        ref = E.RefinerPlan(RWx, B, frames, frames)
        g = torch.Generator().manual_seed(77 + B + frames)
        aud = torch.randn(B, 768, frames, generator=g).to(dev)
        gt = torch.zeros(B, 376, frames, device=dev)
        mask = torch.zeros(B, 1, frames, device=dev)
        if known:      # second window of config 4: the first `known` frames are the previous window's output
            gt[:, :, :known] = torch.randn(B, 376, known, generator=g).to(dev)
            mask[:, :, :known] = 1
        ids = (torch.arange(B, device=dev) % 4).int()

        def step():
            poses, _ = pred.run(aud, gt, mask, ids)
            ref.run(aud, gt, mask, poses, ids)
        ms = _time_steps(step, steps, warm=3)              # call 1 eager, call 2 captures the CUDA graphs, call 3 replays
        del pred, ref
        torch.cuda.empty_cache()
        return ms

    # ---- the same workload in split-bf16 ("bf16x3") arithmetic: the tolerance-tight number beside the bf16 headline
    Bx = min(args.batch, 512)
    VW3, PW3, RW3 = (E.VQVAEWeights(vq, dev, precision="bf16x3"), E.PredictorWeights(pst, dev, "bf16x3"),
                     E.RefinerWeights(rst, dev, "bf16x3"))
    ms = window_rate(PW3, VW3, RW3, Bx, T, steps=2)
    out["bf16x3"] = {"value": Bx / (ms * 1e-3), "unit": "clips/s", "ms_per_step": ms, "clips_per_gpu": Bx,
                     "tolerance": "max-abs <= 1e-3, mean <= 1e-4 vs the fp32 reference (observed 5e-5 / 1e-5; "
                                  "tests/test_gpu_precise.py)",
                     "arithmetic": "A_hi W_hi + A_hi W_lo + A_lo W_hi on tcgen05 (fp32 accumulate), fp32 attention / LayerNorm"}
    del VW3, PW3, RW3
    torch.cuda.empty_cache()
    # ---- config 4 batch sweep and second window (T = 150, 30 known frames)
    out["cfg4_sweep"] = {}
    for B in (64, 128, 256, 512):
        ms = window_rate(PW, VW, RW, B, T)
        out["cfg4_sweep"][str(B)] = {"clips_per_s": B / (ms * 1e-3), "ms_per_step": ms}
    ms = window_rate(PW, VW, RW, args.batch, 150, known=30)
    out["cfg4_window2"] = {"clips_per_s": args.batch / (ms * 1e-3), "ms_per_step": ms, "frames": 150, "known_frames": 30,
                           "clips_per_gpu": args.batch}
    # ---- config 3: s2g_body_predictor alone at batch 64 (full 8-step infer_on_batch)
    pred = E.PredictorPlan(PW, VW, 64, T)
    g = torch.Generator().manual_seed(3)
    a3, z3, m3 = torch.randn(64, 768, T, generator=g).to(dev), torch.zeros(64, 376, T, device=dev), torch.zeros(64, 1, T, device=dev)
    i3 = (torch.arange(64, device=dev) % 4).int()
    ms = _time_steps(lambda: pred.run(a3, z3, m3, i3), 5, warm=3)
    out["cfg3"] = {"clips_per_s": 64 / (ms * 1e-3), "ms_per_step": ms, "batch": 64, "what": "predictor infer_on_batch, 8 MaskGIT steps + decode"}
    del pred
    # ---- config 1: one 300-frame clip, encode -> PQ search -> decode, batch 1 (latency)
    enc = E.EncoderPlan(VW, 1, 300)
    dec = E.DecoderPlan(VW, 1, enc.Tq, codes=enc.codes)
    x1 = S.synthetic_poses(1, 300, 1001).to(dev)

    def cfg1():
        enc.run(x1)
        dec.run()
    ms = _time_steps(cfg1, 20, warm=3)
    out["cfg1"] = {"ms_per_clip": ms, "clips_per_s": 1e3 / ms, "frames": 300, "batch": 1, "launches": enc.plan.launches + dec.plan.launches,
                   "what": "VQVAE_SC encode (split-bf16) -> PQ search -> fused gather+aft_vq -> decode (bf16), eager launches"}
    # the same 39 launches replayed as one CUDA graph (what a latency-sensitive caller would do with a fixed clip length)
    torch.cuda.synchronize()
    g1 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g1):
        enc.plan.run()
        dec.plan.run()
    ms_g = _time_steps(g1.replay, 50, warm=3)
    out["cfg1"]["ms_per_clip_graph_replay"] = ms_g
    del g1
    del enc, dec
    # ---- config 2: PQ nearest-code search microbench, 256 clips x 22 token steps x 4 groups; 16 distinct z buffers
    # (184 MB > the 126 MB L2) rotate so that every launch reads its z from HBM
    cb = VW.codebooks
    zs = [torch.randn(256, 22, 512, device=dev) for _ in range(16)]
    idx = torch.empty(256, 22, 4, dtype=torch.int64, device=dev)
    L = E._lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    k = [0]

    def cfg2():
        z = zs[k[0] & 15]
        k[0] += 1
        E.check(L.pt_pq_search(z.data_ptr(), 22 * 512, 1, 512, 256, 22, 4, 128, 128, cb.data_ptr(), idx.data_ptr(), None,
                               None, None, st), "pt_pq_search")
    ms = _time_steps(cfg2, 64, warm=16)
    nbytes = 256 * 22 * 512 * 4 + cb.numel() * 4 + idx.numel() * 8
    out["cfg2"] = {"us_per_launch": ms * 1e3, "queries": 256 * 22 * 4, "achieved_gbs": nbytes / (ms * 1e-3) / 1e9,
                   "hbm_peak_gbs": _peaks()["hbm"], "frac": nbytes / (ms * 1e-3) / 1e9 / _peaks()["hbm"],
                   "algorithmic_bytes": nbytes, "z": "16 rotating buffers (184 MB > L2)"}
    del zs
    torch.cuda.empty_cache()
    # ---- strong scaling: 1024 global clips split over the N ranks
    Bs = max(1024 // world, 1)
    ms = _max_over_ranks(window_rate(PW, VW, RW, Bs, T), dev, world)
    out["strong"] = {"clips_per_s": world * Bs / (ms * 1e-3), "ms_per_step": ms, "global_clips": world * Bs, "clips_per_gpu": Bs,
                     "limit": "per-GPU batch shrinks with N: below ~256 clips the 690 dependent launches per window bound the step"}
    # ---- config 5: 60 s utterances (t = 1800: 12 predictor + 13 refiner windows each), 512 utterances over the N ranks
    n_utt = 512 if world > 1 else 64
    s0, s1 = D.shard_range(n_utt, rank, world)
    pipe = D.HolisticPipeline.from_weights(VW, PW, RW)
    Bu = s1 - s0
    a5 = torch.randn(Bu, 768, 1800, device=dev)
    l5 = (torch.arange(Bu, device=dev) % 4).int()
    pipe.windows_demo(a5, l5)                            # builds the plans, captures the graphs
    pipe.windows_demo(a5, l5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    pipe.windows_demo(a5, l5)
    e1.record()
    torch.cuda.synchronize()
    ms = _max_over_ranks(e0.elapsed_time(e1), dev, world)
    out["cfg5"] = {"utterances_per_s": n_utt / (ms * 1e-3), "frames_per_s": n_utt * 1800 / (ms * 1e-3), "ms": ms,
                   "utterances": n_utt, "utterances_per_gpu": Bu, "frames": 1800,
                   "what": "windows_demo: 12 predictor windows + one decode + 13 refiner windows per utterance"}
    del pipe, a5
    torch.cuda.empty_cache()
    # ---- f1: wav2vec2-base features of 60 s utterances (scripts/demo.py:105-111), 8 utterances per GPU
    # per call; random-init weights of the architecture (the checkpoint is not on the box); the reference's own
    # implementation (transformers, fp32) timed on the host cores for a bounded 10 s sample
    if world > 1:          # (building the transformers model costs ~20 s of host time per rank: reported at N = 1 only)
        return out
    try:
        from probtalk_b200 import audio as A
        model = S.make_wav2vec2(11)
        WW = A.Wav2VecWeights(model.state_dict(), dev)
        Bw, Lw = 8, 60 * 16000
        wplan = A.Wav2VecPlan(WW, Bw, Lw)
        wave = S.synthetic_waveform(1, Lw, 5).to(dev).repeat(Bw, 1).contiguous()
        ms = _max_over_ranks(_time_steps(lambda: wplan.run(wave), 5, warm=3), dev, world)
        out["f1_wav2vec2"] = {"audio_seconds_per_s": world * Bw * 60.0 / (ms * 1e-3), "ms": ms, "utterances_per_gpu": Bw,
                              "seconds_per_utterance": 60.0, "frames_50fps": wplan.T, "frames_30fps": wplan.t_out,
                              "launches": wplan.launches, "tflops": wplan.flops / (ms * 1e-3) / 1e12,
                              "what": "processor normalisation + Wav2Vec2Model.last_hidden_state + linear interpolation to 30 fps, bf16"}
        del wplan, wave
        if rank == 0:
            feats = A.Wav2Vec2Features(WW)
            hw = S.synthetic_waveform(1, Lw, 6)[0].pin_memory()
            feats(hw)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                feats(hw).cpu()
            out["f1_wav2vec2"]["e2e_ms_one_utterance_host_to_host"] = (time.perf_counter() - t0) / 5 * 1e3
            xs = S.synthetic_waveform(1, 10 * 16000, 7)
            xs = (xs - xs.mean()) / torch.sqrt(xs.var(unbiased=False) + 1e-7)
            with torch.no_grad():
                model(xs)
                t0 = time.perf_counter()
                model(xs)
                dt = time.perf_counter() - t0
            out["f1_wav2vec2"]["cpu_reference"] = {"audio_seconds_per_s": 10.0 / dt, "sample": "one 10 s utterance", "threads": torch.get_num_threads(),
                                                   "kind": "transformers Wav2Vec2Model, fp32, host cores"}
        del WW, model
        torch.cuda.empty_cache()
    except ImportError as e:                      # transformers missing: the row is reported as not run, nothing else changes
        out["f1_wav2vec2"] = {"unavailable": str(e)}
    return out


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
    shapes (profiles/r02_ncu_gemm_suite.csv, M = 90112; bytes scale with M for the refiner's M = 184320 launches).
    Average over the step's launches whose shape was captured; returns (bytes_per_launch, covered_launches)."""
    import csv
    path = os.path.join(ROOT, "profiles", "r02_ncu_gemm_suite.csv")
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
    """The reference arm on a bounded sample, in a SUBPROCESS: oracle/ref_harness.py redirects every 'cuda' device to the
    CPU for the life of its process, which must not happen inside the GPU arm."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2", "--warmup", "1",
           "--ref-batch", str(args.ref_batch), "--batch", str(args.batch)]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "OMP_NUM_THREADS")}
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
        return json.loads(line)["cpu_baseline"]
    except Exception as e:
        return {"value": None, "unit": "clips/s", "cores": 0, "kind": "port", "sample": "failed: %s" % e}


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
    ap.add_argument("--no-extras", action="store_true", help="skip the extra keys (other BASELINE configs, bf16x3)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
