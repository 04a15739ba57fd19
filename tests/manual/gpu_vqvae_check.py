"""Developer diagnostic: per-stage error of the CUDA PQ-VAE encoder / decoder vs the oracle."""
import os
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import probtalk_oracle as O      # noqa: E402
from probtalk_b200 import engine as E, synthetic as S   # noqa: E402


def stage_refs_enc(x, sd):
    out = {}
    h = O.conv_norm_relu(x, sd, "encoder.project"); out["project"] = h
    for i in (1, 2, 3, 4):
        h = O.res_cnr_stack(h, sd, "encoder._enc_%d" % i); out["_enc_%d" % i] = h
        if i < 4:
            h = O.conv_norm_relu(h, sd, "encoder._down_%d" % i, "down", True); out["_down_%d" % i] = h
    return out


def stage_refs_dec(e, sd):
    out = {}
    h = F.conv1d(e, sd["decoder.aft_vq_conv.weight"], sd["decoder.aft_vq_conv.bias"]); out["aft_vq_conv"] = h
    for i in (1, 2, 3, 4):
        h = O.res_cnr_stack(h, sd, "decoder._dec_%d" % i); out["_dec_%d" % i] = h
        if i < 4:
            h = O.conv_norm_relu(h, sd, "decoder._up_%d" % (i + 1), "up", True); out["_up_%d" % (i + 1)] = h
    return out


def main():
    vq = S.make_vq_state(0)
    B, T = 2, 300
    x = S.synthetic_poses(B, T, 1001)
    W = E.VQVAEWeights(vq, "cuda")
    enc = E.EncoderPlan(W, B, T)
    codes = enc.run(x.cuda())
    torch.cuda.synchronize()
    refs = stage_refs_enc(x, vq)
    for name, buf in enc.stages:
        r = refs[name]
        got = buf.valid().float().transpose(1, 2).cpu()
        print("enc %-10s ref rms %.3f  max err %.4f  rms err %.5f  halo max %.3g" % (
            name, r.pow(2).mean().sqrt(), (got - r).abs().max(), (got - r).pow(2).mean().sqrt(),
            max(buf.t[:, :buf.H].abs().max().item(), buf.t[:, buf.H + buf.T:].abs().max().item())))
    z_ref, idx_ref, rec_ref = O.vqvae_forward(x, vq)
    z = enc.z_channel_first().cpu()
    print("z: ref rms %.3f max err %.4f rms err %.5f" % (z_ref.pow(2).mean().sqrt(), (z - z_ref).abs().max(), (z - z_ref).pow(2).mean().sqrt()))
    mism = (codes.cpu() != idx_ref)
    _, gap = O.pq_search_fp64(z_ref, O.pq_codebooks(vq))
    print("codes: %d / %d differ from the fp32 reference; fp64 gaps at mismatches: %s" % (int(mism.sum()), mism.numel(), gap[mism].tolist()[:10]))
    # search on the CUDA z itself must be exact
    idx_self = O.pq_search(z, O.pq_codebooks(vq))
    print("codes vs oracle search on the CUDA z: %d differ" % int((codes.cpu() != idx_self).sum()))
    dec = E.DecoderPlan(W, B, enc.Tq)
    out = dec.run(idx_ref.cuda())
    torch.cuda.synchronize()
    e = O.pq_gather(idx_ref, O.pq_codebooks(vq)).transpose(1, 2).contiguous()
    refs = stage_refs_dec(e, vq)
    for name, buf in dec.stages:
        r = refs[name]
        got = buf.valid().float().transpose(1, 2).cpu()
        print("dec %-12s ref rms %.3f  max err %.4f  rms err %.5f" % (name, r.pow(2).mean().sqrt(), (got - r).abs().max(), (got - r).pow(2).mean().sqrt()))
    o = out.cpu()
    print("recon (given reference codes): ref rms %.3f max err %.4f mean err %.5f" % (rec_ref.pow(2).mean().sqrt(), (o - rec_ref).abs().max(), (o - rec_ref).abs().mean()))
    # timing
    for Bt in (1, 64, 256):
        enc = E.EncoderPlan(W, Bt, 300); dec = E.DecoderPlan(W, Bt, 37)
        xx = S.synthetic_poses(Bt, 300, 5).cuda()
        for _ in range(3):
            enc.run(xx); dec.run(enc.codes)
        torch.cuda.synchronize()
        st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.record()
        for _ in range(10):
            enc.run(xx); dec.run(enc.codes)
        en.record(); torch.cuda.synchronize()
        ms = st.elapsed_time(en) / 10
        print("B=%d: %.3f ms / enc+pq+dec  -> %.0f clips/s, %.1f TFLOP/s (4.71 GFLOP/clip), launches %d" % (
            Bt, ms, Bt / ms * 1e3, Bt * 4.71e9 / ms / 1e9, enc.plan.launches + dec.plan.launches))


if __name__ == "__main__":
    main()
