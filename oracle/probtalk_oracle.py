"""CPU oracle for ProbTalk's token-level inference hot path  --  TEST INFRASTRUCTURE.

A plain PyTorch-CPU fp32 restatement of the reference algorithm (the reference itself is
100 % Python/PyTorch, fp32, so the restatement is fp32 torch as well).  Every
function cites the reference file:line it follows (paths relative to the reference root).
It works on plain ``state_dict``s with the reference's key names, so the same checkpoint
dictionaries drive the reference, this oracle and the CUDA product path.

Rules (see DESIGN.md "Oracle"):
  * only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
    ``--impl reference`` leg may import this file -- never the product package;
  * parity pin: ``tests/test_oracle_vs_reference.py`` runs the UNMODIFIED reference
    (oracle/ref_harness.py) next to this file in the authoring container, and
    ``tests/golden/*.pt`` (made by oracle/make_golden.py from the reference itself) pin it
    on machines where /root/reference is absent.  The reference ships no tests or golden
    vectors of its own, so those reference-generated fixtures ARE the pin.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------------------


def _sub(sd, prefix):
    """View of the entries under ``prefix.`` with the prefix stripped."""
    p = prefix + "."
    return {k[len(p):]: v for k, v in sd.items() if k.startswith(p)}


def _bn(x, sd, prefix, eps=1e-5):
    # nn.BatchNorm1d in eval mode (nets/spg/vqvae_modules.py:183-184)
    return F.batch_norm(x, sd[prefix + ".running_mean"], sd[prefix + ".running_var"],
                        sd[prefix + ".weight"], sd[prefix + ".bias"], False, 0.0, eps)


# --------------------------------------------------------------------------------------
# PQ-VAE building blocks
# --------------------------------------------------------------------------------------


def conv_norm_relu(x, sd, prefix, sample="none", residual=False, slope=0.2, leaky=True):
    """ConvNormRelu.forward, nets/spg/vqvae_modules.py:193-200 (ctor :114-191).

    act( BN(conv(x)) [+ residual_layer(x)] );  sample: none = k3 s1 p1, down = k4 s2 p1,
    up = ConvTranspose k4 s2 p1.  Dropout is inactive (eval)."""
    w, b = sd[prefix + ".conv.weight"], sd[prefix + ".conv.bias"]
    if sample == "none":
        out = F.conv1d(x, w, b, stride=1, padding=1)
    elif sample == "down":
        out = F.conv1d(x, w, b, stride=2, padding=1)
    elif sample == "up":
        out = F.conv_transpose1d(x, w, b, stride=2, padding=1)
    else:
        raise ValueError(sample)
    out = _bn(out, sd, prefix + ".norm")
    if residual:
        rw, rb = sd[prefix + ".residual_layer.weight"], sd[prefix + ".residual_layer.bias"]
        if sample == "down":
            out = out + F.conv1d(x, rw, rb, stride=2, padding=1)
        elif sample == "up":
            out = out + F.conv_transpose1d(x, rw, rb, stride=2, padding=1)
        else:
            raise ValueError("residual only used with down/up on this path")
    return F.leaky_relu(out, slope) if leaky else F.relu(out)


def res_cnr_stack(x, sd, prefix, layers=2):
    """Res_CNR_Stack.forward, nets/spg/vqvae_modules.py:238-245: two leaky CNRs, then
    ReLU( BN(conv3(h)) + x )  (leaky_out is False at every call site on this path)."""
    h = x
    for i in range(layers):
        h = conv_norm_relu(h, sd, "%s._layers.%d" % (prefix, i))
    h = F.conv1d(h, sd[prefix + ".conv.weight"], sd[prefix + ".conv.bias"], padding=1)
    h = _bn(h, sd, prefix + ".norm")
    return F.relu(h + x)


def encoder_sc(x, sd):
    """EncoderSC.forward, nets/inpainting/vqvae_1d_sc.py:220-235.  x [B,376,T] -> z [B,512,T//8]."""
    h = conv_norm_relu(x, sd, "encoder.project")
    h = res_cnr_stack(h, sd, "encoder._enc_1")
    h = conv_norm_relu(h, sd, "encoder._down_1", "down", True)
    h = res_cnr_stack(h, sd, "encoder._enc_2")
    h = conv_norm_relu(h, sd, "encoder._down_2", "down", True)
    h = res_cnr_stack(h, sd, "encoder._enc_3")
    h = conv_norm_relu(h, sd, "encoder._down_3", "down", True)
    h = res_cnr_stack(h, sd, "encoder._enc_4")
    return F.conv1d(h, sd["encoder.pre_vq_conv.weight"], sd["encoder.pre_vq_conv.bias"])


def decoder_sc(e, sd):
    """DecoderSC.forward, nets/inpainting/vqvae_1d_sc.py:259-274.  e [B,512,T'] -> [B,376,8T']."""
    h = F.conv1d(e, sd["decoder.aft_vq_conv.weight"], sd["decoder.aft_vq_conv.bias"])
    h = res_cnr_stack(h, sd, "decoder._dec_1")
    h = conv_norm_relu(h, sd, "decoder._up_2", "up", True)
    h = res_cnr_stack(h, sd, "decoder._dec_2")
    h = conv_norm_relu(h, sd, "decoder._up_3", "up", True)
    h = res_cnr_stack(h, sd, "decoder._dec_3")
    h = conv_norm_relu(h, sd, "decoder._up_4", "up", True)
    h = res_cnr_stack(h, sd, "decoder._dec_4")
    return F.conv1d(h, sd["decoder.project.weight"], sd["decoder.project.bias"])


def pq_codebooks(sd, groups=4):
    return [sd["vq_layer.embeddings_%d" % g] for g in range(groups)]


def pq_search(z, codebooks):
    """ProductQuantization.get_code_indices + the per-group loop of .forward (eval),
    nets/spg/vqvae_modules.py:546-555 and :470-491.

    z [B, G*D, W] fp32  ->  indices [B, W, G] int64.  Distances are evaluated exactly as the
    reference writes them, ``(sum(x^2) + sum(c^2)) - 2 * x @ C^T`` in fp32, argmin takes the
    first minimum.  ``random_token`` (:582-590) is the identity on the values for r = 0; its
    RNG draws are a side effect that is deliberately not reproduced."""
    B, C, W = z.shape
    G = len(codebooks)
    D = C // G
    flat = z.permute(0, 2, 1).reshape(-1, G, D)
    idx = []
    for g in range(G):
        x, emb = flat[:, g], codebooks[g]
        dist = (torch.sum(x ** 2, dim=1, keepdim=True) + torch.sum(emb ** 2, dim=1)
                - 2.0 * torch.matmul(x, emb.t()))
        idx.append(torch.argmin(dist, dim=1))
    return torch.stack(idx, dim=1).view(B, W, G)


def pq_search_fp64(z, codebooks):
    """fp64 evaluation of the same distances: returns (indices [B,W,G], gap [B,W,G]) where
    gap is the exact top-2 distance gap.  Used to classify near-ties (gap < eps)."""
    B, C, W = z.shape
    G = len(codebooks)
    D = C // G
    flat = z.permute(0, 2, 1).reshape(-1, G, D).double()
    idx, gap = [], []
    for g in range(G):
        x, emb = flat[:, g], codebooks[g].double()
        dist = (x ** 2).sum(1, keepdim=True) + (emb ** 2).sum(1) - 2.0 * x @ emb.t()
        top2 = torch.topk(dist, 2, dim=1, largest=False)
        idx.append(top2.indices[:, 0])
        gap.append(top2.values[:, 1] - top2.values[:, 0])
    return torch.stack(idx, 1).view(B, W, G), torch.stack(gap, 1).view(B, W, G)


def pq_tie_eps(z, codebooks):
    """eps = 8 * ulp32(|x|^2 + |c|^2) per query: the band inside which the
    fp32 reference's own rounding decides the winner."""
    B, C, W = z.shape
    G = len(codebooks)
    D = C // G
    flat = z.permute(0, 2, 1).reshape(-1, G, D)
    eps = []
    for g in range(G):
        mag = (flat[:, g] ** 2).sum(1) + (codebooks[g] ** 2).sum(1).max()
        ulp = torch.from_numpy(np.spacing(mag.numpy().astype(np.float32)))
        eps.append(8.0 * ulp.double())
    return torch.stack(eps, 1).view(B, W, G)


def pq_gather(indices, codebooks):
    """ProductQuantization.quantize_all, nets/spg/vqvae_modules.py:557-564.
    indices [..., G] -> [..., G*D] (group-major concat)."""
    return torch.cat([F.embedding(indices[..., g].long(), codebooks[g])
                      for g in range(len(codebooks))], dim=-1)


def vqvae_forward(x, sd, groups=4):
    """VQVAE_SC.encode -> vq_layer (eval) -> decode, nets/inpainting/vqvae_1d_sc.py:300-324;
    this is vq_teacher.infer_on_audio (nets/inpainting/vq_teacher.py:255-275).
    Returns (z, indices [B,T',G], poses [B,376,8T'])."""
    z = encoder_sc(x, sd)
    cbs = pq_codebooks(sd, groups)
    idx = pq_search(z, cbs)
    e = pq_gather(idx, cbs).transpose(1, 2).contiguous()
    return z, idx, decoder_sc(e, sd)


# --------------------------------------------------------------------------------------
# predictor: audio encoder + gated-conv context encoder
# --------------------------------------------------------------------------------------


def condition_encoder(x, sd):
    """ConditionEncoder.forward (AudEnc), nets/inpainting/vqvae_1d_sc.py:86-91 (ctor :71-84):
    1x1 conv + BN + LeakyReLU(0.1), then three residual 'down' CNRs.  [B,768,T] -> [B,256,T//8]."""
    h = F.conv1d(x, sd["feature_map.0.weight"], sd["feature_map.0.bias"])
    h = F.leaky_relu(_bn(h, sd, "feature_map.1"), 0.1)
    for i in (1, 2, 3):
        h = conv_norm_relu(h, sd, "_down_%d" % i, "down", True)
    return h


_GATED_GEOM = {0: (1, 3), 2: (2, 1), 4: (2, 1), 6: (2, 1)}   # layer -> (stride, padding); others (1,1)


def gated_conv(x, sd, prefix, stride, padding):
    """GatedConv.forward, nets/inpainting/gated_pixelcnn_1d.py:74-83 (gate :30-36):
    out = res( tanh(a) * sigmoid(b) ) + conv(x),  [a|b] = vert_stack(x); conv = Identity
    unless stride == 2."""
    h = F.conv1d(x, sd[prefix + ".vert_stack.weight"], sd[prefix + ".vert_stack.bias"],
                 stride=stride, padding=padding)
    a, b = h.chunk(2, dim=1)
    out = torch.tanh(a) * torch.sigmoid(b)
    out = F.conv1d(out, sd[prefix + ".res.weight"], sd[prefix + ".res.bias"])
    if stride == 2:
        skip = F.conv1d(x, sd[prefix + ".conv.weight"], sd[prefix + ".conv.bias"],
                        stride=stride, padding=padding)
    else:
        skip = x
    return out + skip


def predictor_context(masked_motion, mask, audio, sd):
    """The non-sampling half of PredictorNet.forward, nets/inpainting/predictornet.py:85-97:
    cat(motion, mask) -> 1x1 embed -> 10 gated convs -> cat with the encoded audio.
    Returns condition [B, 768, T']."""
    mm = torch.cat([masked_motion, mask], dim=1)
    mm = F.conv1d(mm, sd["motion_embedding.weight"], sd["motion_embedding.bias"])
    for i in range(10):
        s, p = _GATED_GEOM.get(i, (1, 1))
        mm = gated_conv(mm, sd, "enc.%d" % i, s, p)
    return torch.cat([mm, audio], dim=1)


# --------------------------------------------------------------------------------------
# transformer layer shared by predictor and refiner
# --------------------------------------------------------------------------------------


def sinusoid_table(d_model, max_len):
    """PositionalEncoding.__init__, nets/inpainting/PTransformer.py:224-243: float64 table,
    row 0 = pad (zeros), row p+1 = position p."""
    pe = np.array([[pos / np.power(10000, 2.0 * (j // 2) / d_model) for j in range(d_model)]
                   for pos in range(max_len)])
    pe[:, 0::2] = np.sin(pe[:, 0::2])
    pe[:, 1::2] = np.cos(pe[:, 1::2])
    return torch.cat((torch.zeros([1, d_model]), torch.from_numpy(pe)))


def style_block(h, label_embed, sd, prefix):
    """StylizationBlock_V3.forward, 'sb' branch, nets/inpainting/PTransformer.py:134-140:
    Linear( SiLU( LN(h) * (1 + scale) + shift ) ), [scale|shift] = Linear(SiLU(label_embed)).
    h [B, N, C], label_embed [B or 1, C]."""
    emb = F.linear(F.silu(label_embed), sd[prefix + ".emb_layers_2.1.weight"],
                   sd[prefix + ".emb_layers_2.1.bias"]).unsqueeze(1)
    scale, shift = torch.chunk(emb, 2, dim=2)
    C = h.shape[-1]
    y = F.layer_norm(h, (C,), sd[prefix + ".norm_2.weight"], sd[prefix + ".norm_2.bias"], 1e-5)
    y = y * (1 + scale) + shift
    return F.linear(F.silu(y), sd[prefix + ".out_layers_2.2.weight"], sd[prefix + ".out_layers_2.2.bias"])


def _mha(q_in, kv_in, sd, prefix, nhead=8):
    """nn.MultiheadAttention (packed in_proj, q|k|v order, scale 1/sqrt(d_head), no masks,
    eval) as used at nets/inpainting/PTransformer.py:340,350.  Batch-first [B, N, C]."""
    W, bias = sd[prefix + ".in_proj_weight"], sd[prefix + ".in_proj_bias"]
    C = q_in.shape[-1]
    q = F.linear(q_in, W[:C], bias[:C])
    k = F.linear(kv_in, W[C:2 * C], bias[C:2 * C])
    v = F.linear(kv_in, W[2 * C:], bias[2 * C:])
    B, N, _ = q.shape
    M = k.shape[1]
    dh = C // nhead
    q = q.view(B, N, nhead, dh).transpose(1, 2)
    k = k.view(B, M, nhead, dh).transpose(1, 2)
    v = v.view(B, M, nhead, dh).transpose(1, 2)
    att = torch.softmax((q @ k.transpose(-1, -2)) / math.sqrt(dh), dim=-1)
    o = (att @ v).transpose(1, 2).reshape(B, N, C)
    return F.linear(o, sd[prefix + ".out_proj.weight"], sd[prefix + ".out_proj.bias"])


def decoder_layer(x, memory, label_embed, sd, prefix):
    """DeepSupervisionLayer.forward, nets/inpainting/PTransformer.py:314-372 (post-norm):
      x = norm1(x + style1(self_attn(x)))
      x = norm2(x + style2(cross_attn(x, memory)))
      x = norm3(x + style3(linear2(relu(linear1(x)))))
    Note the attention / FF outputs reach the residual ONLY through the style block
    (:344-346, :354-357, :361-364).  Batch-first here; the reference is sequence-first."""
    C = x.shape[-1]

    def ln(t, name):
        return F.layer_norm(t, (C,), sd[prefix + "." + name + ".weight"], sd[prefix + "." + name + ".bias"], 1e-5)

    a = _mha(x, x, sd, prefix + ".self_attn")
    x = ln(x + style_block(a, label_embed, sd, prefix + ".style_gourp_1"), "norm1")
    a = _mha(x, memory, sd, prefix + ".multihead_attn")
    x = ln(x + style_block(a, label_embed, sd, prefix + ".style_gourp_2"), "norm2")
    f = F.linear(F.relu(F.linear(x, sd[prefix + ".linear1.weight"], sd[prefix + ".linear1.bias"])),
                 sd[prefix + ".linear2.weight"], sd[prefix + ".linear2.bias"])
    x = ln(x + style_block(f, label_embed, sd, prefix + ".style_gourp_3"), "norm3")
    return x


def tokens_to_logits(tokens, label, condition, sd, n_layers=6):
    """PTransformer.tokens_to_logits, nets/inpainting/PTransformer.py:657-694.
    tokens [B,T',G] int (mask id = 128), label [B] int, condition [B,768,T'] -> logits
    [B,T',G,128].  sd = the 'decoder.*' sub-dict of the Predictor state."""
    b, t, g = tokens.shape
    x = F.embedding(tokens.long(), sd["tok_emb.weight"]).reshape(b, t * g, -1)
    label_embed = F.embedding(label.long(), sd["class_embed.weight"])
    PE = sd["PE.position_encoding.weight"]                      # float64 [1801, 512]
    PPE = sd["product_PE.position_encoding.weight"]             # float64 [5, 512]
    c_pe = PE[1:t + 1].unsqueeze(0)                             # positions 0..t-1   (:262-268)
    x_pe = c_pe.repeat_interleave(g, 1) + PPE[1:g + 1].unsqueeze(0).repeat(1, t, 1)   # (:669-674)
    x = (x + x_pe).to(torch.float32)                            # (:679) fp64 add, then round
    cond = F.conv1d(condition, sd["condi_emb.weight"], sd["condi_emb.bias"]).transpose(1, 2)
    cond = (cond + c_pe).to(torch.float32)                      # (:683)
    for l in range(n_layers):
        x = decoder_layer(x, cond, label_embed, sd, "transformer_decoder.layers.%d" % l)
    out = x.transpose(1, 2)
    out = F.conv1d(out, sd["output_conv.0.weight"], sd["output_conv.0.bias"])
    out = F.conv1d(F.relu(out), sd["output_conv.2.weight"], sd["output_conv.2.bias"])
    return out.permute(0, 2, 1).reshape(b, t, g, -1)


# --------------------------------------------------------------------------------------
# MaskGIT sampler with injectable noise
# --------------------------------------------------------------------------------------

MASK_ID = 128
CHOICE_TEMPERATURE = 4.5


def maskgit_mask_len(n_tokens, unknown_now, step, T):
    """mask_len of PTransformer.sample_good, nets/inpainting/PTransformer.py:610-628, cosine
    schedule (:581-582).  unknown_number_in_the_beginning == n_tokens (all tokens start masked)."""
    ratio = 1.0 * (step + 1) / T
    mask_ratio = np.cos(ratio * np.pi / 2)
    ml = torch.floor(torch.full_like(unknown_now, n_tokens, dtype=torch.float64) * mask_ratio)
    return torch.maximum(torch.zeros_like(ml), torch.minimum(unknown_now.double() - 1, ml)).long()


def categorical_from_uniform(logits, u):
    """Inverse-CDF categorical draw: smallest k with cdf[k] > u * cdf[-1], cdf accumulated in
    fp32 over exp(logit - max) in index order.  Stands in for
    ``Categorical(logits=logits).sample()`` (PTransformer.py:607) whose RNG stream cannot be
    reproduced across devices; the CUDA sampler uses the same rule with the same ``u``."""
    p = torch.exp(logits - logits.max(-1, keepdim=True).values)
    K = logits.shape[-1]
    acc = torch.zeros_like(p[..., 0])
    cols = []
    for k in range(K):                      # sequential fp32 running sum, index order
        acc = acc + p[..., k]
        cols.append(acc)
    cdf = torch.stack(cols, dim=-1)
    thr = (u * cdf[..., -1]).unsqueeze(-1)
    k = (cdf <= thr).sum(-1)
    return k.clamp_(max=K - 1)


def maskgit_step(logits, cur_ids, step, T, u, gumbel):
    """One iteration of PTransformer.sample_good, nets/inpainting/PTransformer.py:603-638,
    plus mask_by_random_topk :696-704, with the two random draws injected:
      u      [B,N] uniform(0,1)  -> the categorical sample
      gumbel [B,N] Gumbel(0,1)   -> the confidence noise
    logits [B,N,K] fp32, cur_ids [B,N] int32 -> new cur_ids [B,N] int32."""
    B, N, K = logits.shape
    sampled = categorical_from_uniform(logits, u).to(torch.int32)
    unknown = cur_ids == MASK_ID
    sampled = torch.where(unknown, sampled, cur_ids)
    ratio = 1.0 * (step + 1) / T
    probs = F.softmax(logits, dim=-1)
    sel = torch.take_along_dim(probs, sampled.long().unsqueeze(-1), -1).squeeze(-1)
    sel = torch.where(unknown, sel, torch.full_like(sel, float("inf")))
    mask_len = maskgit_mask_len(N, unknown.sum(-1, keepdim=True), step, T)
    temperature = CHOICE_TEMPERATURE * (1.0 - ratio)
    conf = torch.log(sel) + temperature * gumbel
    sorted_conf, _ = torch.sort(conf, dim=-1)
    cut = torch.take_along_dim(sorted_conf, mask_len, dim=-1)
    masking = conf < cut
    return torch.where(masking, torch.full_like(sampled, MASK_ID), sampled)


def sample_good(label, condition, sd, T, u_all, gumbel_all, groups=4, return_trace=False):
    """PTransformer.sample_good, nets/inpainting/PTransformer.py:590-642, with injected noise
    u_all / gumbel_all [T, B, N].  Returns ids [B, T', G] (int32)."""
    B, Tq = condition.shape[0], condition.shape[2]
    N = Tq * groups
    cur = torch.full((B, N), MASK_ID, dtype=torch.int32)
    trace = []
    for t in range(T):
        logits = tokens_to_logits(cur.reshape(B, Tq, groups), label, condition, sd).reshape(B, N, -1)
        new = maskgit_step(logits, cur, t, T, u_all[t], gumbel_all[t])
        if return_trace:
            trace.append((cur.clone(), logits))
        cur = new
    ids = cur.reshape(B, Tq, groups)
    return (ids, trace) if return_trace else ids


# --------------------------------------------------------------------------------------
# wrappers: predictor / refiner infer_on_batch
# --------------------------------------------------------------------------------------


def predictor_condition(aud, gt_poses, mask, pred_state):
    """First half of predictor.TrainWrapper.infer_on_batch, nets/inpainting/predictor.py:425-437:
    input_poses = gt * mask; audio = AudEnc(aud); condition from PredictorNet."""
    input_poses = gt_poses * mask
    audio = condition_encoder(aud, pred_state["AudEnc"])
    cond = predictor_context(input_poses, mask, audio, pred_state["Predictor"])
    return input_poses, cond


def predictor_decode(codes, input_poses, mask, vq_sd, T_out, groups=4):
    """Second half, nets/inpainting/predictor.py:439-446: quantize_all -> VQ.decode ->
    F.interpolate(linear) to T_out -> blend with the known frames."""
    e = pq_gather(codes, pq_codebooks(vq_sd, groups)).transpose(1, 2)
    pred = decoder_sc(e.contiguous(), vq_sd)
    pred = F.interpolate(pred, size=T_out, mode="linear")
    return input_poses * mask + pred * (1 - mask)


def predictor_infer_on_batch(aud, gt_poses, mask, ids, pred_state, vq_sd, u_all, gumbel_all, T=8,
                             groups=4):
    """predictor.TrainWrapper.infer_to_code, nets/inpainting/predictor.py:597-647, for
    already-batched inputs (aud [B,768,T], gt_poses [B,376,T], mask [B,1,T], ids [B]) with the
    sampler noise injected.  Returns (poses [B,376,T], codes [B,T',G])."""
    input_poses, cond = predictor_condition(aud, gt_poses, mask, pred_state)
    dec = _sub(pred_state["Predictor"], "decoder")
    codes = sample_good(ids, cond, dec, T, u_all, gumbel_all, groups)
    return predictor_decode(codes, input_poses, mask, vq_sd, aud.shape[2], groups), codes


def rtransformer(x, audio, mask, label, sd, n_layers=6):
    """RTransformer.forward, nets/inpainting/RTransformer.py:209-230.
    x [B,376,T], audio [B,768,T], mask [B,1,T], label [1 or B] -> [B,376,T]."""
    T = x.shape[2]
    PE = sd["PE.position_encoding.weight"]
    pe = PE[1:T + 1].unsqueeze(0)
    h = F.conv1d(x, sd["input_conv.weight"], sd["input_conv.bias"])
    h = (h.transpose(1, 2) + pe).to(torch.float32)
    label_embed = F.embedding(label.long(), sd["class_embed.weight"])
    onehot = F.one_hot(mask.to(torch.long), 2).squeeze(1).transpose(1, 2)          # (:218)
    cond = torch.cat([audio, onehot], dim=1)
    cond = F.conv1d(cond, sd["audio_conv.weight"], sd["audio_conv.bias"]).transpose(1, 2)
    cond = (cond + pe).to(torch.float32)
    for l in range(n_layers):
        h = decoder_layer(h, cond, label_embed, sd, "transformer_decoder.layers.%d" % l)
    out = h.transpose(1, 2)
    out = F.conv1d(out, sd["output_conv.0.weight"], sd["output_conv.0.bias"])
    return F.conv1d(F.relu(out), sd["output_conv.2.weight"], sd["output_conv.2.bias"])


def refiner_infer_on_batch(aud, gt_poses, mask, ids, pred_poses, ref_sd):
    """refiner.TrainWrapper.infer_on_batch, nets/inpainting/refiner.py:512-558 (two_stage,
    type 'fbhe') for already-batched inputs.  ids [1] or [B]."""
    T = aud.shape[2]
    pred_poses = F.interpolate(pred_poses, size=T, mode="linear")
    blank = gt_poses * mask + pred_poses * (1 - mask)
    out = rtransformer(blank, aud, mask, ids, ref_sd)
    return gt_poses * mask + out * (1 - mask)


# --------------------------------------------------------------------------------------
# training-side quantizer update
# --------------------------------------------------------------------------------------


def pq_ema_update(z, idx, codebooks, ema_dw, ema_cs, decay, eps, step):
    """EMA codebook update + dead-code reset of ProductQuantization.forward (training, share_code = False),
    nets/spg/vqvae_modules.py:495-534 with ExponentialMovingAverage.forward :271-276.
    z [B, G*D, W], idx [B, W, G]; codebooks / ema_dw: lists of [K, D], ema_cs: list of [K] (the `.hidden` buffers), all
    updated in place; step = the EMA counter after its increment.  Returns mse(x, quantized) (:537)."""
    B, C, W = z.shape
    G = len(codebooks)
    K, D = codebooks[0].shape
    flat_x = z.permute(0, 2, 1).reshape(-1, G, D)
    quantized = torch.cat([F.embedding(idx[..., g].reshape(-1), codebooks[g]) for g in range(G)], dim=-1)
    mse = F.mse_loss(flat_x.reshape(-1, G * D), quantized)
    for g in range(G):
        x = flat_x[:, g]
        encodings = F.one_hot(idx[..., g].reshape(-1), K).float()
        ema_cs[g].sub_((ema_cs[g] - torch.sum(encodings, dim=0)) * (1 - decay))
        cs = ema_cs[g] / (1 - decay ** step)
        n = torch.sum(cs)
        cs = (cs + eps) / (n + K * eps) * n
        dw = torch.matmul(encodings.t(), x)
        ema_dw[g].sub_((ema_dw[g] - dw) * (1 - decay))
        w = (ema_dw[g] / (1 - decay ** step)) / cs.reshape(-1, 1)
        assert x.shape[0] >= K, "the reference's _tile adds torch.randn noise when the batch has fewer rows than codes"
        code_rand = x[:K]
        usage = (cs.view(K, 1) >= 1.0).float()
        ema_dw[g].copy_(usage * ema_dw[g] + (1 - usage) * code_rand)
        ema_cs[g].mul_(usage.view(-1))
        ema_cs[g].add_(torch.ones_like(ema_cs[g]) * (1 - usage).view(-1))
        codebooks[g].copy_(usage * w + (1 - usage) * code_rand)
    return mse
