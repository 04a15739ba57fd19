"""Parameter containers that mirror the reference's module tree (names and state_dict keys).

The classes below carry the SAME attribute names as the reference modules so that reference
checkpoints (``torch.save({'generator': wrapper.state_dict()})``) load unchanged -- SURVEY.md 8(b)
lists the key sets, ``tests/golden/state_dict_manifest.json`` pins them.  They hold parameters
only: ``forward`` is never the compute path.  Compute goes through ``probtalk_b200.engine`` (CUDA
library); calling a container's ``forward`` raises.

reference classes mirrored:
  ConvNormRelu / Res_CNR_Stack / ProductQuantization     nets/spg/vqvae_modules.py:107-245, :437-590
  EncoderSC / DecoderSC / VQVAE_SC / ConditionEncoder    nets/inpainting/vqvae_1d_sc.py:70-91, :197-324
  GatedConv                                              nets/inpainting/gated_pixelcnn_1d.py:39-83
  PositionalEncoding / StylizationBlock_V3 / DeepSupervisionLayer / PTransformer
                                                         nets/inpainting/PTransformer.py:28-165, :214-268, :284-372, :394-455
  PredictorNet                                           nets/inpainting/predictornet.py:13-58
  RTransformer                                           nets/inpainting/RTransformer.py:181-207
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn


class _Container(nn.Module):
    def forward(self, *a, **k):   # pragma: no cover
        raise RuntimeError("%s is a parameter container; run it through probtalk_b200.engine" % type(self).__name__)


class ConvNormRelu(_Container):
    def __init__(self, cin, cout, sample="none", residual=False):
        super().__init__()
        k, s = (3, 1) if sample == "none" else (4, 2)
        conv = nn.ConvTranspose1d if sample == "up" else nn.Conv1d
        if residual:
            self.residual_layer = conv(cin, cout, k, s, 1)
        self.conv = conv(cin, cout, k, s, 1)
        self.norm = nn.BatchNorm1d(cout)
        self.sample = sample


class Res_CNR_Stack(_Container):
    def __init__(self, channels, layers=2):
        super().__init__()
        self._layers = nn.ModuleList([ConvNormRelu(channels, channels) for _ in range(layers)])
        self.conv = nn.Conv1d(channels, channels, 3, 1, 1)
        self.norm = nn.BatchNorm1d(channels)


class EncoderSC(_Container):
    def __init__(self, in_dim, embedding_dim, nh, layers):
        super().__init__()
        self.project = ConvNormRelu(in_dim, nh // 4)
        self._enc_1 = Res_CNR_Stack(nh // 4, layers)
        self._down_1 = ConvNormRelu(nh // 4, nh // 4, "down", True)
        self._enc_2 = Res_CNR_Stack(nh // 4, layers)
        self._down_2 = ConvNormRelu(nh // 4, nh // 2, "down", True)
        self._enc_3 = Res_CNR_Stack(nh // 2, layers)
        self._down_3 = ConvNormRelu(nh // 2, nh, "down", True)
        self._enc_4 = Res_CNR_Stack(nh, layers)
        self.pre_vq_conv = nn.Conv1d(nh, embedding_dim, 1, 1)


class DecoderSC(_Container):
    def __init__(self, out_dim, embedding_dim, nh, layers):
        super().__init__()
        self.aft_vq_conv = nn.Conv1d(embedding_dim, nh, 1, 1)
        self._dec_1 = Res_CNR_Stack(nh, layers)
        self._up_2 = ConvNormRelu(nh, nh // 2, "up", True)
        self._dec_2 = Res_CNR_Stack(nh // 2, layers)
        self._up_3 = ConvNormRelu(nh // 2, nh // 4, "up", True)
        self._dec_3 = Res_CNR_Stack(nh // 4, layers)
        self._up_4 = ConvNormRelu(nh // 4, nh // 4, "up", True)
        self._dec_4 = Res_CNR_Stack(nh // 4, layers)
        self.project = nn.Conv1d(nh // 4, out_dim, 1, 1)


class _EMA(_Container):
    def __init__(self, init):
        super().__init__()
        self.register_buffer("hidden", torch.zeros_like(init))


class ProductQuantizationParams(_Container):
    """Buffers of ProductQuantization (embeddings_{g}, ema_dw.{g}.hidden, ema_cluster_size.{g}.hidden)."""

    def __init__(self, embedding_dim, num_embeddings, num_chunks):
        super().__init__()
        self.num_chunks, self.num_embeddings = num_chunks, num_embeddings
        self.embedding_dim = embedding_dim // num_chunks
        self.ema_dw = nn.ModuleList()
        self.ema_cluster_size = nn.ModuleList()
        for g in range(num_chunks):
            e = torch.empty(num_embeddings, self.embedding_dim)
            nn.init.xavier_uniform_(e)
            self.register_buffer("embeddings_%d" % g, e)
            self.ema_dw.append(_EMA(e))
            self.ema_cluster_size.append(_EMA(torch.zeros(num_embeddings)))


class ConditionEncoder(_Container):
    def __init__(self, in_dim, nh):
        super().__init__()
        self.feature_map = nn.Sequential(nn.Conv1d(in_dim, nh, 1, 1), nn.BatchNorm1d(nh), nn.LeakyReLU(0.1),
                                         nn.Dropout(0.1))
        self._down_1 = ConvNormRelu(nh, nh, "down", True)
        self._down_2 = ConvNormRelu(nh, nh, "down", True)
        self._down_3 = ConvNormRelu(nh, nh, "down", True)


class GatedConv(_Container):
    def __init__(self, dim, kernel, stride, padding):
        super().__init__()
        in_dim = dim // 2 if (stride == 2 and dim > 128) else dim
        self.vert_stack = nn.Conv1d(in_dim, dim * 2, kernel, stride, padding)
        self.conv = nn.Conv1d(in_dim, dim, kernel, stride, padding) if stride == 2 else nn.Identity()
        self.res = nn.Conv1d(dim, dim, 1)
        self.geom = (in_dim, dim, kernel, stride, padding)


def sinusoid_table(d_model, max_len):
    """float64 sinusoid table with a leading all-zero pad row (row p+1 = position p)."""
    pos = np.arange(max_len, dtype=np.float64)[:, None]
    j = np.arange(d_model)
    # same expression order as the reference so the fp64 values are identical
    tab = pos / np.power(10000, 2.0 * (j // 2) / d_model)
    tab[:, 0::2] = np.sin(tab[:, 0::2])
    tab[:, 1::2] = np.cos(tab[:, 1::2])
    return torch.cat((torch.zeros([1, d_model]), torch.from_numpy(tab)))


class PositionalEncoding(_Container):
    def __init__(self, d_model, max_seq_len):
        super().__init__()
        self.position_encoding = nn.Embedding(max_seq_len + 1, d_model)
        self.position_encoding.weight = nn.Parameter(sinusoid_table(d_model, max_seq_len), requires_grad=False)


class StylizationBlock_V3(_Container):
    def __init__(self, dim):
        super().__init__()
        self.emb_layers_2 = nn.Sequential(nn.SiLU(), nn.Linear(dim, dim * 2))
        self.norm_2 = nn.LayerNorm(dim)
        self.out_layers_2 = nn.Sequential(nn.SiLU(), nn.Dropout(0.0), nn.Linear(dim, dim))


class DeepSupervisionLayer(_Container):
    def __init__(self, d_model, nhead, groups, dim_ff=2048):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d_model, nhead, dropout=0.1)
        self.multihead_attn = nn.MultiheadAttention(d_model, nhead, dropout=0.1)
        self.linear1 = nn.Linear(d_model, dim_ff)
        self.linear2 = nn.Linear(dim_ff, d_model)
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(d_model), nn.LayerNorm(d_model), nn.LayerNorm(d_model)
        self.style_gourp_1 = StylizationBlock_V3(d_model)     # (sic) -- the reference's spelling is part of the key
        self.style_gourp_2 = StylizationBlock_V3(d_model)
        self.style_gourp_3 = StylizationBlock_V3(d_model)
        if groups > 1:
            for i in (1, 2, 3):
                p = nn.Parameter(torch.zeros(1, d_model, groups - 1))
                nn.init.xavier_uniform_(p)
                setattr(self, "pad_%d" % i, p)


class _LayerStack(_Container):
    def __init__(self, make, n):
        super().__init__()
        self.layers = nn.ModuleList([make() for _ in range(n)])


class PTransformer(_Container):
    def __init__(self, groups, condi_dim, dim, n_layers, num_code, n_classes):
        super().__init__()
        self.groups, self.num_code, self.mask_id = groups, num_code, num_code
        self.tok_emb = nn.Embedding(num_code + 1, dim)
        self.class_embed = nn.Embedding(n_classes, dim)
        self.condi_emb = nn.Conv1d(condi_dim, dim, 1)
        self.PE = PositionalEncoding(dim, 1800)
        self.product_PE = PositionalEncoding(dim, groups)
        self.transformer_decoder = _LayerStack(lambda: DeepSupervisionLayer(dim, 8, groups), n_layers)
        self.output_conv = nn.Sequential(nn.Conv1d(dim, 512, 1), nn.ReLU(True), nn.Conv1d(512, num_code, 1))


class PredictorNet(_Container):
    GEOM = [(7, 1, 3), (3, 1, 1), (4, 2, 1), (3, 1, 1), (4, 2, 1), (3, 1, 1), (4, 2, 1), (3, 1, 1), (3, 1, 1), (3, 1, 1)]

    def __init__(self, mot_dim, num_code, dim, n_classes, groups, maskgit_T):
        super().__init__()
        self.T = maskgit_T
        self.motion_embedding = nn.Conv1d(mot_dim + 1, dim // 4, 1, 1)
        self.enc = nn.ModuleList()
        d = dim // 4
        for i, (k, s, p) in enumerate(self.GEOM):
            if i in (4, 6):
                d *= 2
            self.enc.append(GatedConv(d, k, s, p))
        self.decoder = PTransformer(groups, d + 256, 512, 6, num_code, n_classes)


class RTransformer(_Container):
    def __init__(self, input_dim, condi_dim, dim, n_layers, n_classes):
        super().__init__()
        self.class_embed = nn.Embedding(n_classes, dim)
        self.PE = PositionalEncoding(dim, 1800)
        self.input_conv = nn.Conv1d(input_dim, dim, 1, 1, 0)
        self.audio_conv = nn.Conv1d(condi_dim + 2, dim, 1, 1, 0)
        self.transformer_decoder = _LayerStack(lambda: DeepSupervisionLayer(dim, 8, 1), n_layers)
        self.output_conv = nn.Sequential(nn.Conv1d(dim, 1024, 1), nn.ReLU(True), nn.Conv1d(1024, input_dim, 1))


class VQVAE_SC(_Container):
    def __init__(self, in_dim, embedding_dim, num_embeddings, num_hiddens=1024, num_residual_layers=2, groups=4):
        super().__init__()
        self.in_dim, self.embedding_dim, self.num_embeddings, self.groups = in_dim, embedding_dim, num_embeddings, groups
        self.encoder = EncoderSC(in_dim, embedding_dim, num_hiddens, num_residual_layers)
        self.vq_layer = ProductQuantizationParams(embedding_dim, num_embeddings, groups)
        self.decoder = DecoderSC(in_dim, embedding_dim, num_hiddens, num_residual_layers)
