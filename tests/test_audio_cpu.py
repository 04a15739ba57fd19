"""Host-side logic of the wav2vec2 path (probtalk_b200/audio.py) that needs no GPU: frame arithmetic and weight packing
against transformers' Wav2Vec2Model (the module scripts/demo.py:76-77 instantiates)."""
import pytest
import torch
import torch.nn.functional as F


@pytest.fixture(scope="module")
def hf_model():
    from probtalk_b200 import synthetic as S
    return S.make_wav2vec2(11)


def test_feat_lengths_match_transformers(hf_model):
    from probtalk_b200 import audio as A
    for L in (400, 401, 16000, 48000, 160007, 960000):
        assert A.feat_lengths(L)[-1] == int(hf_model._get_feat_extract_output_lengths(torch.tensor(L)))


def test_weight_packing(hf_model):
    """The six zero-padded 128-column tiles of pos_conv reproduce the grouped, weight-normalised convolution; QKV rows are
    [q | k | v]; conv taps are tap-major."""
    from probtalk_b200 import audio as A
    W = A.Wav2VecWeights(hf_model.state_dict(), "cpu")
    conv = hf_model.encoder.pos_conv_embed.conv
    g = torch.Generator().manual_seed(0)
    T = 40
    x = torch.randn(1, 768, T, generator=g)
    with torch.no_grad():
        ref = conv(x)[:, :, :T]                                            # 'same' padding, last frame dropped
    xp = F.pad(x, (64, 64)).transpose(1, 2)[0]                             # halo rows [T + 128, 768]
    win = xp.unfold(0, 128, 1)[:T]                                         # [T, 768, 128]: win[t, c, j] = xp[t + j, c]
    got = torch.empty(T, 768)
    for j, off in enumerate(A.W2V_POS_OFFSETS):
        w = W.pos                                                          # one [768, 128 taps * 192] matrix, tile j = rows 128 j ..
        assert w.mat.shape == (768, 128 * 192) and w.n_valid == 768
        a = win[:, off:off + 192, :].permute(0, 2, 1).reshape(T, -1)       # tap-major, like the GEMM's K order
        got[:, 128 * j:128 * (j + 1)] = a @ w.mat[128 * j:128 * (j + 1)].float().t() + w.bias[128 * j:128 * (j + 1)]
    # bf16 weights: compare against the same convolution with bf16-rounded weights
    with torch.no_grad():
        ref16 = F.conv1d(x, conv.weight.bfloat16().float(), conv.bias, padding=64, groups=16)[:, :, :T]
    # (g v / |v| evaluated in another order flips the bf16 rounding of a few weights: 2^-9 relative each)
    assert (got.t()[None] - ref16).abs().max().item() < 5e-3
    assert (got.t()[None] - ref).abs().max().item() < 0.05
    lw, hl = W.layers[3], hf_model.encoder.layers[3]
    assert torch.equal(lw["qkv"].mat[768:1536].float(), hl.attention.k_proj.weight.detach().bfloat16().float())
    assert torch.equal(lw["qkv"].bias[1536:2304], hl.attention.v_proj.bias.detach())
    c3 = hf_model.feature_extractor.conv_layers[3].conv.weight.detach()    # [512, 512, 3]
    assert torch.equal(W.convs[2].mat.float().view(512, 3, 512)[:, 2], c3[:, :, 2].bfloat16().float())


def test_refuses_other_architectures(hf_model):
    from probtalk_b200 import audio as A
    sd = dict(hf_model.state_dict())
    sd["feature_extractor.conv_layers.1.layer_norm.weight"] = torch.ones(512)           # 'layer'-norm feature encoder
    with pytest.raises(ValueError):
        A.Wav2VecWeights(sd, "cpu")
