"""PQ nearest-code search + gather on the GPU (through the C ABI) vs the oracle.

Bar (BASELINE.json north_star / SURVEY.md 8d cfg 2): indices bit-exact with the reference's fp32
path, except at near-ties whose exact (fp64) top-2 distance gap is below
eps = 8 * ulp32(|x|^2 + |c|^2); every mismatch must be such a near-tie."""
import os

import pytest
import torch

from oracle import probtalk_oracle as O
from probtalk_b200 import synthetic as S

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.pt")


def _codebooks(kind, seed):
    g = torch.Generator().manual_seed(seed)
    if kind == "xavier":                                        # the reference's init, vqvae_modules.py:463-464
        return (torch.rand(4, 128, 128, generator=g) * 2 - 1) * (6.0 / 256) ** 0.5
    return torch.randn(4, 128, 128, generator=g)


def _check(z, cb, idx_gpu):
    cbs = [cb[g] for g in range(cb.shape[0])]
    ref = O.pq_search(z, cbs)
    idx64, gap = O.pq_search_fp64(z, cbs)
    eps = O.pq_tie_eps(z, cbs)
    idx_gpu = idx_gpu.cpu()
    bad = idx_gpu != ref
    # every disagreement with the fp32 reference is a near-tie ...
    assert bool((gap[bad] < eps[bad]).all()), "mismatch outside the near-tie band: %s" % gap[bad].max()
    # ... and away from near-ties the answer is the exact argmin
    clear = gap >= eps
    assert torch.equal(idx_gpu[clear], idx64[clear])
    return int(bad.sum())


@pytest.mark.parametrize("scale", [1.0, 0.1])
@pytest.mark.parametrize("cbk", ["xavier", "normal"])
def test_cfg2_index_exact(scale, cbk):
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(1002)
    z = torch.randn(256, 512, 22, generator=g) * scale          # BASELINE config 2: 256 clips x 22 steps x 4 groups
    cb = _codebooks(cbk, 7)
    idx, gap, nres = ops.pq_search(z.cuda(), cb.cuda(), want_gap=True, count_rescored=True)
    torch.cuda.synchronize()
    nbad = _check(z, cb, idx)
    assert nbad <= 8                                           # observed: 0-2 of 22 528
    assert 0 < int(nres) < 0.05 * idx.numel()                   # the fp32 re-score path is exercised but rare
    # the reported gap agrees with the exact one where it is not tiny
    _, gap64 = O.pq_search_fp64(z, [cb[i] for i in range(4)])
    big = gap64 > 1e-2
    assert torch.allclose(gap.cpu()[big].double(), gap64[big], rtol=0.1, atol=1e-3)


def test_golden_cfg2():
    from probtalk_b200 import ops
    gold = torch.load(GOLD, map_location="cpu", weights_only=False)
    vq = S.make_vq_state(gold["seeds"]["vq"])
    cb = torch.stack(O.pq_codebooks(vq))
    c = gold["cfg2"]
    g = torch.Generator().manual_seed(c["seed"])
    for case in c["cases"]:
        z = torch.randn(c["B"], 512, c["W"], generator=g) * case["scale"]
        idx = ops.pq_search(z.cuda(), cb.cuda()).cpu()
        bad = idx != case["codes"].long()
        _, gap = O.pq_search_fp64(z, O.pq_codebooks(vq))
        assert bool((gap[bad] < O.pq_tie_eps(z, O.pq_codebooks(vq))[bad]).all())


def test_tc_scores_within_error_band():
    """The split-bf16 tensor-core scores are within the band the kernel assumes (2.5e-4 |x| max|c|)."""
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(3)
    z = torch.randn(64, 512, 22, generator=g) * 3.0
    cb = _codebooks("normal", 9)
    idx, dump = ops.pq_search(z.cuda(), cb.cuda(), score_dump=True)
    flat = z.permute(0, 2, 1).reshape(-1, 4, 128).double()
    d = dump.cpu().view(-1, 4, 128).double()
    for gi in range(4):
        x, e = flat[:, gi], cb[gi].double()
        exact = (e ** 2).sum(1)[None] - 2 * x @ e.t()
        bound = 2.5e-4 * x.norm(dim=1, keepdim=True) * e.norm(dim=1).max()
        assert bool(((d[:, gi] - exact).abs() <= bound).all())


@pytest.mark.parametrize("B,W", [(1, 1), (1, 37), (3, 18), (7, 129), (130, 1)])
def test_ragged_shapes(B, W):
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + W)
    z = torch.randn(B, 512, W, generator=g)
    cb = _codebooks("xavier", 11)
    _check(z, cb, ops.pq_search(z.cuda(), cb.cuda()))


def test_empty_and_errors():
    from probtalk_b200 import ops
    from probtalk_b200._lib import PtError
    cb = _codebooks("xavier", 1).cuda()
    idx = ops.pq_search(torch.zeros(0, 512, 22, device="cuda"), cb)
    assert idx.shape == (0, 22, 4)
    with pytest.raises(PtError):
        ops.pq_search(torch.zeros(2, 384, 4, device="cuda"), torch.zeros(4, 64, 96, device="cuda"))      # D = 96 is not built
    with pytest.raises(PtError):
        ops.pq_search(torch.zeros(2, 256, 4, device="cuda"), torch.zeros(4, 40, 64, device="cuda"))      # K % 16 != 0
    with pytest.raises(PtError):
        ops.pq_search(torch.zeros(2, 512, 4), _codebooks("xavier", 1))      # CPU tensors: no fallback


def test_exact_ties_take_first_index():
    """Duplicate codes: torch.argmin returns the first minimum and so must the kernel."""
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(5)
    cb = _codebooks("normal", 13)
    cb[:, 77] = cb[:, 5]                 # code 77 duplicates code 5 in every group
    cb[:, 100] = cb[:, 5]
    z = cb[:, 5].reshape(1, 512, 1).repeat(4, 1, 3) + 0.01 * torch.randn(4, 512, 3, generator=g)
    idx = ops.pq_search(z.cuda(), cb.cuda()).cpu()
    assert bool((idx == 5).all())
    assert torch.equal(idx, O.pq_search(z, [cb[i] for i in range(4)]))


def test_channels_last_input_and_gather_roundtrip():
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(17)
    z = torch.randn(512, 512, 22, generator=g)
    cb = _codebooks("xavier", 19)
    zc = z.cuda()
    idx = ops.pq_search(zc, cb.cuda())
    z_cl = zc.permute(0, 2, 1).contiguous().permute(0, 2, 1)       # same logical tensor, channels-last storage
    assert torch.equal(ops.pq_search(z_cl, cb.cuda()), idx)
    q, bad = ops.pq_gather(idx, cb.cuda())
    assert int(bad) == 0
    ref = O.pq_gather(idx.cpu(), [cb[i] for i in range(4)]).transpose(1, 2)
    assert torch.equal(q.cpu(), ref)                              # gather is a bit-exact copy
    # idempotence: quantised vectors map to themselves (property that holds at any size)
    assert torch.equal(ops.pq_search(q, cb.cuda()), idx)
    # int32 indices, bf16 channels-last output, out-of-range indices are counted
    q16, _ = ops.pq_gather(idx.int(), cb.cuda(), channels_last=True, dtype=torch.bfloat16)
    assert torch.equal(q16.float().cpu(), ref.transpose(1, 2).bfloat16().float())
    idx_bad = idx.clone()
    idx_bad[0, 0, 0] = 999
    _, bad = ops.pq_gather(idx_bad, cb.cuda())
    assert int(bad) == 1                                          # one bad (row, group) entry


def test_full_size_properties():
    """4096 clips x 22 steps (16x config 2): checked through size-independent properties only."""
    from probtalk_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(23)
    z = torch.randn(4096, 512, 22, generator=g, device="cuda")
    cb = _codebooks("xavier", 29).cuda()
    idx, gap = ops.pq_search(z, cb, want_gap=True)
    assert int(idx.min()) >= 0 and int(idx.max()) < 128 and bool((gap >= 0).all())
    q, _ = ops.pq_gather(idx, cb)
    # the chosen code is at least as close as 8 random other codes, for every query
    flat = z.permute(0, 2, 1).reshape(-1, 4, 128)
    dq = ((flat - q.permute(0, 2, 1).reshape(-1, 4, 128)) ** 2).sum(-1)
    for j in torch.randint(0, 128, (8,)).tolist():
        dj = ((flat - cb[:, j][None]) ** 2).sum(-1)
        assert bool((dq <= dj + 1e-3).all())
    assert torch.equal(ops.pq_search(q, cb), idx)


@pytest.mark.parametrize("G,K,D,B,W", [(1, 2048, 64, 64, 37), (1, 512, 64, 7, 5), (1, 64, 64, 33, 22), (2, 16, 128, 9, 4),
                                       (4, 144, 128, 16, 22), (3, 128, 64, 256, 22), (1, 2048, 128, 5, 9)])
@pytest.mark.parametrize("scale", [1.0, 0.1])
def test_generalized_shapes_vs_oracle(G, K, D, B, W, scale):
    """J1: VectorQuantizerEMA.get_code_indices (vqvae_modules.py:350-358; smplx_body_vq.py:48-53: one codebook,
    embedding_dim = 64, num_embeddings = code_num = 2048) and other (G, K, D) -- the same formula, the same near-tie bar.
    K > 128 goes through the chunked search (fp32 chunk winners merged by packed atomicMin)."""
    from probtalk_b200 import ops
    g = torch.Generator().manual_seed(G * 1000 + K + D)
    z = torch.randn(B, G * D, W, generator=g) * scale
    cb = torch.randn(G, K, D, generator=g) * (1.0 if K < 1024 else 0.3)
    for zz in (z.cuda(), z.permute(0, 2, 1).contiguous().cuda().permute(0, 2, 1)):      # channel-first and channels-last
        idx = ops.pq_search(zz, cb.cuda())
        torch.cuda.synchronize()
        assert int(idx.min()) >= 0 and int(idx.max()) < K
        _check(z, cb, idx)
    # duplicated codes: the FIRST of two identical codes must win, across chunk boundaries too
    cb2 = cb.clone()
    if K >= 32:
        cb2[:, K - 1] = cb2[:, 3]
        x = cb2[:, 3].reshape(1, G * D, 1).repeat(2, 1, 3)
        idx = ops.pq_search(x.cuda(), cb2.cuda()).cpu()
        assert bool((idx == 3).all())
    # gather with the same codebooks
    q, bad = ops.pq_gather(ops.pq_search(z.cuda(), cb.cuda()), cb.cuda())
    ref = O.pq_gather(O.pq_search(z, [cb[i] for i in range(G)]), [cb[i] for i in range(G)]).transpose(1, 2)
    same = ops.pq_search(z.cuda(), cb.cuda()).cpu() == O.pq_search(z, [cb[i] for i in range(G)])
    assert int(bad) == 0 and q.shape == ref.shape
    if bool(same.all()):
        assert torch.equal(q.cpu(), ref)


def test_vector_quantizer_ema_facade():
    """nets.VectorQuantizerEMA: same buffer names as the reference layer, indices equal to its formula."""
    import probtalk_b200.nets as nets
    vq = nets.VectorQuantizerEMA(64, 2048).cuda()
    assert set(vq.state_dict()) == {"embeddings", "ema_dw", "ema_cluster_size"}
    g = torch.Generator().manual_seed(5)
    vq.embeddings.copy_(torch.randn(2048, 64, generator=g) * 0.3)
    x = torch.randn(3, 64, 50, generator=g)
    flat = x.permute(0, 2, 1).reshape(-1, 64)
    idx = vq.get_code_indices(flat.cuda()).cpu()
    emb = vq.embeddings.cpu()
    dist = (flat ** 2).sum(1, keepdim=True) + (emb ** 2).sum(1) - 2.0 * flat @ emb.t()       # vqvae_modules.py:352-356
    ref = dist.argmin(1)
    d64 = (flat.double() ** 2).sum(1, keepdim=True) + (emb.double() ** 2).sum(1) - 2.0 * flat.double() @ emb.double().t()
    top2 = d64.topk(2, dim=1, largest=False).values
    bad = idx != ref
    assert bool(((top2[:, 1] - top2[:, 0])[bad] < 1e-4).all()) and bad.float().mean() < 0.01
    q, _ = vq(x.cuda())
    assert q.shape == x.shape and torch.equal(q.cpu().permute(0, 2, 1).reshape(-1, 64), emb[idx])
