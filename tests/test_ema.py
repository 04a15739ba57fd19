"""Training-side quantizer update: EMA codebook update + dead-code reset, nets/spg/vqvae_modules.py:495-534.
Golden = consecutive training forwards of the UNMODIFIED reference layer (oracle/make_golden_ema.py)."""
import os

import pytest
import torch

from oracle import probtalk_oracle as O
from oracle.make_golden_ema import make_batches

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ema_vectors.pt")


@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD, map_location="cpu", weights_only=False)


def _close(a, b, tol=2e-5):
    return (a - b).abs().max().item() <= tol * max(1.0, b.abs().max().item())


def test_oracle_ema_matches_reference_golden(gold):
    for case in gold["cases"]:
        c = case["cfg"]
        cb, zs = make_batches(c)
        dw = [torch.zeros_like(x) for x in cb]
        cs = [torch.zeros(c["K"]) for _ in cb]
        for step, (z, ref) in enumerate(zip(zs, case["steps"]), 1):
            idx = O.pq_search(z, cb)
            assert torch.equal(idx, ref["idx"].long())
            mse = O.pq_ema_update(z, idx, cb, dw, cs, 0.99, 1e-5, step)
            assert abs(0.25 * float(mse) - ref["loss"]) <= 1e-6 * max(1.0, ref["loss"])
            for g in range(c["G"]):
                assert _close(cb[g], ref["codebooks"][g], 1e-6) and _close(dw[g], ref["ema_dw"][g], 1e-6)
                assert _close(cs[g], ref["ema_cs"][g], 1e-6)


@pytest.mark.gpu
def test_ema_kernel_matches_reference_golden(gold):
    """pt_pq_ema_update chained over the golden's steps, state carried on the device; fp64 row sums vs the reference's fp32
    one-hot matmul: agreement to ~1e-6 relative.  The dead-code re-seeding (codes whose normalised EMA count is < 1 take
    the batch's first rows) must hit exactly the reference's codes."""
    from probtalk_b200 import ops
    for case in gold["cases"]:
        c = case["cfg"]
        cb0, zs = make_batches(c)
        cb = torch.stack(cb0).cuda().contiguous()
        dw = torch.zeros_like(cb)
        cs = torch.zeros(c["G"], c["K"], device="cuda")
        for step, (z, ref) in enumerate(zip(zs, case["steps"]), 1):
            zc = z.cuda()
            idx = ops.pq_search(zc, cb)
            assert (idx.cpu() == ref["idx"].long()).float().mean() >= 0.999
            idx = ref["idx"].long().cuda().contiguous()                 # carry on from the reference's own decisions
            before = cb.clone()
            mse, dead = ops.pq_ema_update(zc, idx, cb, dw, cs, 0.99, 1e-5, step)
            assert abs(0.25 * float(mse) - ref["loss"]) <= 2e-6 * max(1.0, ref["loss"])
            for g in range(c["G"]):
                assert _close(cb[g].cpu(), ref["codebooks"][g]) and _close(dw[g].cpu(), ref["ema_dw"][g])
                assert _close(cs[g].cpu(), ref["ema_cs"][g])
                # re-seeded codes = rows of the batch, bit-exact
                flat = z.permute(0, 2, 1).reshape(-1, c["G"], c["D"])[:c["K"], g]
                reseeded = (ref["codebooks"][g] == flat).all(1)
                assert int(dead[g]) == int(reseeded.sum())
                assert torch.equal(cb[g].cpu()[reseeded], flat[reseeded])
            assert not torch.equal(before, cb)
        # deterministic: the same chain again gives the same bits
        cb2 = torch.stack(cb0).cuda().contiguous()
        dw2, cs2 = torch.zeros_like(cb2), torch.zeros(c["G"], c["K"], device="cuda")
        for step, (z, ref) in enumerate(zip(zs, case["steps"]), 1):
            ops.pq_ema_update(z.cuda(), ref["idx"].long().cuda().contiguous(), cb2, dw2, cs2, 0.99, 1e-5, step)
        assert torch.equal(cb, cb2) and torch.equal(dw, dw2) and torch.equal(cs, cs2)


@pytest.mark.gpu
def test_ema_update_rejects_small_batches():
    from probtalk_b200 import ops
    from probtalk_b200._lib import PtError
    cb = torch.zeros(1, 128, 64, device="cuda")
    with pytest.raises(PtError):
        ops.pq_ema_update(torch.zeros(2, 64, 5, device="cuda"), torch.zeros(2, 5, 1, dtype=torch.int64, device="cuda"), cb,
                          torch.zeros_like(cb), torch.zeros(1, 128, device="cuda"), 0.99, 1e-5, 1)


@pytest.mark.gpu
def test_pq_layer_training_forward(gold):
    """nets.VQVAE(...).vq_layer.train_forward: search + EMA update + commitment loss through the module's own buffers
    (embeddings_{g}, ema_dw.{g}.hidden, ema_cluster_size.{g}.hidden), against the reference's training forwards."""
    import probtalk_b200.nets as nets
    case = gold["cases"][0]
    c = case["cfg"]
    cb0, zs = make_batches(c)
    vq = nets.VQVAE(376, c["G"] * c["D"], c["K"], 1024, 2, c["G"], device="cuda")
    sd = vq.state_dict()
    for g in range(c["G"]):
        sd["vq_layer.embeddings_%d" % g] = cb0[g]
        sd["vq_layer.ema_dw.%d.hidden" % g] = torch.zeros_like(cb0[g])
        sd["vq_layer.ema_cluster_size.%d.hidden" % g] = torch.zeros(c["K"])
    vq.load_state_dict(sd)
    for z, ref in zip(zs, case["steps"]):
        qo = vq.vq_layer.train_forward(z.cuda())
        assert abs(float(qo.loss) - ref["loss"]) <= 2e-6 * max(1.0, ref["loss"])
        assert qo.quantized.shape == z.shape
        for g in range(c["G"]):
            assert _close(getattr(vq.vq_layer, "embeddings_%d" % g).cpu(), ref["codebooks"][g])
            assert _close(vq.vq_layer.ema_cluster_size[g].hidden.cpu(), ref["ema_cs"][g])
    # the inference plans see the updated codebooks
    idx = vq.vq_layer(zs[0].cuda()).loss.cpu()
    assert torch.equal(idx, O.pq_search(zs[0], case["steps"][-1]["codebooks"])) or (idx != O.pq_search(zs[0], case["steps"][-1]["codebooks"])).float().mean() < 0.002
