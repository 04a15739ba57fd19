"""Host-side logic of the batched driver on CPU: window arithmetic and the world_size-2 gloo sharding path."""
import math
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from probtalk_b200 import driver as D


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 1024, 1025):
        for world in (1, 2, 3, 8):
            blocks = [D.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for a, b in zip(blocks, blocks[1:]):
                assert a[1] == b[0]
            sizes = [e - s for s, e in blocks]
            assert max(sizes) - min(sizes) <= 1


def test_eval_window_bounds_match_reference_formula():
    # scripts/test_holisticbody.py:240-246
    for t in (180, 300, 330, 331, 511, 1800):
        w = D.eval_window_bounds(t)
        assert w[0] == (0, min(180, t))
        n_ref = 1 + math.ceil((t - 180) / 150)
        assert len(w) <= n_ref
        covered = set()
        for i, (s, e) in enumerate(w):
            if i:
                assert s == 150 + 180 * (i - 1)
            covered.update(range(s, e))
        if t <= 330:
            assert covered == set(range(t))


def test_demo_window_bounds():
    # scripts/demo.py:118 with t = 300 -> two windows [0,180), [150,300); refiner [0,176), [146,300)
    assert D.demo_window_bounds(300, 180) == [(0, 180), (150, 300)]
    assert D.demo_window_bounds(300, 176) == [(0, 176), (146, 300)]
    w = D.demo_window_bounds(1800, 180)
    assert len(w) == 12 and w[-1] == (1650, 1800)
    w = D.demo_window_bounds(1800, 176)
    assert len(w) == 13 and w[-1] == (1752, 1800)
    # odd utterance lengths give odd tail windows (RefinerPlan takes any number of frames)
    assert D.demo_window_bounds(1801, 176)[-1] == (1752, 1801)
    assert D.eval_window_bounds(331)[-1] == (330, 331)


def test_plan_cache_is_lru_bounded():
    from probtalk_b200 import engine as E
    c, built = E.PlanCache(capacity=2), []
    for k in ("a", "b", "a", "c", "b"):
        c.get(k, lambda k=k: built.append(k) or k)
    assert built == ["a", "b", "c", "b"] and len(c) == 2 and "a" not in c


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_units, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def compute(s, e):          # stand-in for the per-rank CUDA pipeline: a result block that encodes the unit index
        idx = torch.arange(s, e, dtype=torch.float32)
        return idx[:, None, None] * torch.ones(1, 3, 5) + 0.25 * rank * 0
    out = D.run_sharded(n_units, compute)
    if rank == 0:
        q.put(out)
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_units", [8, 7, 1])
def test_sharded_gather_world2_gloo(n_units):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_units, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert out.shape == (n_units, 3, 5)
    assert torch.equal(out[:, 0, 0], torch.arange(n_units, dtype=torch.float32))
