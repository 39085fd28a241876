"""world_size-2 gloo test (CPU) of the view-pair sharding and the gradient / densification-
statistics reduction that the multi-GPU trainer uses (event_3dgs_b200/parallel.py)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from event_3dgs_b200 import parallel
    torch.manual_seed(0)                       # replicated parameters
    P = 1000
    params = [torch.nn.Parameter(torch.randn(P, 3)), torch.nn.Parameter(torch.randn(P, 15, 3)),
              torch.nn.Parameter(torch.tensor(0.17))]
    pairs = parallel.shard_pairs(6, rank, world)
    # a "loss" whose gradient depends on the pair id: mean over this rank's pairs
    loss = sum(((p * (b + 1)).sum()) for b in pairs for p in params) / len(pairs)
    loss.backward()
    parallel.allreduce_gradients(params, world)
    # statistics of one render per rank
    g = torch.Generator().manual_seed(100 + rank)
    vis = torch.rand(P, generator=g) < 0.5
    vgrad = torch.randn(P, 3, generator=g)
    radii = torch.randint(0, 30, (P,), generator=g, dtype=torch.int32)
    gn, den, rad = parallel.densification_increments(vgrad, vis, radii)
    local = (gn.clone(), den.clone(), rad.clone())
    parallel.allreduce_densification_stats(gn, den, rad)
    # moments of a sharded optimizer: every rank advanced only the rows it owns; sync_moments makes them whole
    sh = parallel.ShardedAdam(torch.device("cpu"))
    q = torch.nn.Parameter(torch.zeros(1001, 3))
    opt = torch.optim.Adam([q], lr=0.1)
    opt.state[q] = {"step": 3, "exp_avg": torch.full((1001, 3), -1.0), "exp_avg_sq": torch.full((1001, 3), -1.0)}
    lo, hi = parallel.shard_range(q.numel(), rank, world)
    opt.state[q]["exp_avg"].view(-1)[lo:hi] = torch.arange(lo, hi, dtype=torch.float32)
    opt.state[q]["exp_avg_sq"].view(-1)[lo:hi] = 2.0 * torch.arange(lo, hi, dtype=torch.float32)
    sh.moments_full = False
    sh.sync_moments([opt])
    moments = (opt.state[q]["exp_avg"].clone(), opt.state[q]["exp_avg_sq"].clone())
    flat = torch.arange(10, dtype=torch.float32) * (rank + 1)
    parallel.allreduce_flat(flat, world)
    torch.save(dict(flat=flat, pairs=pairs, moments=moments, grads=[p.grad for p in params], stats=(gn, den, rad), local=local), out % rank)
    dist.destroy_process_group()


def test_two_rank_gradient_and_stats_reduction(tmp_path):
    world = 2
    out = str(tmp_path / "r%d.pt")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    r = [torch.load(out % i) for i in range(world)]
    assert torch.allclose(r[0]["flat"], torch.arange(10, dtype=torch.float32) * 1.5) and torch.equal(r[0]["flat"], r[1]["flat"])
    assert r[0]["pairs"] == [0, 2, 4] and r[1]["pairs"] == [1, 3, 5]
    # single-process result: mean over all 6 pairs of d/dp sum(p*(b+1)) = mean(b+1) = 3.5
    for g0, g1 in zip(r[0]["grads"], r[1]["grads"]):
        assert torch.equal(g0, g1)
        assert torch.allclose(g0, torch.full_like(g0, 3.5))
    for k in range(3):
        assert torch.equal(r[0]["stats"][k], r[1]["stats"][k])
    for k in range(world):
        assert torch.equal(r[k]["moments"][0].view(-1), torch.arange(3003, dtype=torch.float32))
        assert torch.equal(r[k]["moments"][1].view(-1), 2.0 * torch.arange(3003, dtype=torch.float32))
    assert torch.allclose(r[0]["stats"][0], r[0]["local"][0] + r[1]["local"][0])
    assert torch.allclose(r[0]["stats"][1], r[0]["local"][1] + r[1]["local"][1])
    assert torch.equal(r[0]["stats"][2], torch.maximum(r[0]["local"][2], r[1]["local"][2]))


def test_shard_range_partitions_every_group():
    """parallel.shard_range = the row ranges e3_nvls_sharded_adam gives the ranks: disjoint, ordered, covering."""
    sys.path.insert(0, ROOT)
    from event_3dgs_b200.parallel import shard_range
    for n in (1, 3, 4, 5, 1000, 1001, 3_000_003):
        for w in (1, 2, 3, 8):
            cover = 0
            for k in range(w):
                lo, hi = shard_range(n, k, w)
                assert lo == cover or lo == hi == n, (n, w, k)
                assert lo % 4 == 0 or lo == n
                cover = max(cover, hi)
            assert cover == n
