"""Multi-GPU check of the training loop (torchrun, one rank per GPU): every rank trains the same
synthetic scene with view-pairs sharded over the ranks, gradients and densification statistics
all-reduced, densification replicated. Verifies that the replicas stay bit-identical (same row
count, parameters equal across ranks after densifications) and prints one JSON line from rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/train_multi.py
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("NCCL_DEBUG", "WARN")


def main(config="c2", P=100_000, steps=60, sharded=False):
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    from event_3dgs_b200 import trainer as tr
    scene = syn.make_scene(config, P=P)
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in scene["gt"].items()}
    opt = tr.optimization_defaults(densify_from_iter=10, densification_interval=20, densify_until_iter=10 ** 6)
    t = tr.Trainer(pc, cams, cams, lambda i: gt, opt=opt, event=True, extent=2.6, seed=5, pairs_per_step=2 * world, rank=rank, world=world,
                   sharded_optimizer=sharded)
    ref_leaves = None
    if sharded and world > 1:
        # the same run with the replicated optimizer (all-reduce, then every rank steps every row): the sharded run must
        # land on the same model up to the fp32 order of the gradient sum
        pc_r = pl.GaussianParams(scene["gaussians"], dev)
        t_r = tr.Trainer(pc_r, cams, cams, lambda i: gt, opt=opt, event=True, extent=2.6, seed=5, pairs_per_step=2 * world, rank=rank, world=world)
        for _ in range(steps):
            t_r.step()
        ref_leaves = [p.detach().clone() for p in pc_r.leaves()] + [t_r.c.detach().clone()]
        del t_r
    losses, sizes = [], []
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        losses.append(t.step())
        if "densify" in t.last:
            sizes.append(t.last["densify"]["rows"])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    # replicas identical?
    same = True
    if world > 1:
        rows = torch.tensor([pc._xyz.shape[0]], device=dev)
        lo, hi = rows.clone(), rows.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo.item() == hi.item())
        if same:
            for p in pc.leaves():
                a, b = p.detach().clone(), p.detach().clone()
                dist.all_reduce(a, op=dist.ReduceOp.MIN)
                dist.all_reduce(b, op=dist.ReduceOp.MAX)
                same = same and bool(torch.equal(a, b))
    extra = {}
    if ref_leaves is not None:
        mine = [p.detach() for p in pc.leaves()] + [t.c.detach()]
        ok_rows = all(a.shape == b.shape for a, b in zip(mine, ref_leaves))
        extra = {"sharded_adam_steps": t.sharded_steps, "same_rows_as_replicated_run": ok_rows}
        if ok_rows:
            extra["max_abs_diff_vs_replicated_run_over_max_abs"] = max(float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30) for a, b in zip(mine, ref_leaves))
    if rank == 0:
        print(json.dumps({"world": world, **extra, "pairs_per_step": 2 * world, "steps": steps, "rows_start": P, "rows_after_each_densification": sizes,
                          "replicas_bit_identical": same, "gradient_exchange": t.last_exchange, "loss_first": float(losses[0]), "loss_last": float(losses[-1]),
                          "pairs_per_sec_incl_densification": 2 * world * steps / dt}))
    if world > 1:
        dist.destroy_process_group()
    if not same:
        raise SystemExit("replicas diverged")


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--P", type=int, default=100_000)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--sharded", action="store_true", help="gradient exchange + Adam as one kernel per rank (parallel.ShardedAdam); also runs the replicated optimizer for comparison")
    a = ap.parse_args()
    main(a.config, a.P, a.steps, a.sharded)
