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
    step_check = {"n": 0, "param": 0.0, "moment": 0.0}

    def install_step_check():
        """Every sharded step is recomputed from a snapshot with plain torch: NCCL sum of the same gradient buffers, Adam
        in torch arithmetic; the parameters and moments of the rows a rank owns must agree with that to fp32 rounding (the
        other rows of the parameters are the owners' results, compared across ranks at the end: replicas bit-identical)."""
        from event_3dgs_b200 import parallel
        orig = t.sharded.step

        def checked(pc_, c_, optimizer, optimizer_c, exchange, grad_scale, D):
            items = t.sharded._items(pc_, c_, optimizer, optimizer_c, D)
            g_sum = exchange.buf.clone()
            dist.all_reduce(g_sum)
            snap = []
            for p, g, opt_, grp in items:
                st = opt_.state.get(p, {})
                off = g.data_ptr() - exchange.buf.data_ptr()
                gg = g_sum[off // 4: off // 4 + p.numel()].view(p.shape) * grad_scale
                zero = torch.zeros_like(p)
                snap.append((p, p.detach().clone(), st.get("exp_avg", zero).clone(), st.get("exp_avg_sq", zero).clone(), int(st.get("step", 0)) + 1, gg,
                             grp, opt_))
            orig(pc_, c_, optimizer, optimizer_c, exchange, grad_scale, D)
            for p, p0, m0, v0, tt, gg, grp, opt_ in snap:
                b1, b2 = grp["betas"]
                m = b1 * m0 + (1 - b1) * gg
                v = b2 * v0 + (1 - b2) * gg * gg
                p_ref = p0 - (grp["lr"] / (1 - b1 ** tt)) * (m / (v.sqrt() / (1 - b2 ** tt) ** 0.5 + grp["eps"]))
                # rows whose second moment is at the bottom of the fp32 range (gradients ~1e-15: barely visible Gaussians) have a
                # ratio m / sqrt(v) decided by denormal rounding in ANY Adam, torch's included: compare the well-posed ones
                ok = v > 1e-24
                step_check["ill_posed_fraction"] = max(step_check.get("ill_posed_fraction", 0.0), 1.0 - float(ok.float().mean()))
                d = torch.where(ok, (p.detach() - p_ref).abs(), torch.zeros_like(p_ref))
                lo, hi = parallel.shard_range(p.numel(), rank, world)       # this rank's moments are current on its own rows only
                e_p = (float(d.view(-1)[lo:hi].max()) if hi > lo else 0.0) / max(float(p_ref.abs().max()), 1e-30)
                key = "param_%s" % "x".join(str(k) for k in p.shape[1:]) if p.dim() else "param_c"
                step_check[key] = max(step_check.get(key, 0.0), e_p)
                step_check["param"] = max(step_check["param"], e_p)
                mm = opt_.state[p]["exp_avg"].view(-1)[lo:hi]
                if hi > lo:
                    step_check["moment"] = max(step_check["moment"], float((mm - m.view(-1)[lo:hi]).abs().max()) / max(float(m.abs().max()), 1e-30))
            step_check["n"] += 1
        t.sharded.step = checked
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        losses.append(t.step())
        if sharded and world > 1 and t.sharded is not None and step_check["n"] == 0 and not hasattr(t.sharded.step, "__wrapped_check__"):
            install_step_check()
            t.sharded.step.__dict__["__wrapped_check__"] = True
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
        extra = {"sharded_adam_steps": t.sharded_steps, "same_rows_as_replicated_run": ok_rows,
                 "sharded_steps_recomputed_with_torch": step_check["n"], "max_param_diff_vs_torch_adam_on_nccl_sum": step_check["param"],
                 "max_own_moment_diff_vs_torch": step_check["moment"], "rows_left_out_as_ill_posed": step_check.get("ill_posed_fraction"),
                 "per_tensor": {k: v for k, v in step_check.items() if k.startswith("param_")}}
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
