"""Multi-GPU check of the fused gradient exchange (torchrun, >= 2 ranks): the gradients that the
multimem.red path leaves in every rank's copy must equal local accumulation + NCCL all-reduce (AVG),
and be identical on all ranks. Also times both. Prints one JSON line from rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/check_fused_exchange.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("NCCL_DEBUG", "WARN")


def main(config="c3", P=None, pairs_per_rank=1, steps=15):
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from event_3dgs_b200 import fused, parallel, pipeline as pl, synthetic as syn
    scene = syn.make_scene(config, P=P)
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in scene["gt"].items()}
    bg = torch.zeros(3, device=dev)
    c = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    pairs = pairs_per_rank * world
    mine = parallel.shard_pairs(pairs, rank, world)
    view = lambda step, b: 2 + (7 * (step * pairs + b)) % 80

    def step_nccl(s):
        pc._grad_exchange = None
        for k, b in enumerate(mine):
            v = view(s, b)
            fused.event_iteration_backward(pc, cams[v], cams[v], cams[v + 1], gt, c, bg, accumulate=k > 0)
        if len(mine) > 1:
            pc._grad_flat.div_(len(mine))
        parallel.allreduce_flat(pc._grad_flat, world)

    def step_ex(s, mode):
        ex = parallel.fused_exchange_for(pc, fused.grad_flat_numel(pc, int(pc.active_sh_degree)), dev, mode=mode)
        assert ex is not None, "no multicast support"
        pc._grad_exchange, pc._grad_scale = ex, 1.0 / pairs
        for k, b in enumerate(mine):
            v = view(s, b)
            fused.event_iteration_backward(pc, cams[v], cams[v], cams[v + 1], gt, c, bg, accumulate=k > 0)
        if mode == "push":
            ex.end()
        else:
            ex.allreduce(1.0 / pairs)

    step_pull = lambda s: step_ex(s, "pull")
    step_push = lambda s: step_ex(s, "push")
    step_nccl(0)                   # sizes the workspace (synchronous first call)
    step_nccl(0)
    ref = pc._grad_flat.clone()
    res = {}
    for name, fn in (("pull", step_pull), ("push", step_push)):
        fn(0)
        got = pc._grad_flat.clone()
        torch.cuda.synchronize()
        lo, hi = got.clone(), got.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        res[name] = {"max_abs_diff_over_max_abs": float((got - ref).abs().max() / ref.abs().max()), "identical_on_all_ranks": bool(torch.equal(lo, hi))}

    def timed(fn):
        for s in range(3):
            fn(s + 1)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s + 4)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t = {"nccl": [], "pull": [], "push": []}
    for rep in range(2):
        t["nccl"].append(timed(step_nccl))
        t["pull"].append(timed(step_pull))
        t["push"].append(timed(step_push))
    if rank == 0:
        print(json.dumps({"world": world, "config": config, "pairs_per_step": pairs, "parity_vs_nccl": res, "ms_per_step": t,
                          "grad_floats": int(ref.numel())}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main(*(sys.argv[1:2]))
