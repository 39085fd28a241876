"""Probe of torch symmetric memory / NVLS multicast on this box (torchrun, >= 2 ranks)."""
import json
import os

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as sm

os.environ.setdefault("NCCL_DEBUG", "WARN")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    n = 14_000_000
    t = sm.empty(n, dtype=torch.float32, device=dev)
    hdl = sm.rendezvous(t, dist.group.WORLD)
    out["multicast_ptr"] = int(hdl.multicast_ptr)
    out["has_multicast_support"] = bool(getattr(hdl, "has_multicast_support", lambda *a: None) and hdl.multicast_ptr != 0)
    out["buffer_ptrs"] = [int(p) for p in hdl.buffer_ptrs]
    out["local_ptr"] = int(t.data_ptr())
    t.fill_(rank + 1.0)
    hdl.barrier(channel=0)
    torch.cuda.synchronize()
    try:
        gname = dist.group.WORLD.group_name
        for _ in range(3):
            torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname)
        torch.cuda.synchronize()
        t.fill_(rank + 1.0)
        hdl.barrier(channel=0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname)
        e1.record()
        torch.cuda.synchronize()
        out["multimem_all_reduce_ms"] = e0.elapsed_time(e1)
        out["multimem_ok"] = bool(torch.allclose(t, torch.full_like(t, world * (world + 1) / 2)))
    except Exception as ex:
        out["multimem_error"] = repr(ex)[:300]
    u = torch.full((n,), rank + 1.0, device=dev)
    for _ in range(3):
        dist.all_reduce(u)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dist.all_reduce(u)
    e1.record()
    torch.cuda.synchronize()
    out["nccl_all_reduce_ms"] = e0.elapsed_time(e1)
    e0.record()
    for _ in range(10):
        hdl.barrier(channel=0)
    e1.record()
    torch.cuda.synchronize()
    out["barrier_ms"] = e0.elapsed_time(e1) / 10
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
