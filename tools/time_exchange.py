"""Times the gradient exchange alone (torchrun): NCCL all-reduce vs FusedGradientExchange.allreduce on a
c3-sized flat gradient buffer (14,000,004 floats at SH degree 0; 59,000,004 at degree > 0)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("NCCL_DEBUG", "WARN")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from event_3dgs_b200 import parallel
    out = {"world": world}
    for n in (14_000_004, 59_000_004):
        ex = parallel.FusedGradientExchange(n, dev)
        u = torch.randn(n, device=dev)

        def timed(fn, k=20):
            for _ in range(5):
                fn()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / k], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        ex.buf.copy_(u)
        out[str(n)] = {"nccl_ms": timed(lambda: dist.all_reduce(u, op=dist.ReduceOp.AVG)), "nvls_kernel_with_barriers_ms": timed(lambda: ex.allreduce(1.0 / world)),
                       "barrier_pair_ms": timed(lambda: (ex.hdl.barrier(channel=0), ex.hdl.barrier(channel=0)))}
        del ex
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
