"""Multi-GPU plumbing: event view-pairs are sharded over ranks (one process per GPU, parameters
replicated); the only exchange of a step is the sum of the parameter gradients and of the
densification statistics. The reference has no distributed code at all
(utils/general_utils.py:133 pins cuda:0); what must be reduced is defined by its consumers:
  * gradients of the six parameter groups + the event threshold c (train.py:210-212, 330-332),
  * xyz_gradient_accum += |viewspace grad|, denom += 1, max_radii2D = max(...) over visible
    Gaussians (train.py:319-320, scene/gaussian_model.py:405-407).
Backend: NCCL on GPUs (NVLink 5 / NVSwitch), gloo in the CPU tests.
"""
import torch
import torch.distributed as dist


def shard_pairs(num_pairs, rank, world):
    """Pair b of a step goes to rank b mod world; returns this rank's pair ids."""
    return list(range(rank, num_pairs, world))


def allreduce_gradients(params, world=None, average=True):
    """SUM all-reduce of .grad of every parameter (59 floats per Gaussian + 1 for c), launched as
    one flat bucket per parameter group so NCCL sees a few large messages; divides by the number
    of pairs so that the step equals the mean-loss step of a single process."""
    if not dist.is_initialized():
        return
    world = world or dist.get_world_size()
    if world == 1:
        return
    handles = []
    for p in params:
        if p.grad is None:
            p.grad = torch.zeros_like(p)
        handles.append(dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, async_op=True))
    for h in handles:
        h.wait()
    if average:
        for p in params:
            p.grad.div_(world)


def allreduce_flat(flat, world=None):
    """Mean over ranks of one flat gradient buffer (fused.py keeps all gradients of an iteration in
    one): a single collective; NCCL averages in the reduction itself, gloo sums and divides."""
    if not dist.is_initialized():
        return
    world = world or dist.get_world_size()
    if world == 1:
        return
    if dist.get_backend() == "nccl":
        dist.all_reduce(flat, op=dist.ReduceOp.AVG)
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)


def allreduce_densification_stats(grad_norm_increment, denom_increment, max_radii2D):
    """In-place: SUM of the two accumulator increments, MAX of the screen radii."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    h = [dist.all_reduce(grad_norm_increment, op=dist.ReduceOp.SUM, async_op=True),
         dist.all_reduce(denom_increment, op=dist.ReduceOp.SUM, async_op=True),
         dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, async_op=True)]
    for x in h:
        x.wait()


def densification_increments(viewspace_grad, visibility_filter, radii):
    """Local contributions of one render to the statistics of scene/gaussian_model.py:405-407 and
    train.py:319: (|grad[:, :2]| masked, visibility as float, radii masked)."""
    vis = visibility_filter
    gn = torch.zeros(viewspace_grad.shape[0], 1, device=viewspace_grad.device)
    gn[vis] = torch.norm(viewspace_grad[vis, :2], dim=-1, keepdim=True)
    den = vis.float().unsqueeze(1)
    rad = torch.where(vis, radii.float(), torch.zeros_like(radii, dtype=torch.float32))
    return gn, den, rad
