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


class NoMulticast(RuntimeError):
    """The symmetric-memory rendezvous worked but the system offers no NVLS multicast address."""


class FusedGradientExchange:
    """Gradient exchange over NVLink 5 / NVSwitch with our own kernels instead of an NCCL call.

    The flat gradient buffer of fused.py lives in symmetric memory (torch.distributed's
    _symmetric_memory: plumbing only) that is mapped on every rank together with its NVLS multicast
    alias. Two modes:

    "pull" (default): the backward pass accumulates locally as on one GPU; allreduce() then runs ONE
        kernel per rank (e3_nvls_allreduce): each rank fetches the in-switch SUM of its 1/world of the
        vectors (multimem.ld_reduce) and stores scale * sum into every rank's copy (multimem.st),
        between two cross-rank barriers.
    "push": e3_event_iteration is told the offset to the multicast alias (grad_multicast_delta) and
        its gradient-accumulation kernels reduce every row into ALL ranks' copies with multimem.red
        while the blends of the other views still run; begin() zeroes and fences first, end() fences.
        No kernel follows the backward pass, but every view's partial gradient crosses the fabric
        (3x the bytes, world x the incoming bytes per GPU): measured slower than "pull" and than NCCL
        (profiles/r1_fused_exchange_*.json); kept as the fused compute+collective variant.

    Raises if the box has no multicast support (then use allreduce_flat)."""

    def __init__(self, numel, device, group=None, mode="pull"):
        import torch.distributed._symmetric_memory as symm
        group = group or dist.group.WORLD
        self.buf = symm.empty(numel, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group)
        mc = int(self.hdl.multicast_ptr)
        if mc == 0:
            raise NoMulticast("FusedGradientExchange: no NVLS multicast support on this system")
        self.delta = mc - self.buf.data_ptr()
        self.world, self.rank = int(self.hdl.world_size), int(self.hdl.rank)
        self.numel = numel
        self.mode = mode
        self.device = device

    # ---- push mode ----
    def begin(self):
        self.buf.zero_()
        self.hdl.barrier(channel=0)

    def end(self):
        self.hdl.barrier(channel=0)

    def multicast_ptr_of(self, tensor):
        """NVLS alias of a view into the buffer."""
        return tensor.data_ptr() + self.delta

    # ---- pull mode ----
    def allreduce(self, scale):
        """buf <- scale * sum over ranks of buf, on the current stream."""
        from . import _lib
        st = torch.cuda.current_stream(self.device).cuda_stream
        self.hdl.barrier(channel=0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().e3_nvls_allreduce(self.buf.data_ptr() + self.delta, self.numel, self.rank, self.world, float(scale), st))
        self.hdl.barrier(channel=0)


def shard_range(n, rank, world):
    """Floats [lo, hi) of an n-float parameter group that `rank` owns in e3_nvls_sharded_adam: the rank-th share of
    its 16-byte vectors, ceil(vectors / world) each (the last ranks may own less, or nothing)."""
    nv = (n + 3) // 4
    per = (nv + world - 1) // world
    b = min(per * rank, nv)
    e = min(b + per, nv)
    return min(4 * b, n), min(4 * e, n)


class ShardedAdam:
    """Gradient exchange and optimizer step of the replicated trainer as ONE kernel per rank (e3_nvls_sharded_adam):
    in-switch sum of the rank's 1/world of the gradient vectors -> Adam on those rows -> multicast store of the new
    parameters into every rank's copy ("reduce-scatter -> sharded Adam -> all-gather"). The model's
    six leaf tensors and the event threshold c are re-homed as views into one symmetric-memory buffer (adopt(),
    collective); the gradients already live in the symmetric buffer of FusedGradientExchange. Optimizer state keeps
    torch's layout (full-size exp_avg / exp_avg_sq per parameter), but a rank only advances the rows it owns: call
    sync_moments() before anything that reads or edits whole moment tensors (densification, opacity reset, a
    replicated optimizer step, checkpoints)."""

    LEAVES = (("xyz", "_xyz"), ("features_dc", "_features_dc"), ("opacity", "_opacity"), ("scaling", "_scaling"),
              ("rotation", "_rotation"), ("features_rest", "_features_rest"))

    def __init__(self, device, group=None):
        self.device, self.group = device, group or dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.pbuf = self.hdl = None
        self.delta = 0
        self.moments_full = True

    @staticmethod
    def _rekey(optimizer, old, new):
        for g in optimizer.param_groups:
            for k, q in enumerate(g["params"]):
                if q is old:
                    g["params"][k] = new
        if old in optimizer.state:
            optimizer.state[new] = optimizer.state.pop(old)

    def adopt(self, pc, c, optimizer, optimizer_c):
        """Collective: moves the current parameter values into a fresh symmetric buffer and replaces the model's leaf
        tensors (and c) by views into it; the optimizers' groups and states follow the new tensors. Returns c."""
        import torch.distributed._symmetric_memory as symm
        params = [getattr(pc, attr) for _, attr in self.LEAVES]
        sizes = [(t.numel() + 3) // 4 * 4 for t in params] + [4]
        self.pbuf = symm.empty(sum(sizes), dtype=torch.float32, device=self.device)
        self.pbuf.zero_()
        self.hdl = symm.rendezvous(self.pbuf, self.group)
        mc = int(self.hdl.multicast_ptr)
        if mc == 0:
            raise NoMulticast("ShardedAdam: no NVLS multicast support on this system")
        self.delta = mc - self.pbuf.data_ptr()
        off = 0
        with torch.no_grad():
            for (name, attr), t, sz in zip(self.LEAVES, params, sizes[:-1]):
                view = self.pbuf[off:off + t.numel()].view(t.shape)
                view.copy_(t.detach())
                new = torch.nn.Parameter(view)
                self._rekey(optimizer, t, new)
                setattr(pc, attr, new)
                off += sz
            cview = self.pbuf[off:off + 1].view(c.shape)
            cview.copy_(c.detach())
            cnew = torch.nn.Parameter(cview)
            self._rekey(optimizer_c, c, cnew)
        self.hdl.barrier(channel=0)          # every copy holds the parameters before anyone's multicast store can land
        return cnew

    def owns(self, pc):
        return self.pbuf is not None and pc._xyz.data_ptr() == self.pbuf.data_ptr()

    def _items(self, pc, c, optimizer, optimizer_c, D):
        out = []
        for name, attr in self.LEAVES:
            if name == "features_rest" and D == 0:
                continue
            p = getattr(pc, attr)
            grp = next(g for g in optimizer.param_groups if any(q is p for q in g["params"]))
            out.append((p, pc._grad_views[name], optimizer, grp))
        out.append((c, pc._grad_views["c"], optimizer_c, optimizer_c.param_groups[0]))
        return out

    @torch.no_grad()
    def step(self, pc, c, optimizer, optimizer_c, exchange, grad_scale, D):
        """One fused exchange + Adam step. `exchange` is the FusedGradientExchange whose buffer the backward pass
        accumulated into (pc._grad_views are views of it); grad_scale turns the sum over ranks into the mean."""
        from . import _lib
        if pc._grad_flat is not exchange.buf:
            raise RuntimeError("ShardedAdam.step: the gradients are not in the exchange buffer")
        buckets = {}
        for p, g, opt, grp in self._items(pc, c, optimizer, optimizer_c, D):
            st = opt.state[p]
            if len(st) == 0:
                st["step"] = 0
                st["exp_avg"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
                st["exp_avg_sq"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
            st["step"] = int(st["step"]) + 1
            st.pop("_moments_zero", None)
            b1, b2 = grp["betas"]
            buckets.setdefault((st["step"], float(b1), float(b2)), []).append((p, g, st, float(grp["lr"]), float(grp["eps"])))
        L = _lib.lib()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        exchange.hdl.barrier(channel=0)          # every rank's local gradients are complete
        for (t, b1, b2), items in buckets.items():
            arr = (_lib.AdamGroup * len(items))()
            for k, (p, g, st, lr, eps) in enumerate(items):
                arr[k].param, arr[k].grad = p.data_ptr(), g.data_ptr()
                arr[k].exp_avg, arr[k].exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                arr[k].n, arr[k].lr, arr[k].eps = p.numel(), lr, eps
            with torch.cuda.device(self.device):
                _lib.check(L.e3_nvls_sharded_adam(arr, len(items), t, b1, b2, exchange.delta, self.delta, self.rank, self.world,
                                                  float(grad_scale), stream))
        self.hdl.barrier(channel=0)              # every rank's parameter stores have landed everywhere
        self.moments_full = False

    @torch.no_grad()
    def sync_moments(self, optimizers):
        """Every rank receives, for every parameter, the moment rows their owners advanced (one broadcast per owner
        and tensor; rare: before densification, opacity reset, checkpoints, or a replicated step)."""
        if self.moments_full:
            return
        for opt in optimizers:
            for grp in opt.param_groups:
                for p in grp["params"]:
                    st = opt.state.get(p, None)
                    if not st or "exp_avg" not in st:
                        continue
                    for key in ("exp_avg", "exp_avg_sq"):
                        flat = st[key].view(-1)
                        for r in range(self.world):
                            lo, hi = shard_range(flat.numel(), r, self.world)
                            if hi > lo:
                                dist.broadcast(flat[lo:hi], src=dist.get_global_rank(self.group, r), group=self.group)
        self.moments_full = True


_EXCHANGE_UNAVAILABLE = False


def fused_exchange_for(pc, numel, device, mode="pull"):
    """The FusedGradientExchange matching the model's current gradient layout, created collectively
    (every rank must call this in the same step) when the model changed size; None when the process
    group has one rank or the system has no NVLS multicast."""
    global _EXCHANGE_UNAVAILABLE
    if _EXCHANGE_UNAVAILABLE or not dist.is_initialized() or dist.get_world_size() == 1 or dist.get_backend() != "nccl":
        return None
    ex = getattr(pc, "_grad_exchange_cache", None)
    if ex is not None and ex.numel == numel and ex.mode == mode:
        return ex
    pc._grad_exchange_cache = None
    try:
        ex = FusedGradientExchange(numel, device, mode=mode)
    except NoMulticast as e:
        # the one expected reason (same outcome on every rank of a homogeneous box): said once, then NCCL is used.
        # Anything else -- a failed rendezvous, an allocation error -- is a fault and propagates.
        import warnings
        warnings.warn("event_3dgs_b200.parallel: %s; gradients are exchanged with NCCL instead" % e)
        _EXCHANGE_UNAVAILABLE = True
        return None
    pc._grad_exchange_cache = ex
    return ex


def allreduce_densification_stats(grad_norm_increment, denom_increment, max_radii2D):
    """In-place: SUM of the two accumulator increments, MAX of the screen radii -- as ONE collective over a
    packed 12-byte-per-Gaussian buffer: every rank contributes its three rows [3, P] and receives all ranks'
    (all-gather, 12 P bytes per rank), then sums / maximises locally. One launch instead of three reductions; the
    rows are tiny against the gradient buffer, so the gathered volume does not matter."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    world = dist.get_world_size()
    P = max_radii2D.shape[0]
    packed = torch.stack((grad_norm_increment.reshape(P), denom_increment.reshape(P), max_radii2D.reshape(P).to(grad_norm_increment.dtype)))
    gathered = torch.empty(world * 3, P, dtype=packed.dtype, device=packed.device)
    dist.all_gather_into_tensor(gathered, packed.contiguous())
    gathered = gathered.view(world, 3, P)
    grad_norm_increment.copy_(gathered[:, 0].sum(0).view_as(grad_norm_increment))
    denom_increment.copy_(gathered[:, 1].sum(0).view_as(denom_increment))
    max_radii2D.copy_(gathered[:, 2].max(0).values.view_as(max_radii2D).to(max_radii2D.dtype))


def densification_increments(viewspace_grad, visibility_filter, radii):
    """Local contributions of one render to the statistics of scene/gaussian_model.py:405-407 and
    train.py:319: (|grad[:, :2]| masked, visibility as float, radii masked)."""
    vis = visibility_filter
    gn = torch.zeros(viewspace_grad.shape[0], 1, device=viewspace_grad.device)
    gn[vis] = torch.norm(viewspace_grad[vis, :2], dim=-1, keepdim=True)
    den = vis.float().unsqueeze(1)
    rad = torch.where(vis, radii.float(), torch.zeros_like(radii, dtype=torch.float32))
    return gn, den, rad
