"""The training loop of the reference's train.py (training(), :45-335) over the fused B200 path:
per iteration  update_learning_rate -> (every 1000 its) oneupSHdegree -> pick a view index ->
event iteration (3 renders + event loss + backward, train.py:144-212) or frame iteration (L1 +
SSIM, :213-296) -> densification statistics / densify_and_prune / reset_opacity (:317-327) ->
Adam steps (:330-332, optimizer_c :212) — with every stage a single call into libevent3dgs.

Same schedule constants and quirks as the reference: OptimizationParams defaults
(arguments/__init__.py:76-93), event mode forces opacity_reset_interval = 10000 (train.py:126),
index = randint(2, len-4) in event mode and randint(2, len-3) otherwise (:125,131), the held-out
indices 5,25,45,65,85 are replaced by index-1 (:137-139), size_threshold = 20 once iteration >
opacity_reset_interval (:323).

Multi-GPU (the reference has none): `pairs_per_step` view-pairs per optimizer step,
pair b on rank b mod world; gradients averaged with ONE all-reduce over the flat gradient buffer;
the densification statistics are reduced (SUM, SUM, MAX) before they are applied; every rank draws
the same split samples (generator seeded from the iteration), so the replicas stay identical.

No CPU fallback: everything here needs the CUDA library.
"""
import random
from types import SimpleNamespace

import numpy as np
import torch

from . import densify as dn
from . import fused, parallel
from .optim import FusedAdam

HELD_OUT = (5, 25, 45, 65, 85)


def optimization_defaults(**over):
    """OptimizationParams of arguments/__init__.py:76-93."""
    d = dict(iterations=30_000, position_lr_init=0.00016, position_lr_final=0.0000016, position_lr_delay_mult=0.01,
             position_lr_max_steps=30_000, feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001,
             percent_dense=0.01, lambda_dssim=0.2, densification_interval=100, opacity_reset_interval=3000,
             densify_from_iter=500, densify_until_iter=15_000, densify_grad_threshold=0.0002, random_background=False)
    d.update(over)
    return SimpleNamespace(**d)


def expon_lr(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """get_expon_lr_func of utils/general_utils.py:38-62, as a plain function of the step."""
    if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
        return 0.0
    if lr_delay_steps > 0:
        delay = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
    else:
        delay = 1.0
    t = np.clip(step / max_steps, 0, 1)
    return float(delay * np.exp(np.log(lr_init) * (1 - t) + np.log(lr_final) * t))


def training_setup(pc, opt, spatial_lr_scale=1.0):
    """GaussianModel.training_setup (scene/gaussian_model.py:149-175): six named groups, lr 0 default, eps 1e-15."""
    groups = [
        {"params": [pc._xyz], "lr": opt.position_lr_init * spatial_lr_scale, "name": "xyz"},
        {"params": [pc._features_dc], "lr": opt.feature_lr, "name": "f_dc"},
        {"params": [pc._features_rest], "lr": opt.feature_lr / 20.0, "name": "f_rest"},
        {"params": [pc._opacity], "lr": opt.opacity_lr, "name": "opacity"},
        {"params": [pc._scaling], "lr": opt.scaling_lr, "name": "scaling"},
        {"params": [pc._rotation], "lr": opt.rotation_lr, "name": "rotation"},
    ]
    return FusedAdam(groups, lr=0.0, eps=1e-15)


def pick_index(rng, n_views, event):
    """train.py:123-139."""
    index = rng.randint(2, n_views - 4) if event else rng.randint(2, n_views - 3)
    if index in HELD_OUT:
        index -= 1
    return index


class Trainer:
    """cams_train / cams_event: lists of cameras (scene.getTrainCameras / getEventCameras);
    gt_for(index) -> the ground-truth frames of that index, for event mode the dict
    {"intensity", "event_now", "event_next"[, "blurry"]} of [3,H,W] device tensors (float32 or uint8),
    otherwise one [3,H,W] frame."""

    def __init__(self, pc, cams_train, cams_event, gt_for, opt=None, event=True, gray=True, deblur=False, extent=1.0,
                 spatial_lr_scale=None, white_background=False, pairs_per_step=1, seed=0, rank=0, world=1, fused_exchange=True, dedupe_views=True,
                 lazy_check=True, sharded_optimizer=False):
        self.pc, self.cams_train, self.cams_event, self.gt_for = pc, cams_train, cams_event, gt_for
        self.opt = opt or optimization_defaults()
        self.event, self.gray, self.deblur = event, gray, deblur
        self.extent = float(extent)
        self.white_background = white_background
        self.rank, self.world = rank, world
        self.fused_exchange = fused_exchange
        # render the intensity view and the first event view once whenever they are the same camera (they are,
        # in the reference's datasets): loss, gradients and densification statistics are those of the
        # three-render iteration (fused.event_iteration_backward, dedupe_views); False forces three renders
        self.dedupe_views = dedupe_views
        if pairs_per_step % world:
            raise ValueError("pairs_per_step must be a multiple of the number of ranks")
        self.pairs_per_step = pairs_per_step
        self.my_pairs = parallel.shard_pairs(pairs_per_step, rank, world)
        dev = pc._xyz.device
        self.dev = dev
        self.bg = torch.tensor([1.0, 1.0, 1.0] if white_background else [0.0, 0.0, 0.0], device=dev)
        self.spatial_lr_scale = self.extent if spatial_lr_scale is None else spatial_lr_scale
        self.optimizer = training_setup(pc, self.opt, self.spatial_lr_scale)
        self.c = torch.nn.Parameter(torch.tensor(0.17, device=dev))          # train.py:71-73
        self.optimizer_c = FusedAdam([self.c], lr=0.1)
        self.state = dn.DensificationState(pc._xyz.shape[0], dev)
        self.rng = random.Random(seed)           # shared by all ranks: the global batch does not depend on world
        self.seed = seed
        self.iteration = 0
        self.rest_skipped = 0                    # Adam steps of f_rest that were pure no-ops (see _step_optimizers)
        # world > 1: gradient exchange and Adam as one kernel per rank (parallel.ShardedAdam) on the iterations that
        # neither densify nor reset the opacity; the others take the replicated route on synchronised moments
        self.sharded_optimizer = bool(sharded_optimizer) and world > 1 and event
        self.sharded = None
        self.sharded_steps = 0
        self.lazy_check = lazy_check             # fused.event_iteration_backward(lazy_check=...): host up to two steps ahead
        if fused_exchange == "push" and world > 1 and self.opt.densify_until_iter > 0:
            # multimem.red applies every rank's additions to each copy in arrival order: the copies of the gradient
            # agree to rounding, not bit for bit, and replicas that densify on their own copy may then disagree on the
            # new row count (and hang in the next collective). "pull" (one in-switch sum, broadcast) and NCCL give
            # identical copies.
            raise ValueError("fused_exchange='push' leaves the replicas' gradients equal only to fp32 rounding; it cannot be "
                             "combined with densification (set opt.densify_until_iter = 0, or use 'pull')")
        self.late_overflows = 0
        if event:
            self.opt.opacity_reset_interval = 10000          # train.py:126
        self.last = {}

    # ---- one iteration = train.py:95-332 ----
    def step(self):
        pc, opt = self.pc, self.opt
        self.iteration += 1
        it = self.iteration
        for g in self.optimizer.param_groups:                 # update_learning_rate, gaussian_model.py:177-183
            if g["name"] == "xyz":
                g["lr"] = expon_lr(it, opt.position_lr_init * self.spatial_lr_scale, opt.position_lr_final * self.spatial_lr_scale,
                                   0, opt.position_lr_delay_mult, opt.position_lr_max_steps)
        if it % 1000 == 0 and pc.active_sh_degree < pc.max_sh_degree:
            pc.active_sh_degree += 1
        densify_now = it < opt.densify_until_iter and it > opt.densify_from_iter and it % opt.densification_interval == 0
        reset_now = it < opt.densify_until_iter and (it % opt.opacity_reset_interval == 0 or (self.white_background and it == opt.densify_from_iter))
        if self.sharded_optimizer:
            try:
                if self.sharded is None:
                    self.sharded = parallel.ShardedAdam(self.dev)
                if not self.sharded.owns(pc):        # first step, and after every densification (collective)
                    self.sharded.sync_moments([self.optimizer, self.optimizer_c])
                    self.c = self.sharded.adopt(pc, self.c, self.optimizer, self.optimizer_c)
            except parallel.NoMulticast as e:
                import warnings
                warnings.warn("event_3dgs_b200.trainer: %s; replicated optimizer instead" % e)
                self.sharded_optimizer, self.sharded = False, None
        n_views = len(self.cams_train)
        indices = [pick_index(self.rng, n_views, self.event) for _ in range(self.pairs_per_step)]
        bg = torch.rand(3, device=self.dev) if opt.random_background else self.bg
        loss_sum, pkg0 = None, None
        # gradient exchange fused into the backward kernels (parallel.FusedGradientExchange) whenever the
        # no-sync workspace path is active on this model; otherwise local accumulation + one NCCL all-reduce
        exchange = None
        # every rank decides from the camera IT renders first; the ranks agree because all cameras of a run have one
        # image size (checked below after the backward: the gradients must have landed in the exchange buffer)
        if self.world >= (2 if self.sharded_optimizer else 4) and self.event and self.fused_exchange and fused._workspace_for(pc, self.cams_train[indices[self.my_pairs[0]]]) is not None:
            exchange = parallel.fused_exchange_for(pc, fused.grad_flat_numel(pc, int(pc.active_sh_degree)), self.dev,
                                                   mode=self.fused_exchange if isinstance(self.fused_exchange, str) else "pull")
        pc._grad_exchange = exchange
        pc._grad_scale = 1.0 / self.pairs_per_step
        for k, b in enumerate(self.my_pairs):
            index = indices[b]
            if self.event:
                call = lambda: fused.event_iteration_backward(
                    pc, self.cams_train[index], self.cams_event[index], self.cams_event[index + 1], self.gt_for(index), self.c, bg,
                    deblur=self.deblur, return_pkg=(k == 0), accumulate=k > 0, dedupe_views=self.dedupe_views,
                    lazy_check=self.lazy_check)
                try:
                    r = call()
                except fused.Workspace.Overflow:
                    # lazy_check: an iteration two steps back produced more instances than its buffers held (they
                    # have been enlarged since); its image and gradients missed the excess. Counted, not repeated.
                    self.late_overflows += 1
                    r = call()
            else:
                if k > 0:
                    raise RuntimeError("several pairs per rank and step are implemented for the event mode only")
                r = fused.frame_iteration_backward(pc, self.cams_train[index], self.gt_for(index), bg,
                                                   lambda_dssim=opt.lambda_dssim, gray=self.gray, return_pkg=True)
            if k == 0:
                loss, pkg0 = r
            else:
                loss = r
            loss_sum = loss if loss_sum is None else loss_sum + loss
        if exchange is not None and pc._grad_flat is not exchange.buf:
            raise RuntimeError("event_3dgs_b200.trainer: the backward pass did not accumulate into the exchange buffer "
                               "(a camera of another image size than the workspace?); the gradient exchange would reduce stale data")
        use_sharded = (self.sharded_optimizer and exchange is not None and exchange.mode == "pull" and not densify_now and not reset_now
                       and it < opt.iterations)
        if use_sharded:
            pass                                # exchange and Adam happen together below (ShardedAdam.step)
        elif exchange is not None and exchange.mode == "push":
            exchange.end()                      # barrier: every rank's reductions have landed in every copy
        elif exchange is not None:
            exchange.allreduce(1.0 / self.pairs_per_step)     # one in-switch all-reduce kernel, mean over all pairs
        else:
            if len(self.my_pairs) > 1:
                pc._grad_flat.div_(len(self.my_pairs))
            if self.world > 1:
                parallel.allreduce_flat(pc._grad_flat, self.world)
        self.last_exchange = "nvls sharded adam" if use_sharded else (("nvls " + exchange.mode) if exchange is not None else ("nccl all-reduce" if self.world > 1 else "none"))
        if self.sharded is not None and not use_sharded:
            self.sharded.sync_moments([self.optimizer, self.optimizer_c])       # a replicated step / state surgery follows
        self.last = {"loss": loss_sum / len(self.my_pairs), "indices": indices, "num_rendered": pkg0["num_rendered"]}

        if it < opt.densify_until_iter:                        # train.py:317-327
            self._add_stats(pkg0)
            if it > opt.densify_from_iter and it % opt.densification_interval == 0:
                size_threshold = 20 if it > opt.opacity_reset_interval else None
                gen = torch.Generator(device=self.dev).manual_seed(self.seed * 1_000_003 + it)
                # As in the reference, densification runs BEFORE optimizer.step() and replaces all six
                # parameters by new ones without .grad, so the Adam step of this iteration skips them.
                self.last["densify"] = dn.densify_and_prune(pc, self.optimizer, self.state, opt.densify_grad_threshold, 0.005,
                                                            self.extent, size_threshold, percent_dense=opt.percent_dense, generator=gen)
            if it % opt.opacity_reset_interval == 0 or (self.white_background and it == opt.densify_from_iter):
                dn.reset_opacity(pc, self.optimizer)
                pc._opacity.grad = None        # replace_tensor_to_optimizer: the new opacity tensor has no grad
        if use_sharded:
            D = int(pc.active_sh_degree)
            rest = pc._features_rest
            if D == 0 and self.optimizer.moments_are_zero(rest):
                self.optimizer.advance_without_update(rest)         # as in _step_optimizers
                self.rest_skipped += 1
            self.sharded.step(pc, self.c, self.optimizer, self.optimizer_c, exchange, 1.0 / self.pairs_per_step, D)
            self.sharded_steps += 1
        elif it < opt.iterations:
            self._step_optimizers()
        self.optimizer.zero_grad(set_to_none=True)
        self.optimizer_c.zero_grad(set_to_none=True)
        return self.last["loss"]

    def _add_stats(self, pkg):
        """Densification statistics of train.py:317-320 from the FIRST view pair of this rank's share of the step
        (the reference has one pair per step); further pairs of a rank add gradients but no statistics."""
        grad, radii = pkg["viewspace_points"].grad, pkg["radii"]
        if self.world == 1:
            dn.add_densification_stats(self.state, grad, radii)
            return
        inc = dn.DensificationState(radii.shape[0], self.dev)
        dn.add_densification_stats(inc, grad, radii)
        parallel.allreduce_densification_stats(inc.xyz_gradient_accum, inc.denom, inc.max_radii2D)
        self.state.xyz_gradient_accum.add_(inc.xyz_gradient_accum)
        self.state.denom.add_(inc.denom)
        torch.maximum(self.state.max_radii2D, inc.max_radii2D, out=self.state.max_radii2D)

    def _step_optimizers(self):
        pc = self.pc
        rest = pc._features_rest
        # While active_sh_degree == 0 the gradient of features_rest is identically zero and so are its
        # two moments: the Adam update of that group (76 % of the optimizer's bytes) is the identity.
        # Only its step counter advances, so that the bias corrections agree with the reference's
        # once the degree is raised.
        if rest.grad is not None and pc.active_sh_degree == 0 and self.optimizer.moments_are_zero(rest):
            rest.grad = None
            self.optimizer.advance_without_update(rest)
            self.rest_skipped += 1
        self.optimizer.step()
        if self.event:
            self.optimizer_c.step()

    def train(self, iterations, log_every=0):
        losses = []
        for _ in range(iterations):
            loss = self.step()
            if log_every and self.iteration % log_every == 0:
                losses.append((self.iteration, float(loss)))
        return losses
