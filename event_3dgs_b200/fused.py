"""Fused event iteration: the reference's train.py:144-212 hot section (3 renders + event loss +
loss.backward()) with the PyTorch glue folded into the CUDA kernels.

  * activations + SH -> colour inside the preprocess kernels (e3_rasterize_forward_raw /
    e3_rasterize_backward_raw) instead of ~60 ATen launches per render and their autograd backward;
  * the event loss and its three backward seeds in two streaming kernels (e3_event_loss) instead
    of ~50 launches; the dead ssim_gray evaluation (lambda_dssim = 0, train.py:177-178) is gone;
  * the three backward passes add into ONE set of gradient tensors (no autograd graph at all).

The result is what `loss.backward()` leaves in the reference: .grad on the six leaf tensors and on
the event threshold c, plus the first render's screen-space gradient for the densification
statistics (train.py:317-320). The drop-in path (pipeline.py) stays the reference-shaped one.
"""
import ctypes
import math

import torch

from . import _lib
from .rasterizer import _ALLOC_CB, _slots


class RenderState:
    __slots__ = ("R", "color", "radii", "geom", "binning", "img", "cam", "tanfovx", "tanfovy", "channels")


class Workspace:
    """Caller-owned state buffers of the no-sync path (e3_rasterize_forward_raw_ws): one slot per
    render that is alive at the same time (3 per event iteration), kept for the whole run. The
    binning buffers hold `cap` instances; num_rendered of every forward arrives asynchronously in
    pinned host memory and is checked against `cap` once per iteration (check())."""

    def __init__(self, P, W, H, device, cap, slots=3):
        L = _lib.lib()
        self.P, self.W, self.H, self.device, self.slots = P, W, H, device, slots
        u8 = dict(dtype=torch.uint8, device=device)
        self.geom = [torch.empty(L.e3_geometry_bytes(P), **u8) for _ in range(slots)]
        self.img = [torch.empty(L.e3_image_bytes(W, H), **u8) for _ in range(slots)]
        self.r_host = torch.zeros(slots, 2, dtype=torch.int32).pin_memory()
        self.events = [torch.cuda.Event() for _ in range(slots)]
        self.used = [False] * slots
        # single-call iteration (e3_event_iteration): seeds, loss scratch, and -- twice, used alternately -- the
        # pinned words the forwards' num_rendered arrive in and one event after the forwards
        self.seeds = {}
        self.loss_scratch = torch.empty(8, dtype=torch.float64, device=device)
        self.r_host2 = torch.zeros(2, 3, 2, dtype=torch.int32).pin_memory()
        self.fw_events = [torch.cuda.Event(), torch.cuda.Event()]
        with torch.cuda.device(device):
            for e in self.fw_events:
                e.record(torch.cuda.current_stream(device))      # creates the underlying cudaEvent_t
        self.args = _lib.EventIterationArgs()
        self.parity = 0
        self.pending = [False, False]
        self.cap_used = [0, 0]           # capacity the iteration of each buffer set was enqueued with
        self.last_Rs = None
        self.overflows = 0
        self.resize(cap)

    def resize(self, cap):
        L = _lib.lib()
        self.cap = int(max(cap, 4096))
        self.bin = [torch.empty(L.e3_binning_bytes(self.cap), dtype=torch.uint8, device=self.device) for _ in range(self.slots)]

    def seed(self, k, channels):
        t = self.seeds.get((k, channels))
        if t is None:
            t = self.seeds[(k, channels)] = torch.empty(channels, self.H, self.W, dtype=torch.float32, device=self.device)
        return t

    def matches(self, P, W, H, device):
        return (self.P, self.W, self.H, self.device) == (P, W, H, device)

    def check(self):
        """Wait for the R read-backs of the slots used since the last check (they precede the loss
        and backward kernels in the stream, so this does not drain the pipeline). Returns the
        measured R per slot; raises Overflow after enlarging the buffers if one exceeded cap."""
        Rs, worst = [], 0
        for k in range(self.slots):
            if self.used[k]:
                self.events[k].synchronize()
                Rs.append(int(self.r_host[k, 0]))
                self.used[k] = False
        worst = max(Rs) if Rs else 0
        if worst > self.cap:
            self.resize(int(worst * 1.5))
            raise Workspace.Overflow(worst)
        return Rs

    def harvest(self, par):
        """Single-call iterations: wait for the forwards of the iteration that used buffer set `par` (they precede
        its loss and backward kernels in the stream) and read their num_rendered. Raises Overflow after enlarging
        the binning buffers if one exceeded the capacity, and enlarges them ahead of need above 85 % of it."""
        self.fw_events[par].synchronize()
        self.pending[par] = False
        Rs = [int(self.r_host2[par, k, 0]) for k in range(3)]
        self.last_Rs = Rs
        worst = max(Rs)
        if worst > self.cap_used[par]:
            self.overflows += 1
            if worst > self.cap:
                self.resize(int(worst * 1.5))
            raise Workspace.Overflow(worst)
        if worst > 0.85 * self.cap:
            self.resize(int(worst * 1.3) + 4096)
        return Rs

    def drain(self):
        """Harvest every outstanding single-call iteration (call it before trusting the results of a run of
        lazily checked iterations). Raises Overflow if one of them exceeded the capacity."""
        err = None
        for par in (self.parity, self.parity ^ 1):
            if self.pending[par]:
                try:
                    self.harvest(par)
                except Workspace.Overflow as ex:
                    err = ex
        if err is not None:
            raise err
        return self.last_Rs

    class Overflow(RuntimeError):
        pass


def _forward_raw(pc, cam, bg, scaling_modifier=1.0, debug=False, channels=3, ws=None, slot=0):
    if ws is not None:
        return _forward_raw_ws(pc, cam, bg, scaling_modifier, debug, channels, ws, slot)
    L = _lib.lib()
    xyz = pc._xyz
    P = xyz.shape[0]
    dev = xyz.device
    H, W = int(cam.image_height), int(cam.image_width)
    st = RenderState()
    st.cam = cam
    st.tanfovx = math.tan(cam.FoVx * 0.5)
    st.tanfovy = math.tan(cam.FoVy * 0.5)
    st.radii = torch.empty(P, dtype=torch.int32, device=dev)
    st.channels = channels
    empty = torch.empty(0, dtype=torch.uint8, device=dev)
    if P == 0:
        st.R, st.color, st.geom, st.binning, st.img = 0, torch.zeros(channels, H, W, device=dev), empty, empty, empty
        return st
    st.color = torch.empty(channels, H, W, dtype=torch.float32, device=dev)
    D = int(pc.active_sh_degree)
    R = ctypes.c_int(0)
    _slots.clear()
    _slots["device"] = dev
    with torch.cuda.device(dev):
        code = L.e3_rasterize_forward_raw(
            _ALLOC_CB, 1, _ALLOC_CB, 2, _ALLOC_CB, 3, P, D, channels, bg.data_ptr(), W, H, xyz.data_ptr(), pc._features_dc.data_ptr(),
            pc._features_rest.data_ptr(), pc._opacity.data_ptr(), pc._scaling.data_ptr(), float(scaling_modifier),
            pc._rotation.data_ptr(), cam.world_view_transform.data_ptr(), cam.full_proj_transform.data_ptr(),
            cam.camera_center.data_ptr(), st.tanfovx, st.tanfovy, st.color.data_ptr(), st.radii.data_ptr(), int(debug),
            torch.cuda.current_stream(dev).cuda_stream, ctypes.byref(R))
    st.geom, st.binning, st.img = _slots.get(1, empty), _slots.get(2, empty), _slots.get(3, empty)
    _slots.clear()
    _lib.check(code)
    st.R = R.value
    return st


def _forward_raw_ws(pc, cam, bg, scaling_modifier, debug, channels, ws, slot):
    L = _lib.lib()
    xyz = pc._xyz
    P = xyz.shape[0]
    dev = xyz.device
    H, W = int(cam.image_height), int(cam.image_width)
    st = RenderState()
    st.cam, st.channels = cam, channels
    st.tanfovx = math.tan(cam.FoVx * 0.5)
    st.tanfovy = math.tan(cam.FoVy * 0.5)
    st.radii = torch.empty(P, dtype=torch.int32, device=dev)
    st.color = torch.empty(channels, H, W, dtype=torch.float32, device=dev)
    st.geom, st.binning, st.img, st.R = ws.geom[slot], ws.bin[slot], ws.img[slot], ws.cap
    stream = torch.cuda.current_stream(dev)
    with torch.cuda.device(dev):
        code = L.e3_rasterize_forward_raw_ws(
            st.geom.data_ptr(), st.binning.data_ptr(), ws.cap, st.img.data_ptr(), ws.r_host[slot].data_ptr(), P,
            int(pc.active_sh_degree), channels, bg.data_ptr(), W, H, xyz.data_ptr(), pc._features_dc.data_ptr(),
            pc._features_rest.data_ptr(), pc._opacity.data_ptr(), pc._scaling.data_ptr(), float(scaling_modifier),
            pc._rotation.data_ptr(), cam.world_view_transform.data_ptr(), cam.full_proj_transform.data_ptr(),
            cam.camera_center.data_ptr(), st.tanfovx, st.tanfovy, st.color.data_ptr(), st.radii.data_ptr(), int(debug),
            stream.cuda_stream)
    _lib.check(code)
    ws.events[slot].record(stream)
    ws.used[slot] = True
    return st


def _backward_raw(pc, st, bg, dL_dpix, grads, accumulate, d_mean2D=None, scaling_modifier=1.0, debug=False, stats_seed=None):
    """stats_seed: second seed whose screen-space gradient alone is written to d_mean2D (e3_rasterize_backward_raw_dual)."""
    L = _lib.lib()
    xyz = pc._xyz
    P = xyz.shape[0]
    if P == 0:
        return
    dev = xyz.device
    cam = st.cam
    D = int(pc.active_sh_degree)
    fn = L.e3_rasterize_backward_raw if stats_seed is None else L.e3_rasterize_backward_raw_dual
    seeds = (dL_dpix.data_ptr(),) if stats_seed is None else (dL_dpix.data_ptr(), stats_seed.data_ptr())
    with torch.cuda.device(dev):
        code = fn(
            P, D, st.channels, st.R, bg.data_ptr(), int(cam.image_width), int(cam.image_height), xyz.data_ptr(),
            pc._features_dc.data_ptr(), pc._features_rest.data_ptr(), pc._opacity.data_ptr(), pc._scaling.data_ptr(),
            float(scaling_modifier), pc._rotation.data_ptr(), cam.world_view_transform.data_ptr(),
            cam.full_proj_transform.data_ptr(), cam.camera_center.data_ptr(), st.tanfovx, st.tanfovy, st.radii.data_ptr(),
            st.geom.data_ptr(), st.binning.data_ptr() if st.R > 0 else None, st.img.data_ptr(), *seeds,
            int(bool(accumulate)), grads["xyz"].data_ptr(), grads["features_dc"].data_ptr(),
            grads["features_rest"].data_ptr() if D > 0 else None, grads["opacity"].data_ptr(), grads["scaling"].data_ptr(),
            grads["rotation"].data_ptr(), None if d_mean2D is None else d_mean2D.data_ptr(), int(debug),
            torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(code)


def event_loss_and_seeds(image, image_now, image_next, gt, c, deblur=False, upstream=1.0):
    """e3_event_loss: returns (out5 tensor [loss, dL/dc, loss1, loss2, f], seeds dict). image_now and
    image_next are both [3,H,W] or both [1,H,W] (luma renders)."""
    L = _lib.lib()
    dev = image.device
    _, H, W = image.shape
    ec = int(image_now.shape[0])
    if ec != int(image_next.shape[0]) or ec not in (1, 3):
        raise RuntimeError("event_loss_and_seeds: the two event renders must both have 1 or both have 3 channels")
    seeds = {"image": torch.empty_like(image), "image_now": torch.empty_like(image_now), "image_next": torch.empty_like(image_next)}
    out = torch.empty(5, dtype=torch.float32, device=dev)
    scratch = torch.empty(8, dtype=torch.float64, device=dev)
    blur = gt["blurry"] if deblur else None
    frames = [gt["event_now"], gt["event_next"], gt["intensity"]] + ([blur] if blur is not None else [])
    u8 = frames[0].dtype == torch.uint8
    for f in frames:
        if f.dtype != (torch.uint8 if u8 else torch.float32) or not f.is_contiguous() or tuple(f.shape) != (3, H, W):
            raise RuntimeError("event_loss_and_seeds: GT frames must be contiguous [3,H,W], all float32 or all uint8")
    with torch.cuda.device(dev):
        code = L.e3_event_loss(W, H, image.data_ptr(), image_now.data_ptr(), image_next.data_ptr(),
                               gt["event_now"].data_ptr(), gt["event_next"].data_ptr(), gt["intensity"].data_ptr(),
                               None if blur is None else blur.data_ptr(), int(u8), c.data_ptr(), float(upstream), ec,
                               scratch.data_ptr(), 64,
                               seeds["image"].data_ptr(), seeds["image_now"].data_ptr(), seeds["image_next"].data_ptr(),
                               out.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(code)
    return out, seeds


def same_view(cam_a, cam_b):
    """True when two camera objects describe the same pose, intrinsics and image size, judged from
    HOST-side data only (no device sync). The reference builds its train, event and blurry camera
    lists from the same COLMAP extrinsics/intrinsics (scene/dataset_readers.py:145-160), so
    train_cam[i] and event_cam[i] are the same view and train.py renders it twice per iteration."""
    if cam_a is cam_b:
        return True
    try:
        import numpy as np
        if cam_a.R is None or cam_a.T is None or cam_b.R is None or cam_b.T is None:
            return False
        return (int(cam_a.image_width) == int(cam_b.image_width) and int(cam_a.image_height) == int(cam_b.image_height) and
                float(cam_a.FoVx) == float(cam_b.FoVx) and float(cam_a.FoVy) == float(cam_b.FoVy) and
                np.array_equal(np.asarray(cam_a.R), np.asarray(cam_b.R)) and
                np.array_equal(np.asarray(cam_a.T), np.asarray(cam_b.T)) and
                np.array_equal(np.asarray(getattr(cam_a, "trans", 0.0)), np.asarray(getattr(cam_b, "trans", 0.0))) and
                float(getattr(cam_a, "scale", 1.0)) == float(getattr(cam_b, "scale", 1.0)))
    except AttributeError:
        return False


def event_iteration_backward(pc, cam_train, cam_now, cam_next, gt, c, bg, deblur=False, return_pkg=False,
                             dedupe_views=False, gray_event_views=True, gt_ready=None, accumulate=False, no_sync=True,
                             lazy_check=False):
    """train.py:144-212 up to and including loss.backward(): sets .grad on pc's six leaves and on c,
    returns the loss (0-dim device tensor). With return_pkg also the first render's package
    (render, viewspace_points with .grad, visibility_filter, radii) that train.py:317-320 consumes.

    dedupe_views (opt-in): when the intensity camera and the first event camera are the same view
    (see same_view; they are in the reference's datasets) it is rendered ONCE. Its backward blend carries
    two seeds (e3_rasterize_backward_raw_dual): the sum of the intensity and the event term for the
    parameter gradients, the intensity term alone for the screen-space gradient handed to the
    densification statistics -- so loss, parameter gradients and statistics all equal the three-render
    iteration up to fp32 summation order. Off by default only because the headline metric counts three
    renders per view pair.

    gray_event_views: render the two event views with one channel, the luma that
    differentialable_event_simu takes of them anyway (utils/loss_utils.py:24-28,240-241); the
    algebraic rewrite behind the reference's --gray idea. Equal to the
    3-channel route up to fp32 rounding; False reproduces the 3-channel renders exactly.

    gt_ready: optional torch.cuda.Event recorded after the GT frames were copied to the device on
    another stream; the loss waits for it, so the host-to-device copy overlaps the three renders.

    no_sync: after the first call (which reads num_rendered back like the reference and sizes a
    persistent Workspace from it) no forward blocks the host any more; see Workspace.

    lazy_check: the host does not wait for this iteration's num_rendered at all; they are read (and compared
    with the workspace capacity) two calls later, when their buffer set comes up for reuse, so the host runs up to
    two iterations ahead of the GPU (at event-camera resolutions an iteration is shorter than its enqueue).
    An iteration that exceeded the capacity is then REPORTED late -- Workspace.Overflow from a later call or from
    Workspace.drain(), after the buffers were enlarged -- instead of being repeated; the capacity is kept 30 % above
    the largest count seen and grown ahead of need, so this is an exceptional event. Default off: the iteration
    waits for its own forwards (not for its backward) and repeats itself on overflow.

    accumulate: add this view-pair's gradients to the ones the previous call left (several pairs per
    optimizer step, e.g. BASELINE config c5 with 64 pairs per step); the returned loss is this
    pair's. The caller divides by the number of pairs (or lets parallel.allreduce_flat average over
    ranks and scales the rest)."""
    if accumulate and getattr(pc, "_grad_flat", None) is None:
        raise RuntimeError("accumulate=True needs the gradients of a previous call")
    for t in (pc._xyz, pc._features_dc, pc._features_rest, pc._scaling, pc._rotation, pc._opacity, c, bg):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError("event_3dgs_b200.fused: tensors must be contiguous float32 CUDA tensors")
    if dedupe_views and same_view(cam_train, cam_now):
        ws = _workspace_for(pc, cam_train) if no_sync and pc._xyz.shape[0] > 0 else None
        if ws is not None:
            return _event_iteration_single_call(pc, (cam_train, cam_now, cam_next), gt, c, bg, deblur, return_pkg, 1, gt_ready,
                                                accumulate, ws, shared=True, lazy=lazy_check)
        if accumulate:
            raise RuntimeError("accumulate=True with dedupe_views needs the workspace of a previous call")
        return _event_iteration_two_renders(pc, cam_train, cam_next, gt, c, bg, deblur, return_pkg, gt_ready, no_sync)
    ec = 1 if gray_event_views else 3
    ws = _workspace_for(pc, cam_train) if no_sync and pc._xyz.shape[0] > 0 else None
    if ws is not None:
        return _event_iteration_single_call(pc, (cam_train, cam_now, cam_next), gt, c, bg, deblur, return_pkg, ec, gt_ready,
                                            accumulate, ws, lazy=lazy_check)
    sts = [_forward_raw(pc, cam_train, bg, ws=ws, slot=0), _forward_raw(pc, cam_now, bg, channels=ec, ws=ws, slot=1),
           _forward_raw(pc, cam_next, bg, channels=ec, ws=ws, slot=2)]
    if ws is None and no_sync and pc._xyz.shape[0] > 0:
        _remember_workspace(pc, cam_train, max(s.R for s in sts))
    if gt_ready is not None:
        torch.cuda.current_stream(pc._xyz.device).wait_event(gt_ready)
    out, seeds = event_loss_and_seeds(sts[0].color, sts[1].color, sts[2].color, gt, c, deblur)
    D = int(pc.active_sh_degree)
    grads = pc._grad_views if accumulate else _alloc_grads(pc, c, D)
    d_mean2D = None if accumulate else torch.empty_like(pc._xyz)
    _backward_raw(pc, sts[0], bg, seeds["image"], grads, accumulate, d_mean2D)
    _backward_raw(pc, sts[1], bg, seeds["image_now"], grads, True)
    _backward_raw(pc, sts[2], bg, seeds["image_next"], grads, True)
    if ws is not None:
        try:
            Rs = ws.check()
        except Workspace.Overflow:
            # more instances than the workspace was sized for: it has been enlarged, redo the step
            if accumulate:
                raise RuntimeError("event_3dgs_b200.fused: workspace overflow during an accumulating step; repeat the optimizer step")
            return event_iteration_backward(pc, cam_train, cam_now, cam_next, gt, c, bg, deblur, return_pkg, dedupe_views,
                                            gray_event_views, None, accumulate, no_sync)
        for s_, r_ in zip(sts, Rs):
            s_.R = r_
    if accumulate:
        grads["c"].add_(out[1:2])
        return out[0]
    pc._xyz.grad = grads["xyz"]
    pc._features_dc.grad = grads["features_dc"]
    pc._features_rest.grad = grads["features_rest"]
    pc._opacity.grad = grads["opacity"]
    pc._scaling.grad = grads["scaling"]
    pc._rotation.grad = grads["rotation"]
    _finish_c_grad(c, grads, out)
    loss = out[0]
    if return_pkg:
        vsp = torch.zeros_like(pc._xyz)
        vsp.grad = d_mean2D
        pkg = {"render": sts[0].color, "viewspace_points": vsp, "visibility_filter": sts[0].radii > 0,
               "radii": sts[0].radii, "num_rendered": [s.R for s in sts], "loss_terms": out}
        return loss, pkg
    return loss


def _event_iteration_single_call(pc, cams3, gt, c, bg, deblur, return_pkg, ec, gt_ready, accumulate, ws, shared=False, lazy=False):
    """The whole iteration through ONE C call (e3_event_iteration) on the persistent workspace: no per-render
    Python work, and the only host wait is for the forwards' num_rendered (none at all with lazy=True, see
    event_iteration_backward). Falls back to itself once if the workspace overflowed (lazy=False)."""
    L = _lib.lib()
    if lazy:
        if ws.pending[ws.parity]:
            ws.harvest(ws.parity)          # the iteration before the previous one; its forwards finished long ago
    else:
        ws.drain()
    par = ws.parity
    dev = pc._xyz.device
    P, D = pc._xyz.shape[0], int(pc.active_sh_degree)
    H, W = ws.H, ws.W
    blur = gt["blurry"] if deblur else None
    frames = [gt["event_now"], gt["event_next"], gt["intensity"]] + ([blur] if blur is not None else [])
    u8 = frames[0].dtype == torch.uint8
    for f in frames:
        if f.dtype != (torch.uint8 if u8 else torch.float32) or not f.is_contiguous() or tuple(f.shape) != (3, H, W) or f.device != dev:
            raise RuntimeError("event_3dgs_b200.fused: GT frames must be contiguous [3,H,W] on the model's device, all float32 or all uint8")
    ex = getattr(pc, "_grad_exchange", None)        # parallel.FusedGradientExchange of the multi-GPU trainer, or None
    mc = ex if (ex is not None and ex.mode == "push") else None
    if mc is not None:
        if not accumulate:
            mc.begin()                                # zero my copy; barrier: every rank's copy is zeroed
            grads = _alloc_grads(pc, c, D, flat=mc.buf)
        else:
            grads = pc._grad_views
    elif ex is not None:                              # "pull": accumulate locally, but in the symmetric buffer
        grads = pc._grad_views if accumulate else _alloc_grads(pc, c, D, flat=ex.buf)
    else:
        grads = pc._grad_views if accumulate else _alloc_grads(pc, c, D)
    # shared: the first event view is the intensity view (same camera): it is rendered once (channels 0 =
    # "not rendered"), the next view as luma; the backward blend of the shared view carries two seeds
    chans = (3, 0, 1) if shared else (3, ec, ec)
    colors = [torch.empty(ch, H, W, dtype=torch.float32, device=dev) if ch else None for ch in chans]
    radii = [torch.empty(P, dtype=torch.int32, device=dev) if ch else None for ch in chans]
    out = torch.empty(5, dtype=torch.float32, device=dev)
    d_mean2D = None if (accumulate and not shared) else torch.empty_like(pc._xyz)
    if shared:
        ws.r_host2[par, 1].zero_()
    a = ws.args
    a.P, a.active_sh_degree, a.width, a.height = P, D, W, H
    a.binning_capacity, a.scale_modifier, a.accumulate = ws.cap, 1.0, int(bool(accumulate) or mc is not None)
    a.grad_multicast_delta = mc.delta if mc is not None else 0
    a.loss_upstream = float(getattr(pc, "_grad_scale", 1.0)) if mc is not None else 1.0
    a.background = bg.data_ptr()
    a.xyz, a.features_dc, a.features_rest = pc._xyz.data_ptr(), pc._features_dc.data_ptr(), pc._features_rest.data_ptr()
    a.opacity_logits, a.log_scales, a.raw_rotations = pc._opacity.data_ptr(), pc._scaling.data_ptr(), pc._rotation.data_ptr()
    a.d_xyz, a.d_features_dc = grads["xyz"].data_ptr(), grads["features_dc"].data_ptr()
    a.d_features_rest = grads["features_rest"].data_ptr() if D > 0 else None
    a.d_opacity_logits, a.d_log_scales, a.d_raw_rotations = grads["opacity"].data_ptr(), grads["scaling"].data_ptr(), grads["rotation"].data_ptr()
    a.d_mean2D = None if d_mean2D is None else d_mean2D.data_ptr()
    for k, cam in enumerate(cams3):
        v = a.views[k]
        v.viewmatrix, v.projmatrix, v.cam_pos = cam.world_view_transform.data_ptr(), cam.full_proj_transform.data_ptr(), cam.camera_center.data_ptr()
        v.tan_fovx, v.tan_fovy, v.channels = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), chans[k]
        v.geometry_buffer, v.binning_buffer, v.image_buffer = ws.geom[k].data_ptr(), ws.bin[k].data_ptr(), ws.img[k].data_ptr()
        if chans[k]:
            v.out_color, v.radii, v.seed = colors[k].data_ptr(), radii[k].data_ptr(), ws.seed(k, chans[k]).data_ptr()
        else:       # only the seed buffer of the skipped view is used: the total seed of the shared view
            v.out_color, v.radii, v.seed = None, None, ws.seed(k, 3).data_ptr()
        v.num_rendered_host = ws.r_host2[par, k].data_ptr()
    a.gt_now, a.gt_next, a.gt_intensity = gt["event_now"].data_ptr(), gt["event_next"].data_ptr(), gt["intensity"].data_ptr()
    a.gt_blur = None if blur is None else blur.data_ptr()
    a.gt_is_u8, a.c, a.out5, a.loss_scratch = int(u8), c.data_ptr(), out.data_ptr(), ws.loss_scratch.data_ptr()
    a.forwards_done_event = ws.fw_events[par].cuda_event
    stream = torch.cuda.current_stream(dev)
    a.wait_before_loss_event = gt_ready.cuda_event if gt_ready is not None else None
    with torch.cuda.device(dev):
        code = L.e3_event_iteration(ctypes.byref(a), stream.cuda_stream)
    _lib.check(code)
    ws.pending[par] = True
    ws.cap_used[par] = ws.cap
    ws.parity = par ^ 1
    try:
        Rs = (ws.last_Rs or [0, 0, 0]) if lazy else ws.harvest(par)
    except Workspace.Overflow:
        if mc is not None:
            raise RuntimeError("event_3dgs_b200.fused: workspace overflow during a step whose gradients were already reduced "
                               "into the other ranks; enlarge the workspace (Workspace.resize) and repeat the optimizer step")
        if accumulate:
            raise RuntimeError("event_3dgs_b200.fused: workspace overflow during an accumulating step; repeat the optimizer step")
        return _event_iteration_single_call(pc, cams3, gt, c, bg, deblur, return_pkg, ec, None, accumulate, ws, shared)
    if shared:
        Rs = [Rs[0], Rs[2]]
    if mc is not None:
        # dL/dc comes from the loss kernel (already scaled by loss_upstream): reduce it into every rank's copy too
        with torch.cuda.device(dev):
            _lib.check(L.e3_multimem_add(mc.multicast_ptr_of(grads["c"]), out[1:2].data_ptr(), 1, 1.0, stream.cuda_stream))
        if accumulate:
            return out[0]
    elif accumulate:
        grads["c"].add_(out[1:2])
        return out[0]
    pc._xyz.grad = grads["xyz"]
    pc._features_dc.grad = grads["features_dc"]
    pc._features_rest.grad = grads["features_rest"]
    pc._opacity.grad = grads["opacity"]
    pc._scaling.grad = grads["scaling"]
    pc._rotation.grad = grads["rotation"]
    if mc is not None:
        c.grad = grads["c"].view(c.shape)
    else:
        _finish_c_grad(c, grads, out)
    if return_pkg:
        vsp = torch.zeros_like(pc._xyz)
        vsp.grad = d_mean2D
        return out[0], {"render": colors[0], "viewspace_points": vsp, "visibility_filter": radii[0] > 0, "radii": radii[0],
                        "num_rendered": Rs, "loss_terms": out}
    return out[0]


def _workspace_for(pc, cam):
    ws = getattr(pc, "_e3_ws", None)
    if ws is not None and ws.matches(pc._xyz.shape[0], int(cam.image_width), int(cam.image_height), pc._xyz.device):
        return ws
    return None


def _remember_workspace(pc, cam, r_max):
    """First (synchronous) step done: size the persistent workspace from the measured num_rendered."""
    pc._e3_ws = Workspace(pc._xyz.shape[0], int(cam.image_width), int(cam.image_height), pc._xyz.device, int(r_max * 1.3) + 4096)


def _event_iteration_two_renders(pc, cam_a, cam_next, gt, c, bg, deblur, return_pkg, gt_ready=None, no_sync=True):
    st_a = _forward_raw(pc, cam_a, bg)
    st_n = _forward_raw(pc, cam_next, bg)
    if no_sync and pc._xyz.shape[0] > 0:
        _remember_workspace(pc, cam_a, max(st_a.R, st_n.R))
    if gt_ready is not None:
        torch.cuda.current_stream(pc._xyz.device).wait_event(gt_ready)
    out, seeds = event_loss_and_seeds(st_a.color, st_a.color, st_n.color, gt, c, deblur)
    # one backward blend of the shared view with both seeds: parameter gradients from their sum, the
    # screen-space gradient (densification statistics) from the intensity term alone, as in the reference
    total = seeds["image_now"].add_(seeds["image"])
    D = int(pc.active_sh_degree)
    grads = _alloc_grads(pc, c, D)
    d_mean2D = torch.empty_like(pc._xyz)
    _backward_raw(pc, st_a, bg, total, grads, False, d_mean2D, stats_seed=seeds["image"])
    _backward_raw(pc, st_n, bg, seeds["image_next"], grads, True)
    pc._xyz.grad = grads["xyz"]
    pc._features_dc.grad = grads["features_dc"]
    pc._features_rest.grad = grads["features_rest"]
    pc._opacity.grad = grads["opacity"]
    pc._scaling.grad = grads["scaling"]
    pc._rotation.grad = grads["rotation"]
    _finish_c_grad(c, grads, out)
    if return_pkg:
        vsp = torch.zeros_like(pc._xyz)
        vsp.grad = d_mean2D
        return out[0], {"render": st_a.color, "viewspace_points": vsp, "visibility_filter": st_a.radii > 0,
                        "radii": st_a.radii, "num_rendered": [st_a.R, st_n.R], "loss_terms": out}
    return out[0]


def grad_flat_numel(pc, D):
    """Elements of the flat gradient buffer _alloc_grads lays out for this model at SH degree D."""
    ts = [pc._xyz, pc._features_dc, pc._opacity, pc._scaling, pc._rotation] + ([pc._features_rest] if D > 0 else [])
    return sum((t.numel() + 3) // 4 * 4 for t in ts) + 4


def _alloc_grads(pc, c, D, flat=None):
    """All gradients of one iteration live in ONE flat buffer (segments padded to 16 bytes), so that
    the multi-GPU trainer exchanges them with a single NCCL all-reduce (parallel.allreduce_flat);
    the per-parameter tensors handed to .grad are views into it. Left on pc as `_grad_flat`."""
    names = [("xyz", pc._xyz), ("features_dc", pc._features_dc), ("opacity", pc._opacity), ("scaling", pc._scaling),
             ("rotation", pc._rotation)]
    if D > 0:
        names.append(("features_rest", pc._features_rest))
    sizes = [(t.numel() + 3) // 4 * 4 for _, t in names]
    if flat is None:
        flat = torch.empty(sum(sizes) + 4, dtype=torch.float32, device=pc._xyz.device)
        fresh = True
    else:           # caller-owned (symmetric-memory) buffer, already zeroed
        if flat.numel() != sum(sizes) + 4:
            raise RuntimeError("gradient buffer has %d elements, the model needs %d" % (flat.numel(), sum(sizes) + 4))
        fresh = False
    grads, off = {}, 0
    for (n, t), sz in zip(names, sizes):
        grads[n] = flat[off:off + t.numel()].view(t.shape)
        if fresh and sz != t.numel():
            flat[off + t.numel():off + sz].zero_()
        off += sz
    grads["c"] = flat[off:off + 1]
    if fresh:
        flat[off + 1:off + 4].zero_()
    if D == 0:
        grads["features_rest"] = _zeros_like_cached(pc._features_rest)
    pc._grad_flat = flat
    pc._grad_views = grads
    return grads


def _finish_c_grad(c, grads, out):
    grads["c"].copy_(out[1:2])
    c.grad = grads["c"].view(c.shape)


_ZERO_CACHE = {}


def _zeros_like_cached(t):
    """At SH degree 0 the gradient of features_rest is identically zero; hand out one shared
    zero tensor instead of filling 180 bytes per Gaussian every iteration. Treat it as read-only."""
    key = (t.device, tuple(t.shape))
    z = _ZERO_CACHE.get(key)
    if z is None:
        z = torch.zeros_like(t)
        _ZERO_CACHE.clear()
        _ZERO_CACHE[key] = z
    return z


def l1_ssim_loss_and_seed(image, gt_image, lambda_dssim=0.2, gray=False, upstream=1.0):
    """e3_l1_ssim_loss: the loss of the non-event branches of train.py (:213-223 with gray=True,
    :292-296 with gray=False). Returns (out3 tensor [loss, L1, SSIM], dL/dimage [3,H,W])."""
    L = _lib.lib()
    dev = image.device
    _, H, W = image.shape
    u8 = gt_image.dtype == torch.uint8
    if tuple(image.shape) != (3, H, W) or tuple(gt_image.shape) != (3, H, W) or not image.is_contiguous() or \
            not gt_image.is_contiguous() or image.dtype != torch.float32 or (not u8 and gt_image.dtype != torch.float32):
        raise RuntimeError("l1_ssim_loss_and_seed: image float32 [3,H,W], gt float32 or uint8 [3,H,W], contiguous")
    nb = L.e3_l1_ssim_scratch_bytes(W, H, int(bool(gray)))
    scratch = torch.empty((nb + 7) // 8, dtype=torch.float64, device=dev)
    seed = torch.empty_like(image)
    out = torch.empty(3, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.e3_l1_ssim_loss(W, H, int(bool(gray)), image.data_ptr(), gt_image.data_ptr(), int(u8), float(lambda_dssim),
                                     float(upstream), scratch.data_ptr(), scratch.numel() * 8, seed.data_ptr(), out.data_ptr(),
                                     torch.cuda.current_stream(dev).cuda_stream))
    return out, seed


def frame_iteration_backward(pc, cam, gt_image, bg, lambda_dssim=0.2, gray=False, return_pkg=False, no_sync=True):
    """The non-event iteration of train.py (one render, L1 + SSIM loss, loss.backward()): the
    --gray branch (:213-223) with gray=True, the RGB branch (:292-296) with gray=False. Sets .grad
    on the six leaf tensors of pc and returns the loss."""
    # after the first call (which reads num_rendered back and sizes a persistent Workspace) nothing
    # blocks the host any more, exactly as in event_iteration_backward
    ws = _workspace_for(pc, cam) if (no_sync and pc._xyz.shape[0] > 0) else None
    st = _forward_raw(pc, cam, bg, ws=ws, slot=0)
    if ws is None and no_sync and pc._xyz.shape[0] > 0:
        _remember_workspace(pc, cam, st.R)
    out, seed = l1_ssim_loss_and_seed(st.color, gt_image, lambda_dssim, gray)
    D = int(pc.active_sh_degree)
    dummy_c = torch.zeros(1, device=pc._xyz.device)
    grads = _alloc_grads(pc, dummy_c, D)
    d_mean2D = torch.empty_like(pc._xyz)
    _backward_raw(pc, st, bg, seed, grads, False, d_mean2D)
    grads["c"].zero_()
    if ws is not None:
        try:
            st.R = ws.check()[0]
        except Workspace.Overflow:      # more instances than the workspace holds: it has been enlarged, redo the step
            return frame_iteration_backward(pc, cam, gt_image, bg, lambda_dssim, gray, return_pkg, no_sync)
    pc._xyz.grad = grads["xyz"]
    pc._features_dc.grad = grads["features_dc"]
    pc._features_rest.grad = grads["features_rest"]
    pc._opacity.grad = grads["opacity"]
    pc._scaling.grad = grads["scaling"]
    pc._rotation.grad = grads["rotation"]
    if return_pkg:
        vsp = torch.zeros_like(pc._xyz)
        vsp.grad = d_mean2D
        return out[0], {"render": st.color, "viewspace_points": vsp, "visibility_filter": st.radii > 0, "radii": st.radii,
                        "num_rendered": [st.R], "loss_terms": out}
    return out[0]
