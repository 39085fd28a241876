"""Host-side mirror of the reference's rasterizer binding, on top of the C ABI.

Same names, argument order, return values and error behaviour as
submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py
(rasterize_gaussians :21-42, _RasterizeGaussians :44-155, GaussianRasterizationSettings
:157-169, GaussianRasterizer :171-220) and the tensor glue of rasterize_points.cu:35-217.
torch is used for device memory, streams and autograd plumbing only.
"""
import ctypes
import threading
import os
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib

LAST_NUM_RENDERED = 0     # R of the most recent forward (bench.py reads it for the roofline)

# ---- opt-in: forward without the host round trip (E3_DROPIN_NO_SYNC=1 or NO_SYNC = True) ----
# The reference reads num_rendered back in every forward (rasterizer_impl.cu:281) to size the binning
# buffer; that blocking read is what keeps the host from running ahead of the ~100 small ATen launches
# (activations, eval_sh, loss) between two renders. With NO_SYNC the first _LEARN calls per
# (P, W, H, device) run as the reference does and record R; afterwards the binning buffer is sized
# for _HEADROOM x the largest R seen, nothing blocks, and R is verified after the fact (it arrives in
# pinned memory). If a render ever exceeded its buffer, a RuntimeError says so at the next call:
# its image missed the instances beyond the capacity. `num_rendered` handed to the backward pass is
# then the capacity, not R. Off by default: the default keeps the reference's behaviour exactly.
NO_SYNC = os.environ.get("E3_DROPIN_NO_SYNC", "0") == "1"
_LEARN, _HEADROOM, _RING = 8, 2.0, 32
_cap_stats = {}           # (P, W, H, device) -> [calls, max R]
_ring = {}                # device -> {"host": pinned int32 [_RING, 2], "events": [...], "meta": [...], "next": int}


def _ring_for(dev):
    r = _ring.get(dev)
    if r is None:
        r = _ring[dev] = {"host": torch.zeros(_RING, 2, dtype=torch.int32).pin_memory(), "events": [torch.cuda.Event() for _ in range(_RING)],
                          "meta": [None] * _RING, "next": 0}
    return r


def _verify_slot(r, k, block):
    """Checks the read-back of ring slot k (if any is outstanding); returns False if it has not arrived yet."""
    meta = r["meta"][k]
    if meta is None:
        return True
    if not r["events"][k].query():
        if not block:
            return False
        r["events"][k].synchronize()
    key, cap = meta
    r["meta"][k] = None
    R, prefiltered = int(r["host"][k, 0]), int(r["host"][k, 1])
    global LAST_NUM_RENDERED
    LAST_NUM_RENDERED = R
    st = _cap_stats.setdefault(key, [0, 0])
    st[1] = max(st[1], R)
    if prefiltered:
        raise RuntimeError("Point is filtered although prefiltered is set. This shouldn't happen!")
    if R > cap:
        raise RuntimeError("event_3dgs_b200 (E3_DROPIN_NO_SYNC): an earlier render produced %d instances but its binning buffer "
                           "held %d; its image and gradients are incomplete. The capacity has been raised; repeat the step "
                           "(or unset E3_DROPIN_NO_SYNC)." % (R, cap))
    return True


def verify_no_sync(block=True):
    """Verifies every outstanding no-sync forward (call it before trusting results, e.g. once per epoch)."""
    for r in _ring.values():
        for k in range(_RING):
            _verify_slot(r, k, block)

# ---- workspace callbacks (replace resizeFunctional, rasterize_points.cu:27-33) ----
# The buffers handed out during ONE forward call, per host thread: the C call runs the callback synchronously on
# the calling thread, so a thread-local dictionary is private to the call in flight (one thread per GPU).
class _Slots(threading.local):
    def __init__(self):
        self.d = {}

    def clear(self):
        self.d = {}

    def __setitem__(self, k, v):
        self.d[k] = v

    def __getitem__(self, k):
        return self.d[k]

    def get(self, k, default=None):
        return self.d.get(k, default)


_slots = _Slots()


def _alloc(ctx, nbytes):
    dev = _slots["device"]
    t = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
    _slots[int(ctx) if ctx else 0] = t
    return t.data_ptr()


_ALLOC_CB = _lib.ALLOC_FN(_alloc)


def _ptr(t):
    """data pointer, or NULL for the reference's 'empty tensor means absent' convention."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _check_device(dev, **tensors):
    """Every non-empty tensor must live on the device of means3D (a host pointer handed to a
    kernel would fault asynchronously; the reference has the same requirement, implicitly)."""
    for name, t in tensors.items():
        if t is not None and t.numel() != 0 and t.device != dev:
            raise RuntimeError("event_3dgs_b200: `%s` is on %s but means3D is on %s" % (name, t.device, dev))


def _f32c(t):
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


def _forward_impl(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3Ds_precomp,
                  viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, sh_degree, campos,
                  prefiltered, debug):
    """RasterizeGaussiansCUDA, rasterize_points.cu:35-115."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("event_3dgs_b200: tensors must live on a CUDA device (there is no CPU path)")
    L = _lib.lib()
    P = means3D.size(0)
    H, W = int(image_height), int(image_width)
    dev = means3D.device
    _check_device(dev, bg=bg, colors_precomp=colors_precomp, opacities=opacities, scales=scales, rotations=rotations,
                  cov3Ds_precomp=cov3Ds_precomp, viewmatrix=viewmatrix, projmatrix=projmatrix, sh=sh, campos=campos)
    have_colors = colors_precomp is not None and colors_precomp.numel() != 0
    C = int(colors_precomp.size(1)) if have_colors else 3
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
    radii = torch.empty(P, dtype=torch.int32, device=dev)
    empty = torch.empty(0, dtype=torch.uint8, device=dev)
    if P == 0:
        return 0, torch.zeros(C, H, W, dtype=torch.float32, device=dev), radii, empty, empty, empty
    out_color = torch.empty(C, H, W, dtype=torch.float32, device=dev)
    keep = [_f32c(x) for x in (bg, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                              viewmatrix, projmatrix, campos)]
    bg_, means_, sh_, col_, op_, sc_, rot_, cov_, vm_, pm_, cam_ = keep
    key = (P, W, H, dev)
    stats = _cap_stats.setdefault(key, [0, 0])
    if NO_SYNC and not debug and stats[0] >= _LEARN and stats[1] > 0:
        r = _ring_for(dev)
        for k in range(_RING):                      # harvest what has arrived (non-blocking)
            _verify_slot(r, k, False)
        k = r["next"]
        r["next"] = (k + 1) % _RING
        _verify_slot(r, k, True)                    # the slot about to be reused must be done
        cap = int(_HEADROOM * stats[1]) + 4096
        u8 = dict(dtype=torch.uint8, device=dev)
        geom = torch.empty(L.e3_geometry_bytes(P), **u8)
        img = torch.empty(L.e3_image_bytes(W, H), **u8)
        binning = torch.empty(L.e3_binning_bytes(cap), **u8)
        tstream = torch.cuda.current_stream(dev)
        with torch.cuda.device(dev):
            code = L.e3_rasterize_forward_ws(
                geom.data_ptr(), binning.data_ptr(), cap, img.data_ptr(), r["host"][k].data_ptr(),
                P, int(sh_degree), M, C, _ptr(bg_), W, H, _ptr(means_), _ptr(sh_), _ptr(col_), _ptr(op_), _ptr(sc_),
                float(scale_modifier), _ptr(rot_), _ptr(cov_), _ptr(vm_), _ptr(pm_), _ptr(cam_), float(tanfovx),
                float(tanfovy), int(bool(prefiltered)), out_color.data_ptr(), radii.data_ptr(), 0, tstream.cuda_stream)
        _lib.check(code)
        r["events"][k].record(tstream)
        r["meta"][k] = (key, cap)
        stats[0] += 1
        return cap, out_color, radii, geom, binning, img
    R = ctypes.c_int(0)
    _slots.clear()
    _slots["device"] = dev
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        code = L.e3_rasterize_forward(
            _ALLOC_CB, 1, _ALLOC_CB, 2, _ALLOC_CB, 3,
            P, int(sh_degree), M, C, _ptr(bg_), W, H, _ptr(means_), _ptr(sh_), _ptr(col_), _ptr(op_), _ptr(sc_),
            float(scale_modifier), _ptr(rot_), _ptr(cov_), _ptr(vm_), _ptr(pm_), _ptr(cam_), float(tanfovx),
            float(tanfovy), int(bool(prefiltered)), out_color.data_ptr(), radii.data_ptr(), int(bool(debug)),
            stream, ctypes.byref(R))
    geom, binning, img = _slots.get(1, empty), _slots.get(2, empty), _slots.get(3, empty)
    _slots.clear()
    _lib.check(code)
    global LAST_NUM_RENDERED
    LAST_NUM_RENDERED = R.value
    stats[0] += 1
    stats[1] = max(stats[1], R.value)
    return R.value, out_color, radii, geom, binning, img


def _backward_impl(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3Ds_precomp, viewmatrix,
                   projmatrix, tanfovx, tanfovy, grad_out_color, sh, sh_degree, campos, geomBuffer, num_rendered,
                   binningBuffer, imgBuffer, debug):
    """RasterizeGaussiansBackwardCUDA, rasterize_points.cu:117-196."""
    L = _lib.lib()
    P = means3D.size(0)
    C, H, W = grad_out_color.shape
    dev = means3D.device
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
    have_scales = scales is not None and scales.numel() != 0
    f32 = dict(dtype=torch.float32, device=dev)
    new = torch.zeros if P == 0 else torch.empty      # the kernel writes every row when P > 0
    dL_dmeans3D = new(P, 3, **f32)
    dL_dmeans2D = new(P, 3, **f32)
    dL_dcolors = new(P, C, **f32)
    dL_dopacity = new(P, 1, **f32)
    dL_dcov3D = new(P, 6, **f32)
    dL_dsh = new(P, M, 3, **f32)
    dL_dscales = new(P, 3, **f32) if have_scales else torch.zeros(P, 3, **f32)
    dL_drotations = new(P, 4, **f32) if have_scales else torch.zeros(P, 4, **f32)
    if P != 0:
        keep = [_f32c(x) for x in (bg, means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp, viewmatrix,
                                  projmatrix, campos, grad_out_color)]
        bg_, means_, sh_, col_, sc_, rot_, cov_, vm_, pm_, cam_, gout_ = keep
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            code = L.e3_rasterize_backward(
                P, int(sh_degree), M, int(num_rendered), int(C), _ptr(bg_), int(W), int(H), _ptr(means_), _ptr(sh_),
                _ptr(col_), _ptr(sc_), float(scale_modifier), _ptr(rot_), _ptr(cov_), _ptr(vm_), _ptr(pm_), _ptr(cam_),
                float(tanfovx), float(tanfovy), radii.data_ptr(), _ptr(geomBuffer), _ptr(binningBuffer),
                _ptr(imgBuffer), gout_.data_ptr(), dL_dmeans2D.data_ptr(), None, dL_dopacity.data_ptr(),
                dL_dcolors.data_ptr(), dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), _ptr(dL_dsh),
                dL_dscales.data_ptr() if have_scales else None, dL_drotations.data_ptr() if have_scales else None,
                int(bool(debug)), stream)
        _lib.check(code)
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible, rasterize_points.cu:198-217."""
    P = means3D.size(0)
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P != 0:
        m, v, p = _f32c(means3D), _f32c(viewmatrix), _f32c(projmatrix)
        with torch.cuda.device(means3D.device):
            _lib.check(_lib.lib().e3_mark_visible(P, m.data_ptr(), v.data_ptr(), p.data_ptr(), present.data_ptr(),
                                                  torch.cuda.current_stream(means3D.device).cuda_stream))
    return present


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        args = (raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
                raster_settings.image_height, raster_settings.image_width, sh, raster_settings.sh_degree,
                raster_settings.campos, raster_settings.prefiltered, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _forward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _forward_impl(*args)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations, raster_settings.scale_modifier,
                cov3Ds_precomp, raster_settings.viewmatrix, raster_settings.projmatrix, raster_settings.tanfovx,
                raster_settings.tanfovy, grad_out_color, sh, raster_settings.sh_degree, raster_settings.campos,
                geomBuffer, num_rendered, binningBuffer, imgBuffer, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _backward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            out = _backward_impl(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = out
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   raster_settings)
