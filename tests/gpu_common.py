"""Helpers of the -m gpu tests: run the same numpy inputs through (a) the product path via the
drop-in API (which calls the C ABI), (b) the reference CUDA build (oracle/_ref), (c) the C oracle."""
import numpy as np
import torch

from event_3dgs_b200 import inspect_state as ins
from event_3dgs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, _backward_impl,
                                        _forward_impl)

DEV = "cuda:0"


def T(x):
    return None if x is None else torch.as_tensor(np.ascontiguousarray(x)).to(DEV)


def ours_forward(inp, use_sh=False, sh_degree=0, cov3D_precomp=None, debug=False, prefiltered=False):
    """Calls RasterizeGaussiansCUDA's replacement directly; returns tensors + opaque buffers."""
    t = {k: T(v) for k, v in inp.items() if isinstance(v, np.ndarray)}
    e = torch.Tensor([])
    colors = e if use_sh else t["colors"]
    sh = t["shs"] if use_sh else e
    scales, rots, cov = (e, e, T(cov3D_precomp)) if cov3D_precomp is not None else (t["scales"], t["rotations"], e)
    args = (t["bg"], t["means3D"], colors, t["opacities"], scales, rots, 1.0, cov, t["viewmatrix"], t["projmatrix"],
            inp["tanfovx"], inp["tanfovy"], inp["H"], inp["W"], sh, sh_degree, t["campos"], prefiltered, debug)
    R, color, radii, geom, binning, img = _forward_impl(*args)
    return dict(R=R, color=color, radii=radii, geom=geom, binning=binning, img=img, t=t, args=args,
                colors=colors, sh=sh, scales=scales, rots=rots, cov=cov, sh_degree=sh_degree)


def ours_backward(fw, inp, dL):
    t = fw["t"]
    out = _backward_impl(t["bg"], t["means3D"], fw["radii"], fw["colors"], fw["scales"], fw["rots"], 1.0, fw["cov"],
                         t["viewmatrix"], t["projmatrix"], inp["tanfovx"], inp["tanfovy"], T(dL), fw["sh"],
                         fw["sh_degree"], t["campos"], fw["geom"], fw["R"], fw["binning"], fw["img"], False)
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations"]
    return dict(zip(names, out))


def ours_state(fw, inp):
    P, W, H, R = inp["means3D"].shape[0], inp["W"], inp["H"], fw["R"]
    s = dict(depths=ins.geometry(fw["geom"], P, "depths"), rec=ins.geometry(fw["geom"], P, "rec").view(P, 12),
             cov3D=ins.geometry(fw["geom"], P, "cov3D").view(P, 6), tiles_touched=ins.geometry(fw["geom"], P, "tiles_touched"),
             final_T=ins.image(fw["img"], W, H, "final_T"), n_contrib=ins.image(fw["img"], W, H, "n_contrib"),
             ranges=ins.image(fw["img"], W, H, "ranges").view(-1, 2))
    if R > 0:
        s["point_list"] = ins.binning(fw["binning"], R, "point_list")
        s["tile_ids"] = ins.binning(fw["binning"], R, "tile_ids")
    else:
        s["point_list"] = torch.zeros(0, dtype=torch.int32, device=DEV)
        s["tile_ids"] = torch.zeros(0, dtype=torch.int32, device=DEV)
    return s


def sorted_keys(state, radii):
    """Rebuild the reference's 64-bit sorted keys (tile << 32 | depth bits) from our state."""
    pl = state["point_list"].long()
    dbits = state["depths"].view(torch.int32)[pl].long() & 0xffffffff
    return (state["tile_ids"].long() << 32) | dbits


def ref_forward(inp, use_sh=False, sh_degree=0, cov3D_precomp=None):
    from oracle import ref_cuda
    r = ref_cuda.RefRasterizer()
    t = {k: T(v) for k, v in inp.items() if isinstance(v, np.ndarray)}
    kw = dict(scales=t["scales"], rotations=t["rotations"])
    if cov3D_precomp is not None:
        kw = dict(cov3D_precomp=T(cov3D_precomp))
    if use_sh:
        kw.update(shs=t["shs"], sh_degree=sh_degree)
    else:
        kw.update(colors_precomp=t["colors"])
    R, out, radii = r.forward(t["means3D"], t["opacities"].reshape(-1).contiguous(), t["viewmatrix"], t["projmatrix"],
                              t["campos"], t["bg"], inp["W"], inp["H"], inp["tanfovx"], inp["tanfovy"], **kw)
    torch.cuda.synchronize()
    return r, R, out, radii
