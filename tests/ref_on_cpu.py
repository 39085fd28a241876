"""TEST INFRASTRUCTURE: runs the REFERENCE's own Python (train.py, scene/, gaussian_renderer/, utils/ under
/root/reference, unmodified and not copied) on a machine without a GPU, on top of this repository's drop-in packages.

Why it exists: the reference tree is only mounted in the build container, which has no GPU; the GPU box only
receives this repository (reference sources may not be copied into it). So the claim "train.py --event --gray --deblur
runs unchanged on the drop-in" is checked here on CPU, in two halves:
  * everything ABOVE the C boundary is the real thing: the reference's train.py / Scene / GaussianModel / render() /
    loss_utils import `diff_gaussian_rasterization`, `simple_knn`, `plyfile`, `lpips`, `torchgeometry` from
    event_3dgs_b200/dropin and run through event_3dgs_b200.rasterizer (GaussianRasterizationSettings,
    GaussianRasterizer, _RasterizeGaussians with its autograd plumbing);
  * BELOW the boundary the three calls that would enter libevent3dgs.so (_forward_impl, _backward_impl,
    mark_visible) and distCUDA2 are answered by the CPU oracle (oracle/raster_oracle.py), which restates the same
    reference kernels; the CUDA kernels themselves are held against that oracle and the reference build by the -m gpu
    tests.
The reference hard-codes CUDA in its Python (`.cuda()`, device="cuda", torch.cuda.Event, torch.cuda.set_device):
install() redirects those to the CPU from outside, the reference's files stay untouched.

    python tests/ref_on_cpu.py train <dataset> <out> [train.py arguments ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("E3_REFERENCE_ROOT", "/root/reference")
DROPIN = os.path.join(ROOT, "event_3dgs_b200", "dropin")


def available():
    return os.path.isfile(os.path.join(REF, "train.py"))


def install(core="oracle"):
    """Path order of INTEGRATION.md section 2 (repository root, drop-in directory, then the reference), the CPU
    redirections, and the oracle behind the C boundary. core="record": the boundary calls are recorded in CALLS and
    answered with zeros instead (for comparing what arrives at the boundary)."""
    for p in (REF, DROPIN, ROOT):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    import numpy as np
    import torch

    # ---- CUDA spelled out in the reference's Python -> CPU ----
    def cpu_dev(d):
        return "cpu" if (isinstance(d, str) and d.startswith("cuda")) or (isinstance(d, torch.device) and d.type == "cuda") else d

    def strip(fn):
        def wrapped(*a, **k):
            if "device" in k:
                k["device"] = cpu_dev(k["device"])
            return fn(*a, **k)
        return wrapped
    for name in ("zeros", "ones", "empty", "full", "tensor", "rand", "randn", "zeros_like", "ones_like", "arange", "eye",
                 "as_tensor", "empty_like", "full_like", "randn_like", "rand_like"):
        setattr(torch, name, strip(getattr(torch, name)))
    _to = torch.Tensor.to

    def to(self, *a, **k):
        a = tuple(cpu_dev(x) for x in a)
        if "device" in k:
            k["device"] = cpu_dev(k["device"])
        return _to(self, *a, **k)
    torch.Tensor.to = to
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    _device = torch.device
    torch.cuda.set_device = lambda *a, **k: None
    torch.cuda.empty_cache = lambda: None
    torch.Tensor.get_device = lambda self: 0

    class Event:
        def __init__(self, *a, **k):
            import time
            self._t, self._time = 0.0, time

        def record(self, *a):
            self._t = self._time.perf_counter()

        def elapsed_time(self, other):
            return (other._t - self._t) * 1e3

        def synchronize(self):
            pass
    torch.cuda.Event = Event
    del _device

    # ---- below the C boundary: the CPU oracle ----
    from oracle import raster_oracle as ro
    from event_3dgs_b200 import knn as e3_knn
    from event_3dgs_b200 import rasterizer as rz
    states = {}
    CALLS.clear()

    def np_(t):
        return None if t is None or t.numel() == 0 else t.detach().cpu().numpy()

    def fwd(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3Ds_precomp, viewmatrix, projmatrix,
            tanfovx, tanfovy, image_height, image_width, sh, sh_degree, campos, prefiltered, debug):
        if means3D.dim() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        CALLS.append(("forward", dict(bg=bg, means3D=means3D, colors_precomp=colors_precomp, opacities=opacities, scales=scales,
                                      rotations=rotations, scale_modifier=scale_modifier, cov3Ds_precomp=cov3Ds_precomp,
                                      viewmatrix=viewmatrix, projmatrix=projmatrix, tanfovx=tanfovx, tanfovy=tanfovy,
                                      image_height=image_height, image_width=image_width, sh=sh, sh_degree=sh_degree, campos=campos,
                                      prefiltered=prefiltered, debug=debug)))
        P, H, W = means3D.size(0), int(image_height), int(image_width)
        e = torch.empty(0, dtype=torch.uint8)
        if core == "record" or P == 0:
            return 0, torch.zeros(3, H, W), torch.zeros(P, dtype=torch.int32), e, e, e
        kw = dict(scales=np_(scales), rotations=np_(rotations)) if np_(cov3Ds_precomp) is None else dict(cov3D_precomp=np_(cov3Ds_precomp))
        if np_(colors_precomp) is not None:
            kw["colors_precomp"] = np_(colors_precomp)
        else:
            kw.update(shs=np_(sh), sh_degree=int(sh_degree))
        st = ro.rasterize_forward(np_(means3D), np_(opacities), np_(viewmatrix), np_(projmatrix), np_(campos), np_(bg), W, H,
                                  float(tanfovx), float(tanfovy), scale_modifier=float(scale_modifier), prefiltered=bool(prefiltered), **kw)
        key = len(states) + 1
        states[key] = st
        handle = torch.tensor([key], dtype=torch.int64)
        return (int(st["num_rendered"]), torch.from_numpy(np.ascontiguousarray(st["out_color"])),
                torch.from_numpy(np.ascontiguousarray(st["radii"]).astype(np.int32)), handle, e, e)

    def bwd(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3Ds_precomp, viewmatrix, projmatrix, tanfovx,
            tanfovy, grad_out_color, sh, sh_degree, campos, geomBuffer, num_rendered, binningBuffer, imgBuffer, debug):
        CALLS.append(("backward", dict(grad_out_color=grad_out_color, num_rendered=num_rendered)))
        P = means3D.size(0)
        M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
        z = lambda *s: torch.zeros(*s)
        if core == "record" or P == 0 or geomBuffer.numel() == 0:
            return z(P, 3), z(P, 3), z(P, 1), z(P, 3), z(P, 6), z(P, M, 3), z(P, 3), z(P, 4)
        st = states.pop(int(geomBuffer[0]))
        g = ro.rasterize_backward(st, np_(grad_out_color))
        t = lambda k: torch.from_numpy(np.ascontiguousarray(g[k]))
        return (t("dL_dmeans2D"), t("dL_dcolors"), t("dL_dopacity"), t("dL_dmeans3D"), t("dL_dcov3D"), t("dL_dsh"), t("dL_dscales"),
                t("dL_drotations"))

    def vis(means3D, viewmatrix, projmatrix):
        return torch.from_numpy(np.asarray(ro.mark_visible(np_(means3D), np_(viewmatrix), np_(projmatrix))).astype(bool))

    def dist2(points):
        return torch.from_numpy(ro.dist2(np_(points)))

    rz._forward_impl, rz._backward_impl, rz.mark_visible = fwd, bwd, vis
    e3_knn.distCUDA2 = dist2
    return states


CALLS = []


def run_train(argv):
    """The reference's train.py as __main__ with the given command line."""
    import runpy
    install()
    os.chdir(REF)                 # train.py imports its sibling packages by relative top-level names
    sys.argv = ["train.py"] + list(argv)
    runpy.run_path(os.path.join(REF, "train.py"), run_name="__main__")


if __name__ == "__main__":
    if len(sys.argv) >= 2 and sys.argv[1] == "train":
        run_train(sys.argv[2:])
    else:
        raise SystemExit(__doc__)
