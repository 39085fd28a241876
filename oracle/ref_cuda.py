"""TEST INFRASTRUCTURE ONLY: ctypes driver of oracle/_ref/libref_dgr.so -- the UNMODIFIED
reference rasterizer (submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu) built for
sm_100a by oracle/Makefile, behind the small C shim oracle/ref_cabi.cu.

It is (a) the bit-exact parity target of the -m gpu tests (radii, sorted keys, tile ranges,
n_contrib, image) and (b) the reference-CUDA timing leg of bench.py. Needs a GPU; the .so is
prebuilt in the container (where /root/reference exists) and travels to the GPU box.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref_dgr.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.ref_create.restype = ctypes.c_void_p
        L.ref_destroy.argtypes = [ctypes.c_void_p]
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        L.ref_forward.restype = i
        L.ref_forward.argtypes = [vp, i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, i, vp, vp, i]
        L.ref_backward.restype = i
        L.ref_backward.argtypes = [vp, i, i, i, i, vp, i, i, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp,
                                   vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i]
        L.ref_state_ptr.restype = vp
        L.ref_state_ptr.argtypes = [vp, i]
        L.ref_mark_visible.argtypes = [i, vp, vp, vp, vp]
        L.ref_copy_state.restype = i
        L.ref_copy_state.argtypes = [vp, i, vp, ctypes.c_size_t]
        _lib = L
    return _lib


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


class RefRasterizer:
    """One handle = one set of the reference's three opaque buffers (reused across calls)."""

    def __init__(self):
        self.h = lib().ref_create()

    def close(self):
        if self.h:
            lib().ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, means3D, opacities, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy, scales=None,
                rotations=None, cov3D_precomp=None, colors_precomp=None, shs=None, sh_degree=0, scale_modifier=1.0,
                prefiltered=False, debug=False):
        P = means3D.shape[0]
        dev = means3D.device
        M = 0 if shs is None else shs.shape[1]
        out = torch.zeros(3, H, W, dtype=torch.float32, device=dev)       # rasterize_points.cu:68
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        self.args = dict(P=P, D=sh_degree, M=M, W=W, H=H, bg=bg, means3D=means3D, shs=shs, colors=colors_precomp,
                         scales=scales, mod=scale_modifier, rotations=rotations, cov3D=cov3D_precomp, vm=viewmatrix,
                         pm=projmatrix, campos=campos, tx=tanfovx, ty=tanfovy, radii=radii)
        R = lib().ref_forward(self.h, P, sh_degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors_precomp),
                              _p(opacities), _p(scales), scale_modifier, _p(rotations), _p(cov3D_precomp),
                              _p(viewmatrix), _p(projmatrix), _p(campos), tanfovx, tanfovy, int(prefiltered),
                              out.data_ptr(), radii.data_ptr(), int(debug))
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.R = R
        return R, out, radii

    def backward(self, dL_dpix, debug=False):
        a = self.args
        P, M, dev = a["P"], a["M"], a["means3D"].device
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)   # rasterize_points.cu:151-159
        g = dict(dL_dmeans3D=z(P, 3), dL_dmeans2D=z(P, 3), dL_dcolors=z(P, 3), dL_dconic=z(P, 2, 2),
                 dL_dopacity=z(P, 1), dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        dL_dpix = dL_dpix.contiguous()
        rc = lib().ref_backward(self.h, P, a["D"], M, self.R, _p(a["bg"]), a["W"], a["H"], _p(a["means3D"]), _p(a["shs"]),
                                _p(a["colors"]), _p(a["scales"]), a["mod"], _p(a["rotations"]), _p(a["cov3D"]),
                                _p(a["vm"]), _p(a["pm"]), _p(a["campos"]), a["tx"], a["ty"], a["radii"].data_ptr(),
                                dL_dpix.data_ptr(), g["dL_dmeans2D"].data_ptr(), g["dL_dconic"].data_ptr(),
                                g["dL_dopacity"].data_ptr(), g["dL_dcolors"].data_ptr(), g["dL_dmeans3D"].data_ptr(),
                                g["dL_dcov3D"].data_ptr(), _p(g["dL_dsh"]) or 0, g["dL_dscales"].data_ptr(),
                                g["dL_drotations"].data_ptr(), int(debug))
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g

    _STATE = {"depths": (0, torch.float32, "P"), "means2D": (1, torch.float32, "2P"), "cov3D": (2, torch.float32, "6P"),
              "conic_opacity": (3, torch.float32, "4P"), "tiles_touched": (4, torch.int32, "P"),
              "point_offsets": (5, torch.int32, "P"), "point_list": (6, torch.int32, "R"),
              "point_list_keys": (7, torch.int64, "R"), "ranges": (8, torch.int32, "2T"),
              "n_contrib": (9, torch.int32, "N"), "accum_alpha": (10, torch.float32, "N"),
              "rgb": (13, torch.float32, "3P"), "clamped": (14, torch.uint8, "3P")}

    def state(self, name):
        """Copy of one array of the reference's opaque state after a forward (device tensor)."""
        which, dt, cnt = self._STATE[name]
        a = self.args
        T = ((a["W"] + 15) // 16) * ((a["H"] + 15) // 16)
        n = {"P": a["P"], "2P": 2 * a["P"], "3P": 3 * a["P"], "4P": 4 * a["P"], "6P": 6 * a["P"], "R": self.R,
             "2T": 2 * T, "N": a["W"] * a["H"]}[cnt]
        out = torch.empty(n, dtype=dt, device=a["means3D"].device)
        if n:
            rc = lib().ref_copy_state(self.h, which, out.data_ptr(), n * out.element_size())
            if rc != 0:
                raise RuntimeError("ref_copy_state failed: %d" % rc)
        return out


# ---- autograd wrapper with the reference's python surface (for the reference-CUDA timing leg) ----
class _RefRasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, colors_precomp, opacities, scales, rotations, settings, handle):
        s = settings
        R, out, radii = handle.forward(means3D.contiguous(), opacities.contiguous(), s.viewmatrix, s.projmatrix, s.campos,
                                       s.bg, s.image_width, s.image_height, s.tanfovx, s.tanfovy, scales=scales.contiguous(),
                                       rotations=rotations.contiguous(), colors_precomp=colors_precomp.contiguous(),
                                       scale_modifier=s.scale_modifier)
        ctx.handle = handle
        ctx.saved_args = handle.args
        ctx.R = R
        ctx.mark_non_differentiable(radii)
        return out, radii

    @staticmethod
    def backward(ctx, grad_out, _):
        h = ctx.handle
        h.args, h.R = ctx.saved_args, ctx.R
        g = h.backward(grad_out)
        return (g["dL_dmeans3D"], g["dL_dmeans2D"], g["dL_dcolors"], g["dL_dopacity"], g["dL_dscales"],
                g["dL_drotations"], None, None)


class RefGaussianRasterizer:
    """GaussianRasterizer-shaped front of the reference CUDA build. The reference keeps the three
    state buffers of every forward alive until its backward, so each live forward needs its own
    handle: `pool` hands them out round-robin (3 renders per event iteration)."""
    pool = []
    cursor = 0

    def __init__(self, raster_settings):
        self.raster_settings = raster_settings

    def __call__(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None):
        cls = RefGaussianRasterizer
        while len(cls.pool) < 4:
            cls.pool.append(RefRasterizer())
        h = cls.pool[cls.cursor % len(cls.pool)]
        cls.cursor += 1
        return _RefRasterize.apply(means3D, means2D, colors_precomp, opacities, scales, rotations, self.raster_settings, h)
