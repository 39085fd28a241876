"""TEST INFRASTRUCTURE ONLY: numpy/ctypes driver of the plain-C rasterizer oracle.

Mirrors the host orchestration of CudaRasterizer::Rasterizer::forward / backward
(submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:198-336, 340-434)
and the tensor glue of rasterize_points.cu:35-196, stage by stage, on numpy arrays.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-f", os.path.join(_HERE, "Makefile"), "liboracle"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        src = [os.path.join(_HERE, f) for f in ("raster_oracle.c", "knn_oracle.c")]
        if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in src):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.orc_get_higher_msb.restype = ctypes.c_uint32
        _LIB.orc_preprocess.restype = ctypes.c_int
    return _LIB


def _p(a):
    if a is None:
        return ctypes.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def set_threads(n):
    """Sets the OpenMP thread count of the oracle's loops (omp_set_num_threads) and returns the count in
    effect; the environment variable alone is read once when libgomp loads and may have been inherited."""
    os.environ["OMP_NUM_THREADS"] = str(n)
    return int(lib().orc_set_threads(int(n)))


def get_higher_msb(n):
    return int(lib().orc_get_higher_msb(ctypes.c_uint32(n)))


def mark_visible(means3D, viewmatrix, projmatrix):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    out = np.zeros(P, dtype=np.uint8)
    lib().orc_mark_visible(P, _p(means3D), _p(_f32(viewmatrix).reshape(-1)), _p(_f32(projmatrix).reshape(-1)), _p(out))
    return out.astype(bool)


def rasterize_forward(means3D, opacities, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy,
                      scales=None, rotations=None, cov3D_precomp=None, colors_precomp=None, shs=None,
                      sh_degree=0, scale_modifier=1.0, prefiltered=False):
    """Rasterizer::forward (rasterizer_impl.cu:198-336). Returns a dict holding every
    intermediate the reference keeps in its three opaque buffers plus the outputs."""
    L = lib()
    means3D = _f32(means3D)
    P = means3D.shape[0]
    opacities = _f32(opacities).reshape(-1)
    scales, rotations, cov3D_precomp = _f32(scales), _f32(rotations), _f32(cov3D_precomp)
    colors_precomp, shs = _f32(colors_precomp), _f32(shs)
    vm = _f32(viewmatrix).reshape(-1)
    pm = _f32(projmatrix).reshape(-1)
    campos = _f32(campos).reshape(-1)
    bg = _f32(bg).reshape(-1)
    C = 3 if colors_precomp is None else int(colors_precomp.shape[1])
    M = 0 if shs is None else int(shs.shape[1])
    focal_y = np.float32(H) / (np.float32(2.0) * np.float32(tanfovy))   # rasterizer_impl.cu:222-223
    focal_x = np.float32(W) / (np.float32(2.0) * np.float32(tanfovx))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    st = dict(P=P, W=W, H=H, C=C, M=M, D=sh_degree, gx=gx, gy=gy, bg=bg, vm=vm, pm=pm, campos=campos,
              tanfovx=float(tanfovx), tanfovy=float(tanfovy), focal_x=float(focal_x), focal_y=float(focal_y),
              means3D=means3D, scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp,
              colors_precomp=colors_precomp, shs=shs, scale_modifier=float(scale_modifier))
    st["radii"] = np.zeros(P, np.int32)
    st["means2D"] = np.zeros((P, 2), np.float32)
    st["depths"] = np.zeros(P, np.float32)
    st["cov3D"] = np.zeros((P, 6), np.float32)
    st["rgb"] = np.zeros((P, 3), np.float32)
    st["clamped"] = np.zeros((P, 3), np.uint8)
    st["conic_opacity"] = np.zeros((P, 4), np.float32)
    st["tiles_touched"] = np.zeros(P, np.uint32)
    st["out_color"] = np.zeros((C, H, W), np.float32)
    st["final_T"] = np.zeros(H * W, np.float32)
    st["n_contrib"] = np.zeros(H * W, np.uint32)
    st["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    st["num_rendered"] = 0
    if P == 0:     # rasterize_points.cu:81 -- all-zero image
        st["point_offsets"] = np.zeros(0, np.uint32)
        st["point_list"] = np.zeros(0, np.uint32)
        st["point_list_keys"] = np.zeros(0, np.uint64)
        return st
    rc = L.orc_preprocess(P, sh_degree, M, _p(means3D), _p(scales), ctypes.c_float(scale_modifier), _p(rotations),
                          _p(opacities), _p(shs), _p(st["clamped"]), _p(cov3D_precomp), _p(colors_precomp), _p(vm),
                          _p(pm), _p(campos), W, H, ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
                          ctypes.c_float(focal_x), ctypes.c_float(focal_y), _p(st["radii"]), _p(st["means2D"]),
                          _p(st["depths"]), _p(st["cov3D"]), _p(st["rgb"]), _p(st["conic_opacity"]), gx, gy,
                          _p(st["tiles_touched"]), int(prefiltered))
    if rc != 0:
        raise RuntimeError("Point is filtered although prefiltered is set")
    offsets = np.zeros(P, np.uint32)
    L.orc_inclusive_sum(P, _p(st["tiles_touched"]), _p(offsets))
    st["point_offsets"] = offsets
    R = int(offsets[-1])
    st["num_rendered"] = R
    keys_u = np.zeros(R, np.uint64)
    vals_u = np.zeros(R, np.uint32)
    L.orc_duplicate_with_keys(P, _p(st["means2D"]), _p(st["depths"]), _p(offsets), _p(keys_u), _p(vals_u),
                              _p(st["radii"]), gx, gy)
    bit = get_higher_msb(gx * gy)
    keys = np.zeros(R, np.uint64)
    vals = np.zeros(R, np.uint32)
    ktmp = np.zeros(R, np.uint64)
    vtmp = np.zeros(R, np.uint32)
    L.orc_sort_pairs(R, _p(keys_u), _p(keys), _p(vals_u), _p(vals), 32 + bit, _p(ktmp), _p(vtmp))
    st["point_list_keys_unsorted"], st["point_list_unsorted"] = keys_u, vals_u
    st["point_list_keys"], st["point_list"] = keys, vals
    if R > 0:
        L.orc_identify_tile_ranges(R, _p(keys), _p(st["ranges"]))
    feat = colors_precomp if colors_precomp is not None else st["rgb"]
    L.orc_render_forward(C, gx, gy, _p(st["ranges"]), _p(vals), W, H, _p(st["means2D"]), _p(feat),
                         _p(st["conic_opacity"]), _p(st["final_T"]), _p(st["n_contrib"]), _p(bg), _p(st["out_color"]))
    return st


def rasterize_backward(st, dL_dpix):
    """Rasterizer::backward (rasterizer_impl.cu:340-434) + the zero-filled gradient tensors of
    rasterize_points.cu:151-159. Returns the reference's 8 gradients by name."""
    L = lib()
    P, C, M, W, H = st["P"], st["C"], st["M"], st["W"], st["H"]
    dL_dpix = _f32(dL_dpix)
    g = dict(dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dmeans2D=np.zeros((P, 3), np.float32),
             dL_dcolors=np.zeros((P, C), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
             dL_dsh=np.zeros((P, M, 3), np.float32), dL_dscales=np.zeros((P, 3), np.float32),
             dL_drotations=np.zeros((P, 4), np.float32))
    if P == 0:
        return g
    feat = st["colors_precomp"] if st["colors_precomp"] is not None else st["rgb"]
    L.orc_render_backward(C, P, st["gx"], st["gy"], _p(st["ranges"]), _p(st["point_list"]), W, H, _p(st["bg"]),
                          _p(st["means2D"]), _p(st["conic_opacity"]), _p(feat), _p(st["final_T"]),
                          _p(st["n_contrib"]), _p(dL_dpix), _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]),
                          _p(g["dL_dopacity"]), _p(g["dL_dcolors"]))
    cov3D = st["cov3D_precomp"] if st["cov3D_precomp"] is not None else st["cov3D"]
    L.orc_cov2d_backward(P, _p(st["means3D"]), _p(st["radii"]), _p(cov3D), ctypes.c_float(st["focal_x"]),
                         ctypes.c_float(st["focal_y"]), ctypes.c_float(st["tanfovx"]), ctypes.c_float(st["tanfovy"]),
                         _p(st["vm"]), _p(g["dL_dconic"]), _p(g["dL_dmeans3D"]), _p(g["dL_dcov3D"]))
    L.orc_preprocess_backward(P, st["D"], M, _p(st["means3D"]), _p(st["radii"]), _p(st["shs"]), _p(st["clamped"]),
                              _p(st["scales"]), _p(st["rotations"]), ctypes.c_float(st["scale_modifier"]),
                              _p(st["pm"]), _p(st["campos"]), _p(g["dL_dmeans2D"]), _p(g["dL_dmeans3D"]),
                              _p(g["dL_dcolors"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]), _p(g["dL_dscales"]),
                              _p(g["dL_drotations"]))
    return g


def dist2(points):
    """simple_knn._C.distCUDA2 contract (scene/gaussian_model.py:134): brute force, exact."""
    pts = _f32(points)
    out = np.zeros(pts.shape[0], np.float32)
    lib().orc_dist2(pts.shape[0], _p(pts), _p(out))
    return out
