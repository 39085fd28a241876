"""ctypes binding of libevent3dgs.so (the C ABI declared in include/event3dgs.h).

There is no fallback: if the CUDA library is missing or fails to load, importing a symbol from
here raises -- the product path never runs on the CPU or through the oracle.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libevent3dgs.so")

ALLOC_FN = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)

_vp, _i, _f, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t

# name -> (restype, argtypes); kept in one table so tests can check it against the header
SIGNATURES = {
    "e3_error_string": (ctypes.c_char_p, [_i]),
    "e3_version": (_i, []),
    "e3_geometry_bytes": (_sz, [_i]),
    "e3_image_bytes": (_sz, [_i, _i]),
    "e3_binning_bytes": (_sz, [_i]),
    "e3_rasterize_forward": (_i, [ALLOC_FN, _vp, ALLOC_FN, _vp, ALLOC_FN, _vp,
                                  _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp,
                                  _vp, _vp, _vp, _f, _f, _i, _vp, _vp, _i, _vp, ctypes.POINTER(_i)]),
    "e3_rasterize_backward": (_i, [_i, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _f, _vp, _vp,
                                   _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                   _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "e3_rasterize_forward_raw": (_i, [ALLOC_FN, _vp, ALLOC_FN, _vp, ALLOC_FN, _vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp,
                                      _vp, _f, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _i, _vp, ctypes.POINTER(_i)]),
    "e3_rasterize_forward_raw_ws": (_i, [_vp, _vp, _sz, _vp, _vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp,
                                         _vp, _f, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _i, _vp]),
    "e3_rasterize_backward_raw": (_i, [_i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _f, _f,
                                       _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "e3_event_loss": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _f, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "e3_l1_ssim_scratch_bytes": (_sz, [_i, _i, _i]),
    "e3_l1_ssim_loss": (_i, [_i, _i, _i, _vp, _vp, _i, _f, _f, _vp, _sz, _vp, _vp, _vp]),
    "e3_mark_visible": (_i, [_i, _vp, _vp, _vp, _vp, _vp]),
    "e3_dist2_scratch_bytes": (_sz, [_i]),
    "e3_dist2": (_i, [_i, _vp, _vp, _vp, _sz, _vp]),
    "e3_scan_scratch_bytes": (_sz, [_sz]),
    "e3_exclusive_scan": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _vp]),
    "e3_exclusive_scan_noalloc_max": (_sz, []),
    "e3_exclusive_scan_noalloc": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "e3_profile_enable": (_i, [ctypes.c_uint]),
    "e3_kernel_launches": (ctypes.c_ulonglong, []),
    "e3_profile_read": (_i, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong), _i]),
    "e3_geometry_ptr": (_vp, [_vp, _i, _i]),
    "e3_binning_ptr": (_vp, [_vp, _i, _i]),
    "e3_image_ptr": (_vp, [_vp, _i, _i, _i]),
}



class AdamGroup(ctypes.Structure):
    """e3_adam_group of include/event3dgs.h."""
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("exp_avg", ctypes.c_void_p),
                ("exp_avg_sq", ctypes.c_void_p), ("n", ctypes.c_ulonglong), ("lr", ctypes.c_float), ("eps", ctypes.c_float)]


class DensifyArray(ctypes.Structure):
    """e3_densify_array of include/event3dgs.h."""
    _fields_ = [("src", ctypes.c_void_p), ("dst", ctypes.c_void_p), ("exp_avg_src", ctypes.c_void_p),
                ("exp_avg_dst", ctypes.c_void_p), ("exp_avg_sq_src", ctypes.c_void_p), ("exp_avg_sq_dst", ctypes.c_void_p),
                ("width", ctypes.c_int), ("kind", ctypes.c_int)]


class View(ctypes.Structure):
    """e3_view of include/event3dgs.h."""
    _fields_ = [("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p), ("cam_pos", ctypes.c_void_p),
                ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float), ("channels", ctypes.c_int),
                ("geometry_buffer", ctypes.c_void_p), ("binning_buffer", ctypes.c_void_p), ("image_buffer", ctypes.c_void_p),
                ("out_color", ctypes.c_void_p), ("radii", ctypes.c_void_p), ("seed", ctypes.c_void_p),
                ("num_rendered_host", ctypes.c_void_p)]


class EventIterationArgs(ctypes.Structure):
    """e3_event_iteration_args of include/event3dgs.h."""
    _fields_ = [("P", ctypes.c_int), ("active_sh_degree", ctypes.c_int), ("width", ctypes.c_int), ("height", ctypes.c_int),
                ("binning_capacity", ctypes.c_ulonglong), ("scale_modifier", ctypes.c_float), ("accumulate", ctypes.c_int),
                ("background", ctypes.c_void_p),
                ("xyz", ctypes.c_void_p), ("features_dc", ctypes.c_void_p), ("features_rest", ctypes.c_void_p),
                ("opacity_logits", ctypes.c_void_p), ("log_scales", ctypes.c_void_p), ("raw_rotations", ctypes.c_void_p),
                ("d_xyz", ctypes.c_void_p), ("d_features_dc", ctypes.c_void_p), ("d_features_rest", ctypes.c_void_p),
                ("d_opacity_logits", ctypes.c_void_p), ("d_log_scales", ctypes.c_void_p), ("d_raw_rotations", ctypes.c_void_p),
                ("d_mean2D", ctypes.c_void_p),
                ("views", View * 3),
                ("gt_now", ctypes.c_void_p), ("gt_next", ctypes.c_void_p), ("gt_intensity", ctypes.c_void_p), ("gt_blur", ctypes.c_void_p),
                ("gt_is_u8", ctypes.c_int), ("c", ctypes.c_void_p), ("out5", ctypes.c_void_p), ("loss_scratch", ctypes.c_void_p),
                ("wait_before_loss_event", ctypes.c_void_p), ("forwards_done_event", ctypes.c_void_p),
                ("loss_upstream", ctypes.c_float), ("grad_multicast_delta", ctypes.c_longlong)]


SIGNATURES["e3_event_iteration"] = (_i, [ctypes.POINTER(EventIterationArgs), _vp])
SIGNATURES["e3_adam_step"] = (_i, [ctypes.POINTER(AdamGroup), _i, _i, ctypes.c_double, ctypes.c_double, _vp])
SIGNATURES["e3_rasterize_forward_ws"] = (_i, [_vp, _vp, _sz, _vp, _vp, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp,
                                             _vp, _vp, _vp, _f, _f, _i, _vp, _vp, _i, _vp])
SIGNATURES["e3_rasterize_backward_raw_dual"] = (_i, [_i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _f, _f,
                                                    _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp])
SIGNATURES["e3_multimem_add"] = (_i, [_vp, _vp, _sz, _f, _vp])
SIGNATURES["e3_nvls_sharded_adam"] = (_i, [ctypes.POINTER(AdamGroup), _i, _i, ctypes.c_double, ctypes.c_double, ctypes.c_longlong,
                                          ctypes.c_longlong, _i, _i, _f, _vp])
SIGNATURES["e3_nvls_allreduce"] = (_i, [_vp, _sz, _i, _i, _f, _vp])
SIGNATURES["e3_densification_stats"] = (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp])
SIGNATURES["e3_densify_scratch_bytes"] = (_sz, [_i])
SIGNATURES["e3_densify_plan"] = (_i, [_i, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _i, _vp, _sz, _vp, _vp])
SIGNATURES["e3_densify_apply"] = (_i, [_i, _i, _i, ctypes.POINTER(DensifyArray), _i, _vp, _vp, _vp, _i, _vp, _sz, _vp])
SIGNATURES["e3_reset_opacity"] = (_i, [_i, _vp, _vp, _vp, _f, _vp])

_lib = None


def build(verbose=False):
    """Compile every .cu under csrc/ for sm_100a into libevent3dgs.so (in-tree)."""
    cmd = ["make", "-f", os.path.join(_HERE, "csrc", "Makefile")]
    subprocess.check_call(cmd, stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "event_3dgs_b200: %s is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)     # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code):
    if code != 0:
        msg = lib().e3_error_string(code)
        raise RuntimeError(msg.decode() if msg else "event3dgs error %d" % code)


STAGES = ["preprocess_fwd", "depth_sort", "offsets_scan", "emit", "tile_sort", "tile_ranges", "render_fwd",
          "render_bwd", "preprocess_bwd", "event_loss", "dist2", "activate"]
NUM_STAGES = 16


def profile_enable(names):
    """Bracket the named stages with CUDA events on the launching stream (names: see STAGES)."""
    mask = 0
    for n in names:
        mask |= 1 << STAGES.index(n)
    check(lib().e3_profile_enable(mask))


def profile_read(reset=True):
    """{stage: (total_ms, launches)} since the last reset."""
    ms = (ctypes.c_double * NUM_STAGES)()
    cnt = (ctypes.c_longlong * NUM_STAGES)()
    check(lib().e3_profile_read(ms, cnt, int(reset)))
    return {n: (ms[i], cnt[i]) for i, n in enumerate(STAGES) if cnt[i]}
