"""Views into the opaque state buffers of a finished forward (for tests and debugging).
Wraps e3_geometry_ptr / e3_binning_ptr / e3_image_ptr of include/event3dgs.h."""
import torch

from . import _lib

_GEOM = {"depths": (0, torch.float32, 1), "rec": (1, torch.float32, 12), "cov3D": (2, torch.float32, 6),
         "tiles_touched": (3, torch.int32, 1), "depth_order": (4, torch.int32, 1), "rgb": (5, torch.float32, 3),
         "clamped": (6, torch.uint8, 3)}
_BIN = {"point_list": (0, torch.int32, 1), "tile_ids": (1, torch.int32, 1)}
_IMG = {"final_T": (0, torch.float32, "N"), "n_contrib": (1, torch.int32, "N"), "ranges": (2, torch.int32, "2T")}


def _view(buf, ptr, count, dtype):
    off = ptr - buf.data_ptr()
    nbytes = count * torch.empty(0, dtype=dtype).element_size()
    assert 0 <= off and off + nbytes <= buf.numel()
    return buf[off:off + nbytes].view(dtype)


def geometry(buf, P, name):
    which, dt, per = _GEOM[name]
    return _view(buf, _lib.lib().e3_geometry_ptr(buf.data_ptr(), P, which), P * per, dt)


def binning(buf, R, name):
    which, dt, per = _BIN[name]
    return _view(buf, _lib.lib().e3_binning_ptr(buf.data_ptr(), R, which), R * per, dt)


def image(buf, W, H, name):
    which, dt, cnt = _IMG[name]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    n = W * H if cnt == "N" else 2 * T
    return _view(buf, _lib.lib().e3_image_ptr(buf.data_ptr(), W, H, which), n, dt)
