"""Adaptive density control on the B200: host side of e3_densification_stats, e3_densify_plan /
e3_densify_apply and e3_reset_opacity (csrc/densify.cu).

Mirrors the reference's GaussianModel methods with the same names and argument meaning
(scene/gaussian_model.py): add_densification_stats :405-407 (+ the max_radii2D update of
train.py:319), densify_and_prune :389-403, reset_opacity :210-213, and the state they keep
(xyz_gradient_accum, denom, max_radii2D :116-121,183-185). The optimizer is edited the way the
reference does (:258-340): each group keeps its single parameter, which is replaced by a new
nn.Parameter, and the state dict entry {"step", "exp_avg", "exp_avg_sq"} moves to the new key.

There is no CPU fallback: without the CUDA library these functions raise.
"""

import numpy as np
import torch
from torch import nn

from . import _lib

# optimizer group name (scene/gaussian_model.py:160-167) -> attribute of the model, row kind of e3_densify_array
GROUPS = (("xyz", "_xyz", 1), ("f_dc", "_features_dc", 0), ("f_rest", "_features_rest", 0), ("opacity", "_opacity", 0),
          ("scaling", "_scaling", 2), ("rotation", "_rotation", 0))


class DensificationState:
    """xyz_gradient_accum [P,1], denom [P,1], max_radii2D [P] (scene/gaussian_model.py:183-185)."""

    def __init__(self, P, device):
        self.reset(P, device)

    def reset(self, P, device):
        buf = torch.zeros(3, P, dtype=torch.float32, device=device)
        self._buf = buf
        self.xyz_gradient_accum = buf[0].view(P, 1)
        self.denom = buf[1].view(P, 1)
        self.max_radii2D = buf[2]


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not (t.is_cuda and t.is_contiguous()):
            raise RuntimeError("event3dgs densification: tensors must be contiguous CUDA tensors")


def add_densification_stats(state, viewspace_point_grad, radii):
    """train.py:317-320 in one launch. `viewspace_point_grad` is viewspace_points.grad [P,3] of the FIRST
    render, `radii` its int32 radii (visibility_filter = radii > 0)."""
    P = radii.shape[0]
    _require_cuda(viewspace_point_grad, radii, state._buf)
    if radii.dtype != torch.int32 or viewspace_point_grad.dtype != torch.float32 or viewspace_point_grad.shape != (P, 3):
        raise RuntimeError("add_densification_stats: radii must be int32 [P], the gradient float32 [P,3]")
    if state.max_radii2D.shape[0] != P:
        raise RuntimeError("add_densification_stats: state holds %d rows, render has %d" % (state.max_radii2D.shape[0], P))
    with torch.cuda.device(radii.device):
        _lib.check(_lib.lib().e3_densification_stats(P, radii.data_ptr(), viewspace_point_grad.data_ptr(),
                                                     state.max_radii2D.data_ptr(), state.xyz_gradient_accum.data_ptr(),
                                                     state.denom.data_ptr(), _stream(radii.device)))


def _group_of(optimizer, name, param):
    if optimizer is None:
        return None
    for group in optimizer.param_groups:
        if group.get("name") == name or any(p is param for p in group["params"]):
            if len(group["params"]) != 1:
                raise RuntimeError("densification expects one parameter per optimizer group (gaussian_model.py:310)")
            return group
    return None


def densify_and_prune(pc, optimizer, state, max_grad, min_opacity, extent, max_screen_size, percent_dense=0.01,
                      normals=None, generator=None, split_n=2):
    """GaussianModel.densify_and_prune(max_grad, min_opacity, extent, max_screen_size). `normals`
    ([split_n * n_selected, 3] standard-normal draws, or a callable n_rows -> tensor) replaces the
    draw of torch.normal (gaussian_model.py:360) when the caller needs a fixed sample, e.g. the
    replicated multi-GPU trainer (every rank must draw the same) and the parity tests;
    by default they come from torch.randn on the device with `generator`.
    Returns the dict of counts {"kept", "cloned", "children_per_draw", "split_selected", "rows"}."""
    L = _lib.lib()
    dev = pc._xyz.device
    P = int(pc._xyz.shape[0])
    params = [getattr(pc, attr) for _, attr, _ in GROUPS]
    _require_cuda(*params, state._buf)
    if state.max_radii2D.shape[0] != P:
        raise RuntimeError("densify_and_prune: statistics hold %d rows, the model %d" % (state.max_radii2D.shape[0], P))
    f32 = lambda v: float(np.float32(v))
    dense_extent = f32(percent_dense * extent)
    big_ws = f32(0.1 * extent) if max_screen_size else -1.0
    with torch.cuda.device(dev), torch.no_grad():
        st = _stream(dev)
        nbytes = int(L.e3_densify_scratch_bytes(P))
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        counts_dev = torch.empty(8, dtype=torch.int32, device=dev)
        _lib.check(L.e3_densify_plan(P, state.xyz_gradient_accum.data_ptr(), state.denom.data_ptr(), pc._scaling.data_ptr(),
                                     pc._opacity.data_ptr(), f32(max_grad), dense_extent, f32(min_opacity), big_ws, split_n,
                                     scratch.data_ptr(), nbytes, counts_dev.data_ptr(), st))
        kept, cloned, children, selected, rows = counts_dev[:5].tolist()      # the one host sync: sizes of the new tensors
        z = None
        if selected > 0:
            z = normals(split_n * selected) if callable(normals) else normals
            if z is None:
                z = torch.randn(split_n * selected, 3, device=dev, generator=generator)
            z = z.to(device=dev, dtype=torch.float32).contiguous()
            if z.shape != (split_n * selected, 3):
                raise RuntimeError("densify_and_prune: normals must be [%d, 3], got %s" % (split_n * selected, tuple(z.shape)))
        arr = (_lib.DensifyArray * len(GROUPS))()
        new_params, new_states, keep_alive = [], [], []
        for k, (name, attr, kind) in enumerate(GROUPS):
            p = getattr(pc, attr)
            width = int(np.prod(p.shape[1:]))
            q = torch.empty((rows,) + tuple(p.shape[1:]), dtype=torch.float32, device=dev)
            arr[k].src, arr[k].dst, arr[k].width, arr[k].kind = p.data_ptr(), q.data_ptr(), width, kind
            group = _group_of(optimizer, name, p)
            stt = optimizer.state.get(group["params"][0], None) if group is not None else None
            ns = None
            if stt is not None and "exp_avg" in stt:
                m, v = stt["exp_avg"].contiguous(), stt["exp_avg_sq"].contiguous()
                nm, nv = torch.empty_like(q), torch.empty_like(q)
                arr[k].exp_avg_src, arr[k].exp_avg_dst = m.data_ptr(), nm.data_ptr()
                arr[k].exp_avg_sq_src, arr[k].exp_avg_sq_dst = v.data_ptr(), nv.data_ptr()
                keep_alive += [m, v]
                ns = dict(stt)
                ns["exp_avg"], ns["exp_avg_sq"] = nm, nv
            new_params.append(q)
            new_states.append((group, ns))
        _lib.check(L.e3_densify_apply(P, rows, selected, arr, len(GROUPS), pc._rotation.data_ptr(), pc._scaling.data_ptr(),
                                      z.data_ptr() if z is not None else None, split_n, scratch.data_ptr(), nbytes, st))
        for (name, attr, _), q, (group, ns) in zip(GROUPS, new_params, new_states):
            newp = nn.Parameter(q.requires_grad_(True))
            if group is not None:
                old = group["params"][0]
                if old in optimizer.state:
                    del optimizer.state[old]
                group["params"][0] = newp
                if ns is not None:
                    optimizer.state[newp] = ns
            setattr(pc, attr, newp)
        state.reset(rows, dev)
        # everything sized by P belongs to the old model
        for stale in ("_grad_flat", "_grad_views", "_e3_ws", "_grad_exchange"):
            if hasattr(pc, stale):
                setattr(pc, stale, None)
    return {"kept": kept, "cloned": cloned, "children_per_draw": children, "split_selected": selected, "rows": rows}


def reset_opacity(pc, optimizer, max_opacity=0.01):
    """GaussianModel.reset_opacity: in place on the parameter, the two moments of its group zeroed."""
    p = pc._opacity
    _require_cuda(p)
    group = _group_of(optimizer, "opacity", p)
    stt = optimizer.state.get(p, None) if group is not None else None
    m = v = None
    if stt is not None and "exp_avg" in stt:
        m, v = stt["exp_avg"], stt["exp_avg_sq"]
        _require_cuda(m, v)
    with torch.cuda.device(p.device), torch.no_grad():
        _lib.check(_lib.lib().e3_reset_opacity(p.shape[0], p.data_ptr(), m.data_ptr() if m is not None else None,
                                               v.data_ptr() if v is not None else None, float(np.float32(max_opacity)),
                                               _stream(p.device)))
