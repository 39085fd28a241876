"""FusedAdam: torch.optim.Adam-compatible optimizer whose step() is ONE kernel launch over all
parameter groups (e3_adam_step). Mirrors what the reference builds in
scene/gaussian_model.py:149-175 (six groups, lr per group, eps = 1e-15) and steps in
train.py:330-332. State layout (state[p] = {"step", "exp_avg", "exp_avg_sq"}) and param_groups
are torch's, so the densification code that edits optimizer state by hand
(scene/gaussian_model.py:258-340) keeps working on it."""

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    def moments_are_zero(self, p):
        """True while p has only ever been advanced by advance_without_update (or never stepped)."""
        st = self.state.get(p, None)
        return st is None or len(st) == 0 or bool(st.get("_moments_zero", False))

    @torch.no_grad()
    def advance_without_update(self, p):
        """For a parameter whose gradient is identically zero and whose moments are zero: the Adam
        update is the identity (m = v = 0, p -= lr * 0 / (0 + eps)); only the step counter moves."""
        if not self.moments_are_zero(p):
            raise RuntimeError("advance_without_update: the moments of this parameter are not zero")
        st = self.state[p]
        if len(st) == 0:
            st["step"] = 0
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        st["_moments_zero"] = True
        st["step"] = int(st["step"]) + 1

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # parameters that share (step, betas) go into one launch
        buckets = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.grad.is_contiguous()):
                    raise RuntimeError("FusedAdam: parameters and gradients must be contiguous float32 CUDA tensors")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] = int(st["step"]) + 1
                st.pop("_moments_zero", None)
                key = (st["step"], float(b1), float(b2), p.device)
                buckets.setdefault(key, []).append((p, st, float(group["lr"]), float(group["eps"])))
        L = _lib.lib()
        for (t, b1, b2, dev), items in buckets.items():
            stream = torch.cuda.current_stream(dev).cuda_stream
            for i in range(0, len(items), 8):
                chunk = items[i:i + 8]
                arr = (_lib.AdamGroup * len(chunk))()
                for k, (p, st, lr, eps) in enumerate(chunk):
                    arr[k].param, arr[k].grad = p.data_ptr(), p.grad.data_ptr()
                    arr[k].exp_avg, arr[k].exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                    arr[k].n, arr[k].lr, arr[k].eps = p.numel(), lr, eps
                with torch.cuda.device(dev):
                    _lib.check(L.e3_adam_step(arr, len(chunk), t, b1, b2, stream))
        return loss
