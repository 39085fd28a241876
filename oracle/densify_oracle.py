"""TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu legs): restatement, in
plain torch on whatever device the inputs live on, of the reference's adaptive density control.
Never imported by the product path.

Follows scene/gaussian_model.py step by step, INCLUDING its three rounds of concatenate / mask
(which the CUDA path collapses into one compaction), so that a disagreement in row order shows up:
  densify_and_prune        :389-403   grads = accum / denom, NaN -> 0; clone; split; prune
  densify_and_clone        :374-387
  densify_and_split        :349-372   (the normal draw is an input: samples = z * stds)
  densification_postfix    :329-347   new rows appended, statistics zeroed at the new size
  cat_tensors_to_optimizer :307-327   moments of new rows = 0
  prune_points / _prune_optimizer :273-305
  reset_opacity            :210-213, inverse_sigmoid utils/general_utils.py:18-19
  add_densification_stats  :405-407 + train.py:319
  build_rotation           utils/general_utils.py:78-99

Pinned by tests/golden/densify_*.npz, which tests/golden/make_golden_densify.py produced by running the
reference's own GaussianModel methods (imported from /root/reference, tensors kept on the CPU).
"""
import torch

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def rotation_matrices(q_raw):
    n = torch.sqrt(q_raw[:, 0] * q_raw[:, 0] + q_raw[:, 1] * q_raw[:, 1] + q_raw[:, 2] * q_raw[:, 2] + q_raw[:, 3] * q_raw[:, 3])
    q = q_raw / n[:, None]
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.zeros(q.shape[0], 3, 3, dtype=q.dtype, device=q.device)
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r * z)
    R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y)
    R[:, 2, 1] = 2 * (y * z + r * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class Model:
    """params / exp_avg / exp_avg_sq: dicts over NAMES of [P, ...] tensors (moments may be None)."""

    def __init__(self, params, exp_avg=None, exp_avg_sq=None, percent_dense=0.01):
        self.p = {k: params[k].clone() for k in NAMES}
        self.m = None if exp_avg is None else {k: exp_avg[k].clone() for k in NAMES}
        self.v = None if exp_avg_sq is None else {k: exp_avg_sq[k].clone() for k in NAMES}
        self.percent_dense = percent_dense
        P, dev = self.p["xyz"].shape[0], self.p["xyz"].device
        self.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
        self.denom = torch.zeros(P, 1, device=dev)
        self.max_radii2D = torch.zeros(P, device=dev)

    # ---- row bookkeeping ----
    def _append(self, new):
        for k in NAMES:
            self.p[k] = torch.cat((self.p[k], new[k]), dim=0)
            if self.m is not None:
                self.m[k] = torch.cat((self.m[k], torch.zeros_like(new[k])), dim=0)
                self.v[k] = torch.cat((self.v[k], torch.zeros_like(new[k])), dim=0)
        P, dev = self.p["xyz"].shape[0], self.p["xyz"].device
        self.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
        self.denom = torch.zeros(P, 1, device=dev)
        self.max_radii2D = torch.zeros(P, device=dev)

    def _drop(self, mask):
        keep = ~mask
        for k in NAMES:
            self.p[k] = self.p[k][keep]
            if self.m is not None:
                self.m[k] = self.m[k][keep]
                self.v[k] = self.v[k][keep]
        self.xyz_gradient_accum = self.xyz_gradient_accum[keep]
        self.denom = self.denom[keep]
        self.max_radii2D = self.max_radii2D[keep]

    # ---- the reference's steps ----
    def clone_step(self, grads, grad_threshold, scene_extent):
        scale_max = torch.max(torch.exp(self.p["scaling"]), dim=1).values
        sel = (torch.norm(grads, dim=-1) >= grad_threshold) & (scale_max <= self.percent_dense * scene_extent)
        self._append({k: self.p[k][sel] for k in NAMES})
        return int(sel.sum())

    def split_step(self, grads, grad_threshold, scene_extent, z, N=2):
        P = self.p["xyz"].shape[0]
        padded = torch.zeros(P, device=grads.device)
        padded[:grads.shape[0]] = grads.squeeze()
        scale_max = torch.max(torch.exp(self.p["scaling"]), dim=1).values
        sel = (padded >= grad_threshold) & (scale_max > self.percent_dense * scene_extent)
        n_sel = int(sel.sum())
        stds = torch.exp(self.p["scaling"])[sel].repeat(N, 1)
        samples = z.to(stds) * stds                                   # torch.normal(mean=0, std=stds)
        rots = rotation_matrices(self.p["rotation"][sel]).repeat(N, 1, 1)
        new = {k: self.p[k][sel].repeat(N, *([1] * (self.p[k].dim() - 1))) for k in NAMES}
        new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.p["xyz"][sel].repeat(N, 1)
        new["scaling"] = torch.log(torch.exp(self.p["scaling"])[sel].repeat(N, 1) / (0.8 * N))
        self._append(new)
        self._drop(torch.cat((sel, torch.zeros(N * n_sel, dtype=torch.bool, device=sel.device))))
        return n_sel

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, z):
        """z: standard-normal draws [2 * n_selected, 3] (or a callable n_rows -> tensor)."""
        grads = self.xyz_gradient_accum / self.denom
        grads[grads.isnan()] = 0.0
        n_clone = self.clone_step(grads, max_grad, extent)
        # the selection of the split step is needed to size z: same expression as split_step
        P = self.p["xyz"].shape[0]
        padded = torch.zeros(P, device=grads.device)
        padded[:grads.shape[0]] = grads.squeeze()
        n_sel = int(((padded >= max_grad) & (torch.max(torch.exp(self.p["scaling"]), dim=1).values > self.percent_dense * extent)).sum())
        zz = z(2 * n_sel) if callable(z) else z
        if zz is None:
            zz = torch.zeros(2 * n_sel, 3)
        self.split_step(grads, max_grad, extent, zz[:2 * n_sel].to(grads.device))
        prune = (torch.sigmoid(self.p["opacity"]) < min_opacity).squeeze(-1)
        if max_screen_size:
            big_vs = self.max_radii2D > max_screen_size
            big_ws = torch.exp(self.p["scaling"]).max(dim=1).values > 0.1 * extent
            prune = prune | big_vs | big_ws
        self._drop(prune)
        return n_clone, n_sel

    def reset_opacity(self):
        x = torch.min(torch.sigmoid(self.p["opacity"]), torch.ones_like(self.p["opacity"]) * 0.01)
        self.p["opacity"] = torch.log(x / (1 - x))
        if self.m is not None:
            self.m["opacity"] = torch.zeros_like(self.p["opacity"])
            self.v["opacity"] = torch.zeros_like(self.p["opacity"])

    def add_densification_stats(self, viewspace_grad, radii):
        vis = radii > 0
        self.max_radii2D[vis] = torch.max(self.max_radii2D[vis], radii[vis].to(self.max_radii2D.dtype))
        self.xyz_gradient_accum[vis] += torch.norm(viewspace_grad[vis, :2], dim=-1, keepdim=True)
        self.denom[vis] += 1
