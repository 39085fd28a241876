"""Generates tests/golden/densify_*.npz by running the REFERENCE's own GaussianModel methods
(scene/gaussian_model.py imported from /root/reference) on CPU tensors. Build container only; the
GPU box never runs this.

The reference hard-codes device="cuda" in its torch.zeros calls and draws with torch.normal; both
are redirected here WITHOUT touching the reference's code: torch.zeros is wrapped to drop the
device keyword, torch.normal is wrapped to return mean + std * z for a recorded standard-normal z
(which is what ATen computes), so that the draw can be replayed by the oracle and by the CUDA path.
plyfile and simple_knn are imported by the module but not used on this path; they are stubbed.

    python tests/golden/make_golden_densify.py
"""
import os
import sys
import types
from types import SimpleNamespace

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

ply = types.ModuleType("plyfile")
ply.PlyData = ply.PlyElement = object
sys.modules.setdefault("plyfile", ply)
knn = types.ModuleType("simple_knn")
knn_c = types.ModuleType("simple_knn._C")
knn_c.distCUDA2 = None
knn._C = knn_c
sys.modules.setdefault("simple_knn", knn)
sys.modules.setdefault("simple_knn._C", knn_c)

_zeros = torch.zeros


def zeros_on_cpu(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


torch.zeros = zeros_on_cpu
DRAWS = []


def replayable_normal(mean, std, **_):
    z = torch.randn(std.shape, generator=torch.Generator().manual_seed(1234 + len(DRAWS)))
    DRAWS.append(z)
    return mean + std * z


torch.normal = replayable_normal

from scene.gaussian_model import GaussianModel  # noqa: E402

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", scaling="_scaling", rotation="_rotation")


def build_model(seed, P, extent):
    g = torch.Generator().manual_seed(seed)
    gm = GaussianModel(3)
    rnd = lambda *s: torch.randn(*s, generator=g)
    gm._xyz = torch.nn.Parameter(rnd(P, 3))
    gm._features_dc = torch.nn.Parameter(rnd(P, 1, 3) * 0.5)
    gm._features_rest = torch.nn.Parameter(rnd(P, 15, 3) * 0.1)
    # raw scales straddling percent_dense * extent (clone vs split) and 0.1 * extent (world-size prune)
    gm._scaling = torch.nn.Parameter(torch.log(torch.tensor(0.01 * extent)) - 0.8 + rnd(P, 3) * 1.2)
    gm._rotation = torch.nn.Parameter(rnd(P, 4))
    gm._opacity = torch.nn.Parameter(rnd(P, 1) * 3.0 - 2.0)          # some below sigmoid^-1(0.005)
    gm.max_radii2D = torch.zeros(P)
    args = SimpleNamespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016,
                           position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025, opacity_lr=0.05,
                           scaling_lr=0.005, rotation_lr=0.001)
    gm.spatial_lr_scale = 1.0
    gm.training_setup(args)
    # two Adam steps with random gradients so that the moments are populated
    for _ in range(2):
        for k in NAMES:
            p = getattr(gm, ATTR[k])
            p.grad = torch.randn(p.shape, generator=g) * 0.01
        gm.optimizer.step()
        gm.optimizer.zero_grad(set_to_none=True)
    return gm, g


def snapshot(gm, prefix):
    out = {}
    for k in NAMES:
        p = getattr(gm, ATTR[k])
        st = gm.optimizer.state[p]
        out[prefix + k] = p.detach().numpy().copy()
        out[prefix + "m_" + k] = st["exp_avg"].numpy().copy()
        out[prefix + "v_" + k] = st["exp_avg_sq"].numpy().copy()
    out[prefix + "accum"] = gm.xyz_gradient_accum.numpy().copy()
    out[prefix + "denom"] = gm.denom.numpy().copy()
    out[prefix + "max_radii2D"] = gm.max_radii2D.numpy().copy()
    return out


def case(seed, P, extent, max_screen_size, iters, poke_inf=0):
    DRAWS.clear()
    gm, g = build_model(seed, P, extent)
    out = snapshot(gm, "in_")
    # train.py:317-320, `iters` times with fresh renders' radii / screen-space gradients
    radii_all, grad_all = [], []
    for _ in range(iters):
        radii = torch.randint(-20, 30, (P,), generator=g, dtype=torch.int32).clamp_min(0)
        vsp = SimpleNamespace(grad=torch.randn(P, 3, generator=g) * 0.00016)
        vis = radii > 0
        gm.max_radii2D[vis] = torch.max(gm.max_radii2D[vis], radii[vis])
        gm.add_densification_stats(vsp, vis)
        radii_all.append(radii.numpy())
        grad_all.append(vsp.grad.numpy())
    out["radii"] = np.stack(radii_all)
    out["viewspace_grad"] = np.stack(grad_all)
    if poke_inf:
        # rows whose gradient norm was accumulated but whose visit count is zero: accum / denom = inf, which the
        # reference neither masks nor clamps (only NaN is zeroed, gaussian_model.py:391) -> they are selected
        seen = torch.nonzero(gm.xyz_gradient_accum[:, 0] > 0)[:poke_inf, 0]
        gm.denom[seen] = 0
    out.update({k.replace("in_", "stats_"): v for k, v in snapshot(gm, "in_").items() if k in ("in_accum", "in_denom", "in_max_radii2D")})
    gm.densify_and_prune(0.0002, 0.005, extent, max_screen_size)
    out.update(snapshot(gm, "out_"))
    out["z"] = DRAWS[0].numpy() if DRAWS else np.zeros((0, 3), np.float32)
    gm.reset_opacity()
    p = gm._opacity
    out["reset_opacity"] = p.detach().numpy().copy()
    out["reset_m"] = gm.optimizer.state[p]["exp_avg"].numpy().copy()
    out["extent"] = np.float32(extent)
    out["max_screen_size"] = np.int32(max_screen_size or 0)
    out["poke_inf"] = np.int32(poke_inf)
    return out


def init_case(seed, P):
    """GaussianModel.create_from_pcd (scene/gaussian_model.py:124-147) run by the reference itself. Its
    distCUDA2 (simple-knn, not in the tree) is replaced by an exact brute-force 3-NN in float64 -- the contract
    of that function --, Tensor.cuda by the identity."""
    import scene.gaussian_model as gmod
    from utils.graphics_utils import BasicPointCloud

    def dist2_bruteforce(pts):
        p = pts.double()
        d = torch.cdist(p, p) ** 2
        d.fill_diagonal_(float("inf"))
        return d.topk(3, largest=False).values.mean(1).float()
    gmod.distCUDA2 = dist2_bruteforce
    cuda, ones = torch.Tensor.cuda, torch.ones
    torch.Tensor.cuda = lambda self, *a, **k: self

    def ones_on_cpu(*a, **k):
        k.pop("device", None)
        return ones(*a, **k)
    torch.ones = ones_on_cpu
    try:
        g = torch.Generator().manual_seed(seed)
        pts = (torch.rand(P, 3, generator=g) * 2.6 - 1.3).numpy()
        cols = torch.rand(P, 3, generator=g).numpy()
        gm = GaussianModel(3)
        import io
        import contextlib
        with contextlib.redirect_stdout(io.StringIO()):
            gm.create_from_pcd(BasicPointCloud(points=pts, colors=cols, normals=np.zeros_like(pts)), 1.0)
    finally:
        torch.Tensor.cuda, torch.ones = cuda, ones
    return dict(points=pts, colors=cols, xyz=gm._xyz.detach().numpy(), features_dc=gm._features_dc.detach().numpy(),
                features_rest=gm._features_rest.detach().numpy(), scaling=gm._scaling.detach().numpy(),
                rotation=gm._rotation.detach().numpy(), opacity=gm._opacity.detach().numpy())


if __name__ == "__main__":
    for name, kw in (("densify_a", dict(seed=1, P=300, extent=4.0, max_screen_size=None, iters=3)),
                     ("densify_b", dict(seed=2, P=400, extent=2.5, max_screen_size=20, iters=2)),
                     ("densify_c", dict(seed=3, P=200, extent=3.0, max_screen_size=20, iters=1, poke_inf=25))):
        d = case(**kw)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
        print(name, "P", kw["P"], "->", d["out_xyz"].shape[0], "rows; split draws", d["z"].shape[0])
    np.savez_compressed(os.path.join(HERE, "init_pcd.npz"), **init_case(7, 1500))
    print("init_pcd written")
