"""Diagnosis: the raw rasterizer backward under the EVENT LOSS's own seeds (spiky: 1 / (luma + 1e-8)) for a
fused_sweep scene: drop-in path vs reference build vs fp64-accumulating C oracle, per view."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_3dgs_b200 import pipeline as pl, synthetic as syn  # noqa: E402
from oracle import raster_oracle as ro  # noqa: E402
from tests import gpu_common as gc  # noqa: E402
from tests.common import oracle_forward, rel_err  # noqa: E402

KEYS = ["dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dscales", "dL_drotations"]
dev = torch.device("cuda", 0)


class Capture:
    last = None

    def __init__(self, raster_settings):
        self.s = raster_settings

    def __call__(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None):
        Capture.last = dict(means3D=means3D, opacities=opacities, colors=colors_precomp, scales=scales, rotations=rotations, s=self.s)
        from event_3dgs_b200.rasterizer import GaussianRasterizer
        return GaussianRasterizer(self.s)(means3D=means3D, means2D=means2D, opacities=opacities, colors_precomp=colors_precomp, scales=scales, rotations=rotations)


for seed in [int(x) for x in sys.argv[1:]]:
    rng = np.random.default_rng(9000 + seed)
    D, deblur = int(rng.integers(0, 4)), bool(rng.integers(0, 2))
    cfg = dict(P=int(rng.integers(50, 30_000)), W=int(rng.integers(17, 400)), H=int(rng.integers(17, 300)), sh_degree=D, deblur=deblur)
    sc = syn.make_scene(cfg, seed=seed)
    g = sc["gaussians"]
    g["scaling"] = (g["scaling"] + rng.normal(0, 0.6, (cfg["P"], 1)) + rng.uniform(-0.5, 1.0)).astype(np.float32)
    cams = [pl.make_camera(c, dev) for c in sc["cameras"][:6]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in sc["gt"].items()}
    bg = torch.tensor(rng.uniform(0.2, 1, 3).astype(np.float32), device=dev)
    pc = pl.GaussianParams(g, dev)
    c = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    v = int(rng.integers(0, 4))
    caps, imgs = [], []
    with torch.no_grad():
        for cam in (cams[v], cams[v], cams[v + 1]):
            pkg = pl.render(cam, pc, pl.PIPE, bg, rasterizer_cls=Capture)
            caps.append(Capture.last)
            imgs.append(pkg["render"].clone().requires_grad_(True))
    loss = pl.event_loss(imgs[0], imgs[1], imgs[2], gt["event_now"], gt["event_next"], gt["intensity"], c, gt["blurry"] if deblur else None)
    loss.backward()
    for k, (cap, img) in enumerate(zip(caps, imgs)):
        s = cap["s"]
        seedimg = img.grad.cpu().numpy()
        inp = dict(means3D=cap["means3D"].detach().cpu().numpy(), opacities=cap["opacities"].detach().cpu().numpy(), scales=cap["scales"].detach().cpu().numpy(),
                   rotations=cap["rotations"].detach().cpu().numpy(), colors=cap["colors"].detach().cpu().numpy(), shs=None, viewmatrix=s.viewmatrix.cpu().numpy(),
                   projmatrix=s.projmatrix.cpu().numpy(), campos=s.campos.cpu().numpy(), bg=bg.cpu().numpy(), W=s.image_width, H=s.image_height,
                   tanfovx=s.tanfovx, tanfovy=s.tanfovy)
        inp = {kk: vv for kk, vv in inp.items() if vv is not None}
        fw = gc.ours_forward(inp)
        go_ = {kk: vv.cpu().numpy() for kk, vv in gc.ours_backward(fw, inp, seedimg).items()}
        ref, R, out, radii = gc.ref_forward(inp)
        gr = {kk: vv.cpu().numpy() for kk, vv in ref.backward(gc.T(seedimg)).items()}
        o = oracle_forward(inp)
        gt_ = ro.rasterize_backward(o, seedimg)
        res = {"seed": seed, "view": k, "seed_max_over_median": float(np.abs(seedimg).max() / max(np.median(np.abs(seedimg)), 1e-30))}
        sw = inp["scales"].astype(np.float64)
        for kk in KEYS:
            sh = gt_[kk].shape
            w = sw if kk == "dL_dscales" else 1.0
            res[kk] = {"ours_vs_oracle": "%.1e" % rel_err(go_[kk].reshape(sh) * w, gt_[kk] * w, 1e-30), "ref_vs_oracle": "%.1e" % rel_err(gr[kk].reshape(sh) * w, gt_[kk] * w, 1e-30),
                       "ours_vs_ref": "%.1e" % rel_err(go_[kk].reshape(sh) * w, gr[kk].reshape(sh) * w, 1e-30)}
        print(json.dumps(res))
        ref.close()
