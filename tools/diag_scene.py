"""Diagnosis helper: one render of a fused_sweep scene through (a) the drop-in path, (b) the reference CUDA build,
(c) the fp64-accumulating C oracle; prints the pairwise errors per gradient array."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_3dgs_b200 import synthetic as syn  # noqa: E402
from oracle import raster_oracle as ro  # noqa: E402
from tests import gpu_common as gc  # noqa: E402
from tests.common import oracle_forward, rel_err, smooth_dL  # noqa: E402

KEYS = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dscales", "dL_drotations"]


def scene(seed):
    rng = np.random.default_rng(9000 + seed)
    D, deblur = int(rng.integers(0, 4)), bool(rng.integers(0, 2))
    cfg = dict(P=int(rng.integers(50, 30_000)), W=int(rng.integers(17, 400)), H=int(rng.integers(17, 300)), sh_degree=D, deblur=deblur)
    sc = syn.make_scene(cfg, seed=seed)
    g = sc["gaussians"]
    g["scaling"] = (g["scaling"] + rng.normal(0, 0.6, (cfg["P"], 1)) + rng.uniform(-0.5, 1.0)).astype(np.float32)
    a = syn.activate(g)
    v = int(np.random.default_rng(9000 + seed).integers(0, 4))
    cam = sc["cameras"][v]
    colors = np.clip(a["shs"][:, 0, :] * syn.SH_C0 + 0.5, 0, None).astype(np.float32)
    return dict(means3D=a["means3D"], opacities=a["opacities"], scales=a["scales"], rotations=a["rotations"], shs=a["shs"], colors=colors,
                viewmatrix=cam["world_view_transform"], projmatrix=cam["full_proj_transform"], campos=cam["camera_center"],
                bg=np.zeros(3, np.float32), W=cfg["W"], H=cfg["H"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"]), cfg


for seed in [int(x) for x in sys.argv[1:]]:
    inp, cfg = scene(seed)
    dL = smooth_dL(3, cfg["H"], cfg["W"], seed)
    fw = gc.ours_forward(inp)
    g = {k: v.cpu().numpy() for k, v in gc.ours_backward(fw, inp, dL).items()}
    ref, R, out, radii = gc.ref_forward(inp)
    gr = {k: v.cpu().numpy() for k, v in ref.backward(gc.T(dL)).items()}
    gr2 = {k: v.cpu().numpy() for k, v in ref.backward(gc.T(dL)).items()}
    o = oracle_forward(inp)
    go = ro.rasterize_backward(o, dL)
    res = {"seed": seed, "cfg": cfg, "R": R, "max_radius": int(radii.max()), "image_equal": bool(torch.equal(fw["color"].view(torch.int32), out.view(torch.int32)))}
    for k in KEYS:
        sh = go[k].shape
        res[k] = {"ours_vs_oracle": "%.1e" % rel_err(g[k].reshape(sh), go[k], 1e-30), "ref_vs_oracle": "%.1e" % rel_err(gr[k].reshape(sh), go[k], 1e-30),
                  "ours_vs_ref": "%.1e" % rel_err(g[k].reshape(sh), gr[k].reshape(sh), 1e-30), "ref_vs_ref": "%.1e" % rel_err(gr2[k].reshape(sh), gr[k].reshape(sh), 1e-30)}
    print(json.dumps(res))
