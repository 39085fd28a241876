"""Diagnosis: which of the two fp32 builds is closer to the truth on LARGE splats? Small scene, gradients of one
render from (a) the drop-in path, (b) the reference CUDA build, (c) the dense PyTorch transcription in float64
(autograd). Errors unweighted and weighted like the leaf gradients are (d log-scale = d scale * scale)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import torch_rasterizer as tr  # noqa: E402
from tests import gpu_common as gc  # noqa: E402
from tests.common import rel_err, scene_inputs, smooth_dL  # noqa: E402

for seed, shift in ((0, 0.0), (1, 1.0), (2, 1.5), (3, 2.0)):
    P, W, H = 1500, 96, 64
    inp = scene_inputs(P, W, H, seed=seed, mean_radius_px=4.0)
    rng = np.random.default_rng(seed)
    inp["scales"] = (inp["scales"] * np.exp(shift + rng.normal(0, 0.6, (P, 1)))).astype(np.float32)
    dL = smooth_dL(3, H, W, seed)
    fw = gc.ours_forward(inp)
    g = {k: v.cpu().numpy() for k, v in gc.ours_backward(fw, inp, dL).items()}
    ref, R, out, radii = gc.ref_forward(inp)
    gr = {k: v.cpu().numpy() for k, v in ref.backward(gc.T(dL)).items()}
    t64 = lambda a: torch.tensor(np.asarray(a, np.float64), requires_grad=True)
    m, op, sc, ro_, col = t64(inp["means3D"]), t64(inp["opacities"]), t64(inp["scales"]), t64(inp["rotations"]), t64(inp["colors"])
    img, rad, aux = tr.rasterize(m, op, sc, ro_, col, torch.tensor(inp["viewmatrix"], dtype=torch.float64).reshape(4, 4),
                                 torch.tensor(inp["projmatrix"], dtype=torch.float64).reshape(4, 4), torch.tensor(inp["bg"], dtype=torch.float64),
                                 W, H, inp["tanfovx"], inp["tanfovy"])
    (img * torch.tensor(dL, dtype=torch.float64)).sum().backward()
    truth = {"dL_dmeans3D": m.grad.numpy(), "dL_dscales": sc.grad.numpy(), "dL_drotations": ro_.grad.numpy(), "dL_dopacity": op.grad.numpy().reshape(-1, 1)}
    res = {"shift": shift, "max_radius": int(radii.max()), "img_vs_truth": "%.1e" % rel_err(fw["color"].cpu().numpy(), img.detach().numpy())}
    s = inp["scales"].astype(np.float64)
    for k in truth:
        wgt = s if k == "dL_dscales" else 1.0
        res[k] = {"ours": "%.1e" % rel_err(g[k].reshape(truth[k].shape), truth[k], 1e-30), "ref": "%.1e" % rel_err(gr[k].reshape(truth[k].shape), truth[k], 1e-30),
                  "ours_vs_ref": "%.1e" % rel_err(g[k].reshape(truth[k].shape), gr[k].reshape(truth[k].shape), 1e-30)}
        if k == "dL_dscales":
            res[k + "*scale"] = {"ours": "%.1e" % rel_err(g[k] * wgt, truth[k] * wgt, 1e-30), "ref": "%.1e" % rel_err(gr[k].reshape(truth[k].shape) * wgt, truth[k] * wgt, 1e-30),
                                 "ours_vs_ref": "%.1e" % rel_err(g[k] * wgt, gr[k].reshape(truth[k].shape) * wgt, 1e-30)}
    print(json.dumps(res))
