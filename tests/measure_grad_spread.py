"""Run-to-run spread of the gradients (float atomics commit in a different order every run): the same
backward pass N times, max over runs of max|g_k - g_0| / max|g_0| per gradient array, for
  ref    : the reference CUDA build (oracle/_ref), 9 global atomics per (pixel, Gaussian) pair
  ours   : e3_rasterize_backward through the drop-in binding
  fused  : the whole event iteration (3 views) through e3_event_iteration
SURVEY.md 8c asks for the parity tolerance to be reported next to this number."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # repo root
from tests import gpu_common as gc  # noqa: E402
from tests.common import rel_err, scene_inputs, smooth_dL  # noqa: E402

KEYS = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dscales", "dL_drotations"]


def spread(runs):
    out = {}
    for k in runs[0]:
        out[k] = max(rel_err(r[k], runs[0][k], 1e-30) for r in runs[1:])
    return out


def main(N=8):
    res = {}
    for name, (P, W, H) in {"20k_200x120": (20_000, 200, 120), "300k_346x260": (300_000, 346, 260)}.items():
        inp = scene_inputs(P, W, H, seed=3)
        dL = smooth_dL(3, H, W, 3)
        fw = gc.ours_forward(inp)
        ours = [{k: v.cpu().numpy() for k, v in gc.ours_backward(fw, inp, dL).items() if k in KEYS} for _ in range(N)]
        entry = {"ours": spread(ours)}
        try:
            ref, R, out, radii = gc.ref_forward(inp)
            refs = [{k: v.cpu().numpy() for k, v in ref.backward(gc.T(dL)).items() if k in KEYS} for _ in range(N)]
            entry["ref"] = spread(refs)
            entry["ours_vs_ref"] = {k: rel_err(ours[0][k].reshape(refs[0][k].shape), refs[0][k], 1e-30) for k in KEYS}
        except Exception as ex:
            entry["ref"] = {"unavailable": repr(ex)[:100]}
        res[name] = entry
    # fused event iteration
    from tests.test_fused_gpu import _setup
    from event_3dgs_b200 import fused
    pl, pc, cams, gt, bg, c = _setup(20_000, 200, 120, 1, False)
    names = ["xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity", "c"]
    runs = []
    for _ in range(N):
        fused.event_iteration_backward(pc, cams[1], cams[1], cams[2], gt, c, bg)
        runs.append({n: p.grad.detach().cpu().numpy().copy() for n, p in zip(names, pc.leaves() + [c])})
    res["fused_iteration_20k_200x120"] = spread(runs)
    print(json.dumps(res, indent=1, default=float))


if __name__ == "__main__":
    main()
