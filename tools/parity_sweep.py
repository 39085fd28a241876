"""One-off stress of the bit-exactness claims: many random scenes (sizes, image shapes, backgrounds, scale
distributions from sub-pixel to frame-filling splats, strongly anisotropic ones, points at / behind the camera) through
the drop-in path and through the reference CUDA build (oracle/_ref); counts scenes in which num_rendered, radii,
point_list, ranges, n_contrib or the image differ in a single bit. The sub-tile masks (exact ellipse test with margins)
are the part of the blend that could, if a margin were too tight, drop a contribution the reference blends.

    python tools/parity_sweep.py [n_scenes]        (GPU box; prints one JSON line)"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import gpu_common as gc  # noqa: E402
from tests.common import scene_inputs  # noqa: E402


def main(n=200):
    bad, total_R = [], 0
    for seed in range(n):
        rng = np.random.default_rng(5000 + seed)
        P = int(rng.integers(1, 20_000))
        W, H = int(rng.integers(17, 500)), int(rng.integers(17, 400))
        inp = scene_inputs(P, W, H, seed=seed, bg=tuple(rng.uniform(0, 1, 3)), mean_radius_px=float(rng.uniform(0.5, 25.0)))
        s = inp["scales"].copy()
        k = max(1, P // 20)
        s[rng.integers(0, P, k)] *= 30.0
        s[rng.integers(0, P, k)] *= 0.01
        a = rng.integers(0, P, k)                       # needles: one axis 100x the others
        s[a, rng.integers(0, 3, k)] *= 100.0
        inp["scales"] = s
        op = inp["opacities"].copy()
        op[rng.integers(0, P, k)] = rng.uniform(0.0039, 0.0045, (k, 1)).astype(np.float32)     # alpha barely above 1/255 at the centre
        op[rng.integers(0, P, k)] = 0.9999
        inp["opacities"] = op
        m = inp["means3D"].copy()
        cam_dir = -inp["campos"] / np.linalg.norm(inp["campos"])
        idx = rng.integers(0, P, k)
        m[idx] = inp["campos"][None, :] + rng.uniform(-0.5, 0.6, (k, 1)).astype(np.float32) * cam_dir[None, :] + \
            rng.normal(0, 0.2, (k, 3)).astype(np.float32)
        inp["means3D"] = m.astype(np.float32)
        fw = gc.ours_forward(inp)
        st = gc.ours_state(fw, inp)
        ref, R, out, radii = gc.ref_forward(inp)
        ok = (fw["R"] == R and torch.equal(fw["radii"], radii) and torch.equal(st["point_list"], ref.state("point_list")) and
              torch.equal(st["ranges"].reshape(-1), ref.state("ranges")) and torch.equal(st["n_contrib"], ref.state("n_contrib")) and
              torch.equal(st["final_T"].view(torch.int32), ref.state("accum_alpha").view(torch.int32)) and
              torch.equal(fw["color"].view(torch.int32), out.view(torch.int32)))
        total_R += R
        if not ok:
            bad.append(seed)
        ref.close()
    print(json.dumps({"scenes": n, "instances": total_R, "scenes_with_any_bit_difference": bad}))


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 200)
