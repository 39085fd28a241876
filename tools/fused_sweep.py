"""One-off stress of the fused event iteration against the REFERENCE route (reference CUDA kernels under the
reference-shaped render()/loss autograd flow) over many random small scenes: SH degree, --deblur, image shape, splat
sizes and backgrounds vary. Reports the worst error per quantity and how many scenes leave BASELINE.md section 5.

    python tools/fused_sweep.py [n_scenes]        (GPU box; prints one JSON line)"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_3dgs_b200 import fused, pipeline as pl, synthetic as syn  # noqa: E402
from oracle import ref_cuda  # noqa: E402
from tests.common import grad_tol, rel_err  # noqa: E402

NAMES = ["xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity", "c"]


def main(n=100):
    dev = torch.device("cuda", 0)
    worst = {k: 0.0 for k in NAMES + ["loss", "viewspace_rows_off", "pixels_off"]}
    failed, unexplained, worst_ratio, worst_dropin = [], [], {}, {}

    def render_ref(cam, pc_, pipe, bg_, **kw):
        return pl.render(cam, pc_, pipe, bg_, rasterizer_cls=ref_cuda.RefGaussianRasterizer)
    for seed in range(n):
        rng = np.random.default_rng(9000 + seed)
        D, deblur = int(rng.integers(0, 4)), bool(rng.integers(0, 2))
        cfg = dict(P=int(rng.integers(50, 30_000)), W=int(rng.integers(17, 400)), H=int(rng.integers(17, 300)), sh_degree=D, deblur=deblur)
        sc = syn.make_scene(cfg, seed=seed)
        g = sc["gaussians"]
        g["scaling"] = (g["scaling"] + rng.normal(0, 0.6, (cfg["P"], 1)) + rng.uniform(-0.5, 1.0)).astype(np.float32)
        cams = [pl.make_camera(c, dev) for c in sc["cameras"][:6]]
        gt = {k: torch.as_tensor(v).to(dev) for k, v in sc["gt"].items()}
        # E3_SWEEP_DARK_BG=1: black background for two thirds of the scenes. The event loss takes log(Y + 1e-8) of the
        # rendered luma, so on a sparse scene over a black background the seed at a pixel covered by a single faint splat
        # is ~1e6 times a normal one, and any last-bit difference in that splat's alpha decides whole gradients (the
        # reference disagrees with a 1-ulp-perturbed copy of itself there). Default: backgrounds >= 0.2.
        if os.environ.get("E3_SWEEP_DARK_BG") and seed % 3:
            bg = torch.zeros(3, device=dev)
        else:
            bg = torch.tensor(rng.uniform(0.2, 1, 3).astype(np.float32), device=dev)
        pc_r, pc = pl.GaussianParams(g, dev), pl.GaussianParams(g, dev)
        c_r = torch.nn.Parameter(torch.tensor(0.17, device=dev))
        c = torch.nn.Parameter(torch.tensor(0.17, device=dev))
        v = int(rng.integers(0, 4))
        # the reference route twice: its own run-to-run spread (float atomics) on this very input
        loss_r, pkg_r = pl.event_iteration(pc_r, cams[v], cams[v], cams[v + 1], gt, c_r, bg, deblur=deblur, render_fn=render_ref)
        loss_r.backward()
        first = [q.grad.clone() for q in pc_r.leaves()]
        for q in pc_r.leaves() + [c_r]:
            q.grad = None
        loss_r, pkg_r = pl.event_iteration(pc_r, cams[v], cams[v], cams[v + 1], gt, c_r, bg, deblur=deblur, render_fn=render_ref)
        loss_r.backward()
        ref_spread = {nme: rel_err(a.cpu().numpy(), q.grad.cpu().numpy(), 1e-30) for nme, a, q in zip(NAMES, first, pc_r.leaves())}
        # the same route over OUR drop-in rasterizer: separates the blend / preprocess kernels from the fused path's own parts
        pc_d = pl.GaussianParams(g, dev)
        c_d = torch.nn.Parameter(torch.tensor(0.17, device=dev))
        loss_d, _ = pl.event_iteration(pc_d, cams[v], cams[v], cams[v + 1], gt, c_d, bg, deblur=deblur)
        loss_d.backward()
        for nme, p_, q_ in zip(NAMES, pc_d.leaves(), pc_r.leaves()):
            if nme == "features_rest" and D == 0:
                continue
            worst_dropin[nme] = max(worst_dropin.get(nme, 0.0), rel_err(p_.grad.cpu().numpy(), q_.grad.cpu().numpy(), 1e-30))
        for call in range(2):
            for p in pc.leaves() + [c]:
                p.grad = None
            loss, pkg = fused.event_iteration_backward(pc, cams[v], cams[v], cams[v + 1], gt, c, bg, deblur=deblur, return_pkg=True,
                                                       gray_event_views=bool(seed % 2), dedupe_views=(seed % 5 == 0))
        ok = True
        e = abs(float(loss) - float(loss_r)) / max(abs(float(loss_r)), 1e-12)
        worst["loss"] = max(worst["loss"], e)
        ok &= e <= 1e-5
        same = (pkg["radii"] == pkg_r["radii"]).cpu().numpy()
        for nme, p, q in zip(NAMES, pc.leaves() + [c], pc_r.leaves() + [c_r]):
            a, b = p.grad.cpu().numpy(), q.grad.cpu().numpy()
            if a.ndim and a.shape[0] == same.shape[0]:
                a, b = a[same], b[same]
            if nme == "features_rest" and D == 0:
                continue
            e = rel_err(a, b, 1e-30)
            worst[nme] = max(worst[nme], e)
            ok &= e < grad_tol(nme, two_atomic_runs=True)
            # beyond the table only where the reference does not reproduce ITSELF either: count those separately
            if e >= grad_tol(nme, two_atomic_runs=True):
                ratio = e / max(ref_spread.get(nme, 0.0), 1e-30)
                worst_ratio[nme] = max(worst_ratio.get(nme, 0.0), ratio)
                if ratio > 5.0:
                    unexplained.append((seed, nme, float(e), float(ref_spread.get(nme, 0.0))))
        va, vb = pkg["viewspace_points"].grad.cpu().numpy()[same], pkg_r["viewspace_points"].grad.cpu().numpy()[same]
        rows = float((np.abs(va - vb).max(axis=1) >= 1e-4 * max(np.abs(vb).max(), 1e-30)).mean()) if va.size else 0.0
        worst["viewspace_rows_off"] = max(worst["viewspace_rows_off"], rows)
        d = (pkg["render"] - pkg_r["render"]).abs()
        px = float((d > 1e-5 * max(float(pkg_r["render"].abs().max()), 1e-30)).float().mean())
        worst["pixels_off"] = max(worst["pixels_off"], px)
        if not ok:
            failed.append(seed)
            if os.environ.get("E3_SWEEP_VERBOSE"):
                errs = {}
                for nme, p, q in zip(NAMES, pc.leaves() + [c], pc_r.leaves() + [c_r]):
                    a, b = p.grad.cpu().numpy(), q.grad.cpu().numpy()
                    errs[nme] = "%.1e" % rel_err(a, b, 1e-30)
                print(seed, cfg, "bg" if seed % 3 == 0 else "", "luma" if seed % 2 else "rgb", "dedupe" if seed % 5 == 0 else "",
                      "loss %.1e" % (abs(float(loss) - float(loss_r)) / abs(float(loss_r))), "radii!=", int((~same).sum()), "px_off %.1e" % px,
                      "maxdiff %.1e" % float(d.max()), errs, file=sys.stderr)
    print(json.dumps({"scenes": n, "worst": worst, "scenes_outside_the_tolerance_table": failed, "worst_dropin_route_vs_reference_route": worst_dropin,
                      "worst_error_over_reference_self_spread_where_outside": worst_ratio,
                      "outside_and_more_than_5x_the_reference_self_spread": unexplained}))


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 100)
