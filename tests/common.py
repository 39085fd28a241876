"""Shared helpers of the test-suite (not collected)."""
import numpy as np

from event_3dgs_b200 import synthetic as syn


def scene_inputs(P, W, H, seed=0, kind="trained", view=3, channels=3, bg=(0.0, 0.0, 0.0), mean_radius_px=4.0):
    """numpy inputs of one rasterizer call with precomputed colours."""
    g = syn.make_gaussians(P, W, seed=seed, kind=kind, mean_radius_px=mean_radius_px)
    a = syn.activate(g)
    cam = syn.make_camera(view, 96, W, H)
    rng = np.random.default_rng(seed + 7)
    colors = rng.uniform(0, 1, (P, channels)).astype(np.float32)
    return dict(means3D=a["means3D"], opacities=a["opacities"], scales=a["scales"], rotations=a["rotations"],
                shs=a["shs"], colors=colors, viewmatrix=cam["world_view_transform"],
                projmatrix=cam["full_proj_transform"], campos=cam["camera_center"],
                bg=np.asarray(bg, np.float32), W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])


def smooth_dL(C, H, W, seed=0):
    rng = np.random.default_rng(seed + 99)
    yy, xx = np.mgrid[0:H, 0:W] / float(max(H, W))
    ph = rng.uniform(0, 6.28, (C, 3))
    out = np.stack([np.sin(5 * xx + ph[c, 0]) * np.cos(4 * yy + ph[c, 1]) + 0.3 * np.sin(9 * (xx + yy) + ph[c, 2])
                    for c in range(C)])
    return out.astype(np.float32)


def oracle_forward(inp, use_sh=False, sh_degree=0, cov3D_precomp=None):
    from oracle import raster_oracle as ro
    kw = dict(scales=inp["scales"], rotations=inp["rotations"])
    if cov3D_precomp is not None:
        kw = dict(cov3D_precomp=cov3D_precomp)
    if use_sh:
        kw.update(shs=inp["shs"], sh_degree=sh_degree)
    else:
        kw.update(colors_precomp=inp["colors"])
    return ro.rasterize_forward(inp["means3D"], inp["opacities"], inp["viewmatrix"], inp["projmatrix"], inp["campos"],
                                inp["bg"], inp["W"], inp["H"], inp["tanfovx"], inp["tanfovy"], **kw)


def rel_err(a, b, floor=0.0):
    """max |a-b| / max(max |b|, floor) -- the 'rel' of BASELINE.json's tolerances (relative to the
    array scale). `floor` guards arrays that are mathematically zero (e.g. dL/drotation of
    isotropic Gaussians), where only rounding noise is left to compare."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.abs(a - b).max() if a.size else 0.0
    s = max(np.abs(b).max() if b.size else 0.0, floor)
    return d / s if s > 0 else d


def scaled_err(a, b, f=1e-2):
    """Scale-aware companion of rel_err: max_i |a_i - b_i| / (|b_i| + f * max|b|). An element as large as the
    array maximum is held relatively; an element far below it is held against f * max|b| instead of against
    max|b|, i.e. 1/f times tighter than rel_err holds it (rel_err lets rows two orders of magnitude below the
    maximum go unchecked). Bars for it are in BASELINE.md section 5."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if not a.size:
        return 0.0
    m = np.abs(b).max()
    if m == 0:
        return float(np.abs(a).max())
    return float((np.abs(a - b) / (np.abs(b) + f * m)).max())


# ---- tolerances (BASELINE.md section 5 holds the table and its justification) ----
IMG_TOL = 1e-5          # image, final_T, loss: relative to the array maximum
GRAD_TOL = 1e-4         # every gradient array against the fp64-accumulating oracle / the reference build
# Two fp32 runs whose float atomics commit in different orders (ours against the reference build, or separately
# executed iterations): the arrays behind the conic -> covariance chain are sums of per-pixel terms that cancel, and
# repeat only to ~1.2e-4 (20k Gaussians) .. 1.6e-4 (c3) in the reference build itself
# (profiles/r1_grad_run_to_run_spread.json, profiles/r2_parity_observed.json ref_vs_ref/...). Since the backward blend
# sends the terms of sparsely covered pairs to the accumulator as per-pixel vector reductions (as the reference does for
# every pair), ours repeats itself to the same level (ours_vs_ours/...), and two such runs differ by up to 3.4e-4 on
# `scaling` under the L1 + SSIM seeds (20k Gaussians, 346x260; 2 of 11 runs above 3e-4).
GRAD_TOL_ATOMIC_CHAIN = 6e-4
# one such run against the exact (fp64-accumulating) oracle: the reference build lands at up to 7e-5 on the test scenes,
# ours between 1e-5 and 7e-5 from run to run
GRAD_TOL_CHAIN_VS_EXACT = 2e-4
CHAIN_ARRAYS = ("dL_dmeans3D", "dL_dscales", "dL_drotations", "xyz", "scaling", "rotation")


# 3M Gaussians at 1920x1080: the reference build repeats ITSELF only to 6.5e-4 (dL_drotations) / 1.2e-4 (dL_dscales)
# from one run to the next there (profiles/r2_parity_observed.json, ref_vs_ref/c5): a handful of ill-conditioned rows
# out of 3M set the maximum. Two such runs against each other are held to ~4x that spread, and to GRAD_TOL on all
# but 1e-4 of the rows (row-wise check in tests/test_timed_sizes_gpu.py).
GRAD_TOL_ATOMIC_CHAIN_C5 = {"dL_drotations": 3e-3, "dL_dscales": 1e-3, "dL_dmeans3D": 3e-4}
# 1M Gaussians at 1280x720 (scene_inputs "trained"): dL_drotations repeats to 1.7e-4 in the reference build itself and
# ours-against-reference was seen between 1.5e-4 and 3.1e-4 from run to run (one or two rows out of 1M set it): same
# rule, ~4x the reference's own spread, and GRAD_TOL on all but 1e-4 of the rows. The leaf gradients of the fused
# iteration at c3 / c4 (three views summed, through the activations) were seen between 5e-5 and 1.7e-4 (scaling) over
# 48 comparisons; the reference route against itself is recorded beside them (ref_route_vs_itself/...).
GRAD_TOL_ATOMIC_CHAIN_C3 = {"dL_drotations": 1e-3, "rotation": 6e-4, "scaling": 6e-4}


def grad_tol(name, two_atomic_runs=False, size=None):
    """two_atomic_runs: both sides are fp32 runs with float atomics; otherwise ours against the fp64 oracle."""
    if not two_atomic_runs and name in CHAIN_ARRAYS:
        return GRAD_TOL_CHAIN_VS_EXACT
    if two_atomic_runs and name in CHAIN_ARRAYS:
        table = GRAD_TOL_ATOMIC_CHAIN_C5 if size == "c5" else GRAD_TOL_ATOMIC_CHAIN_C3 if size in ("c3", "c3init") else {}
        return table.get(name, GRAD_TOL_ATOMIC_CHAIN)
    return GRAD_TOL


# order of GaussianParams.leaves() (+ the event threshold c)
LEAF_NAMES = ["xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity", "c"]


OBSERVED = {}


def observe(name, value):
    """Record a measured error; tests/conftest.py writes them to gpurun_out/parity_observed.json."""
    OBSERVED[name] = float(value)
    return float(value)
