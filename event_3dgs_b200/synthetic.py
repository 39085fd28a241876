"""Seeded synthetic scenes of the BASELINE.json shapes.

Pure numpy, no GPU needed: the same arrays feed the CUDA path, the CPU oracle and the
reference-CUDA build, so every arm of a comparison sees identical inputs.

Conventions restated from the reference (not imported):
  * world_view_transform / full_proj_transform are stored TRANSPOSED, flat index c*4+r = M[r][c]
    (scene/cameras.py:54-57, utils/graphics_utils.py:38-71);
  * camera_center = inverse(world_view_transform)[3, :3] (scene/cameras.py:57);
  * znear 0.01 / zfar 100 (scene/cameras.py:48-49);
  * SH DC from colour: (rgb - 0.5) / C0 (utils/sh_utils.py:114-115).
"""
import math

import numpy as np

SH_C0 = 0.28209479177387814

CONFIGS = {
    # name: (P, W, H, sh_degree, deblur)
    "c1": dict(P=10_000, W=64, H=64, sh_degree=0, deblur=False),
    "c2": dict(P=300_000, W=346, H=260, sh_degree=0, deblur=False),
    "c3": dict(P=1_000_000, W=1280, H=720, sh_degree=0, deblur=False),
    "c4": dict(P=1_000_000, W=1280, H=720, sh_degree=3, deblur=True),
    "c5": dict(P=3_000_000, W=1920, H=1080, sh_degree=0, deblur=False),
}


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (returned un-transposed, float32)."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    P = np.zeros((4, 4), np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(i, n_views, W, H, fovx_deg=60.0, radius=4.0, height=0.6):
    """View i of n_views on a circle of `radius` looking at the origin (COLMAP axes: +z forward,
    +x right, +y down)."""
    ang = 2.0 * math.pi * i / n_views
    c = np.array([radius * math.cos(ang), height * math.sin(3 * ang), radius * math.sin(ang)], np.float64)
    fwd = -c / np.linalg.norm(c)
    up = np.array([0.0, -1.0, 0.0])
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    Rwc = np.stack([right, down, fwd], 0)            # world -> camera rotation (rows = camera axes)
    V = np.eye(4)
    V[:3, :3] = Rwc
    V[:3, 3] = -Rwc @ c
    V = V.astype(np.float32)
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * H / W)
    Pm = projection_matrix(0.01, 100.0, fovx, fovy)
    wvt = np.ascontiguousarray(V.T)                               # transposed storage
    full = np.ascontiguousarray((wvt @ Pm.T).astype(np.float32))  # cameras.py:56
    center = np.linalg.inv(wvt.astype(np.float64))[3, :3].astype(np.float32)
    # R, T in the reference's Camera convention (scene/cameras.py: R = camera-to-world rotation, T = w2c translation)
    return dict(R=np.ascontiguousarray(Rwc.T), T=(-Rwc @ c).copy(), world_view_transform=wvt, full_proj_transform=full, camera_center=center,
                FoVx=fovx, FoVy=fovy, tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5),
                image_width=W, image_height=H)


def make_gaussians(P, W, seed=0, kind="trained", sh_degree=0, gray=True, mean_radius_px=4.0):
    """Raw (pre-activation) parameters in the layout of scene/gaussian_model.py:124-147:
    _xyz [P,3], _features_dc [P,1,3], _features_rest [P,15,3], _scaling [P,3] (log),
    _rotation [P,4] (unnormalised w,x,y,z), _opacity [P,1] (logit)."""
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(-1.3, 1.3, (P, 3)).astype(np.float32)
    fx = W / (2.0 * math.tan(math.radians(60.0) / 2))
    if kind == "init":
        # isotropic, opacity 0.1, identity rotation (gaussian_model.py:134-141); the kNN
        # distance of a uniform cloud ~ (V/P)^(1/3)
        s = 0.55 * (2.6 ** 3 / max(P, 1)) ** (1.0 / 3.0)
        scaling = np.full((P, 3), math.log(s), np.float32)
        rotation = np.zeros((P, 4), np.float32)
        rotation[:, 0] = 1.0
        opacity = np.full((P, 1), math.log(0.1 / 0.9), np.float32)
        rgb = np.full((P, 3), 0.5, np.float32) if gray else rng.uniform(0, 1, (P, 3)).astype(np.float32)
        f_dc = ((rgb - 0.5) / SH_C0).astype(np.float32)[:, None, :]
        f_rest = np.zeros((P, 15, 3), np.float32)
    else:
        # radius ~ 3 sigma_px, sigma_px ~ s * fx / z with z ~ 4
        mu = math.log(mean_radius_px / 3.0 * 4.0 / fx) - 0.5 * 0.7 ** 2
        scaling = rng.normal(mu, 0.7, (P, 3)).astype(np.float32)
        rotation = rng.normal(0, 1, (P, 4)).astype(np.float32)
        opacity = rng.normal(0, 2, (P, 1)).astype(np.float32)
        f_dc = rng.normal(0, 0.5, (P, 1, 3)).astype(np.float32)
        f_rest = rng.normal(0, 0.1, (P, 15, 3)).astype(np.float32)
        if gray:
            f_dc[:] = f_dc[:, :, :1]
    return dict(xyz=xyz, features_dc=f_dc, features_rest=f_rest, scaling=scaling, rotation=rotation,
                opacity=opacity, active_sh_degree=sh_degree)


def activate(g):
    """scene/gaussian_model.py:95-115: exp / normalize / sigmoid / cat."""
    scales = np.exp(g["scaling"]).astype(np.float32)
    n = np.sqrt((g["rotation"].astype(np.float32) ** 2).sum(1, keepdims=True))
    rot = (g["rotation"] / np.maximum(n, 1e-12)).astype(np.float32)
    opac = (1.0 / (1.0 + np.exp(-g["opacity"].astype(np.float64)))).astype(np.float32)
    feats = np.concatenate([g["features_dc"], g["features_rest"]], 1).astype(np.float32)
    return dict(means3D=g["xyz"], scales=scales, rotations=rot, opacities=opac, shs=feats)


def smooth_field(rng, H, W, channels, cells=6):
    """Smooth random field in [0,1], bilinear upsampling of a coarse grid; uint8-quantised like
    the reference's PIL loader (utils/general_utils.py:21-27)."""
    coarse = rng.uniform(0.05, 0.95, (channels, cells + 1, cells + 1))
    ys = np.linspace(0, cells, H, endpoint=False)
    xs = np.linspace(0, cells, W, endpoint=False)
    y0, x0 = ys.astype(int), xs.astype(int)
    fy, fx = (ys - y0)[None, :, None], (xs - x0)[None, None, :]
    a = coarse[:, y0][:, :, x0]
    b = coarse[:, y0][:, :, x0 + 1]
    c = coarse[:, y0 + 1][:, :, x0]
    d = coarse[:, y0 + 1][:, :, x0 + 1]
    img = (a * (1 - fy) * (1 - fx) + b * (1 - fy) * fx + c * fy * (1 - fx) + d * fy * fx)
    return (np.round(img * 255.0) / 255.0).astype(np.float32)


def make_gt_images(W, H, seed=0, gray=True):
    """GT set of one event view-pair: intensity frame, event frames (index, index+1; they differ
    on ~30 % of the pixels so the loss mask is mixed) and a 3-frame box-blurred frame."""
    rng = np.random.default_rng(seed + 1000)
    ch = 1 if gray else 3
    base = smooth_field(rng, H, W, ch)
    nxt = smooth_field(rng, H, W, ch)
    sel = (smooth_field(rng, H, W, 1, cells=9) > 0.62)
    ev_now = base
    ev_next = np.where(sel, np.clip(0.5 * base + 0.5 * nxt, 0, 1), base)
    ev_next = (np.round(ev_next * 255.0) / 255.0).astype(np.float32)
    blur = (np.round(((base + ev_next + nxt) / 3.0) * 255.0) / 255.0).astype(np.float32)
    rep = (lambda a: np.ascontiguousarray(np.repeat(a, 3, 0))) if gray else (lambda a: np.ascontiguousarray(a))
    return dict(intensity=rep(base), event_now=rep(ev_now), event_next=rep(ev_next), blurry=rep(blur))


def make_scene(name_or_cfg, seed=0, kind="trained", n_views=96, P=None):
    cfg = dict(CONFIGS[name_or_cfg]) if isinstance(name_or_cfg, str) else dict(name_or_cfg)
    if P is not None:
        cfg["P"] = P
    gray = not cfg.get("deblur", False)
    g = make_gaussians(cfg["P"], cfg["W"], seed, kind, cfg["sh_degree"], gray=gray)
    cams = [make_camera(i, n_views, cfg["W"], cfg["H"]) for i in range(n_views)]
    gt = make_gt_images(cfg["W"], cfg["H"], seed, gray=gray)
    return dict(cfg=cfg, gaussians=g, cameras=cams, gt=gt)
