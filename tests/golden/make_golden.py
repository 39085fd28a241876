"""Generates tests/golden/*.npz by IMPORTING the reference's own python code from /root/reference
(build container only; the GPU box never runs this). Pins oracle/loss_oracle.py and the host
mirror event_3dgs_b200/pipeline.py to the reference's utils/loss_utils.py and utils/sh_utils.py.

    python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
for name in ("torchgeometry",):                      # imported but unused by utils/loss_utils.py:17
    sys.modules.setdefault(name, types.ModuleType(name))
try:
    import torchvision  # noqa: F401
except Exception:
    sys.modules.setdefault("torchvision", types.ModuleType("torchvision"))
from utils import loss_utils as lu  # noqa: E402
from utils import sh_utils as su  # noqa: E402


def event_case(seed, H, W, deblur, zero_frac):
    g = torch.Generator().manual_seed(seed)
    r = lambda: torch.rand(3, H, W, generator=g, dtype=torch.float32) * 0.9 + 0.05
    image, image_now, image_next = r().requires_grad_(), r().requires_grad_(), r().requires_grad_()
    q = lambda t: torch.round(t * 255) / 255
    gt_now = q(r()[:1].repeat(3, 1, 1))
    gt_next = q(r()[:1].repeat(3, 1, 1))
    same = torch.rand(1, H, W, generator=g) < zero_frac       # pixels without an event
    gt_next = torch.where(same, gt_now, gt_next)
    gt_int, gt_blur = q(r()), q(r())
    c = torch.tensor(0.17 + 0.05 * seed, requires_grad=True)
    # ---- train.py:165-203, verbatim flow ----
    img_diff = lu.differentialable_event_simu(image_now, image_next, False, c)
    gt_image = lu.differentialable_event_simu(gt_now, gt_next, False, 0.17)
    Ll1 = lu.l1_loss(img_diff, gt_image)
    lambda_dssim = 0
    loss1 = (1.0 - lambda_dssim) * Ll1 + lambda_dssim * (1.0 - lu.ssim_gray(img_diff, gt_image))
    Ll1b = lu.l1_loss(image, gt_int)
    loss2 = (1.0 - lambda_dssim) * Ll1b
    Event_weight = 0.9
    mask = (gt_image != 0).float()
    loss = Event_weight * (loss1 * mask).sum() + (1 - Event_weight) * (loss2 * (1 - mask)).sum()
    loss /= (mask.sum() + (1 - mask).sum())
    if deblur:
        blur_alpha = 0.5
        loss = (1.0 - blur_alpha) * loss + blur_alpha * lu.l1_loss(image, gt_blur)
    loss.backward()
    return dict(image=image.detach().numpy(), image_now=image_now.detach().numpy(), image_next=image_next.detach().numpy(),
                gt_now=gt_now.numpy(), gt_next=gt_next.numpy(), gt_int=gt_int.numpy(), gt_blur=gt_blur.numpy(),
                c=np.float32(c.item()), deblur=np.bool_(deblur), loss=np.float32(loss.item()),
                d_image=image.grad.numpy(), d_image_now=image_now.grad.numpy(), d_image_next=image_next.grad.numpy(),
                d_c=np.float32(c.grad.item()), img_diff=img_diff.detach().numpy(), gt_diff=gt_image.numpy())


def ssim_case(seed, H, W, gray):
    """train.py:213-223 (--gray, no --event) and train.py:292-296 (RGB), lambda_dssim = 0.2."""
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.linspace(0, 1, H), torch.linspace(0, 1, W), indexing="ij")
    base = torch.stack([0.5 + 0.4 * torch.sin(6 * xx + k) * torch.cos(5 * yy - k) for k in range(3)])
    gt = torch.round((base + 0.05 * torch.randn(3, H, W, generator=g)).clamp(0, 1) * 255) / 255
    image = (base + 0.1 * torch.randn(3, H, W, generator=g)).clamp(0, 1).requires_grad_()
    lam = 0.2
    if gray:
        Ll1 = lu.l1_loss_gray(image, gt)
        ss = lu.ssim_gray(image, gt)
    else:
        Ll1 = lu.l1_loss(image, gt)
        ss = lu.ssim(image, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - ss)
    loss.backward()
    return dict(image=image.detach().numpy(), gt=gt.numpy(), gray=np.bool_(gray), lam=np.float32(lam),
                loss=np.float32(loss.item()), l1=np.float32(Ll1.item()), ssim=np.float32(ss.item()), d_image=image.grad.numpy())


def sh_case(seed, P):
    g = torch.Generator().manual_seed(seed)
    sh = torch.randn(P, 3, 16, generator=g)
    d = torch.randn(P, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    out = {"sh": sh.numpy(), "dirs": d.numpy()}
    for deg in range(4):
        out["rgb%d" % deg] = su.eval_sh(deg, sh, d).numpy()
    out["rgb2sh"] = su.RGB2SH(torch.linspace(0, 1, 7)).numpy()
    return out


def schedule_golden():
    """tests/golden/schedule.json: the learning-rate schedule (utils/general_utils.py:38-62, called by
    the reference itself) and the view-index sequence of train.py:123-139 under random.seed(0)."""
    import json
    import random
    from utils.general_utils import get_expon_lr_func
    f = get_expon_lr_func(lr_init=0.00016 * 2.6, lr_final=0.0000016 * 2.6, lr_delay_mult=0.01, max_steps=30000)
    g_ = get_expon_lr_func(lr_init=0.01, lr_final=0.0001, lr_delay_steps=500, lr_delay_mult=0.05, max_steps=2000)
    steps = [0, 1, 2, 10, 100, 999, 1000, 5000, 15000, 29999, 30000, 40000]

    def pick(n, event):
        index = random.randint(2, n - 4) if event else random.randint(2, n - 3)
        return index - 1 if index in (5, 25, 45, 65, 85) else index
    random.seed(0)
    ev = [pick(96, True) for _ in range(40)]
    random.seed(0)
    fr = [pick(96, False) for _ in range(40)]
    json.dump({"steps": steps, "lr_xyz": [float(f(s)) for s in steps], "lr_delay": [float(g_(s)) for s in steps],
               "views_event": ev, "views_frame": fr}, open(os.path.join(HERE, "schedule.json"), "w"), indent=1)


def cameras_golden():
    """tests/golden/cameras.npz: the camera tensors of scene/cameras.py:48-57, computed with the reference's own
    utils/graphics_utils.py (getWorld2View2, getProjectionMatrix) for the (R, T, FoV) of a few synthetic views."""
    from utils.graphics_utils import getProjectionMatrix, getWorld2View2
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from event_3dgs_b200 import synthetic as syn
    out = {}
    for n, (i, W, H) in enumerate(((0, 64, 64), (7, 346, 260), (50, 1280, 720))):
        cam = syn.make_camera(i, 96, W, H)
        R, T = cam["R"], cam["T"]
        wvt = torch.tensor(getWorld2View2(R, T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
        proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=cam["FoVx"], fovY=cam["FoVy"]).transpose(0, 1)
        full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        center = wvt.inverse()[3, :3]
        out["view%d_index" % n] = np.array([i, W, H])
        out["view%d_world_view_transform" % n] = wvt.numpy()
        out["view%d_full_proj_transform" % n] = full.numpy()
        out["view%d_camera_center" % n] = center.numpy()
    np.savez_compressed(os.path.join(HERE, "cameras.npz"), **out)


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "event_loss_a.npz"), **event_case(0, 24, 40, False, 0.7))
    np.savez_compressed(os.path.join(HERE, "event_loss_b.npz"), **event_case(1, 17, 33, True, 0.3))
    np.savez_compressed(os.path.join(HERE, "event_loss_c.npz"), **event_case(2, 16, 16, False, 1.0))   # no events at all
    np.savez_compressed(os.path.join(HERE, "eval_sh.npz"), **sh_case(3, 64))
    np.savez_compressed(os.path.join(HERE, "l1_ssim_gray.npz"), **ssim_case(4, 37, 50, True))
    np.savez_compressed(os.path.join(HERE, "l1_ssim_rgb.npz"), **ssim_case(5, 33, 20, False))
    schedule_golden()
    cameras_golden()
    print("golden vectors written to", HERE)

