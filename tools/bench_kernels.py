"""Timing of the streaming kernels beside the blend path (CUDA events, inputs > L2 or rotated):
Adam, event loss, L1+SSIM, densification statistics, distCUDA2 -- with the torch / ATen route the
reference takes for the same operation next to each. Prints one JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_3dgs_b200 import densify as dn, fused, knn, pipeline as pl  # noqa: E402
from event_3dgs_b200.optim import FusedAdam  # noqa: E402


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    dev = torch.device("cuda", 0)
    peak = 6552.3
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    out = {"hbm_peak_gbs": peak}
    P, W, H = 1_000_000, 1280, 720
    torch.manual_seed(0)
    # ---- Adam over the reference's six groups, 59 floats per Gaussian ----
    shapes = [(P, 3), (P, 1, 3), (P, 15, 3), (P, 1), (P, 3), (P, 4)]
    a = [torch.randn(s, device=dev).requires_grad_() for s in shapes]
    b = [x.detach().clone().requires_grad_() for x in a]
    for x, y in zip(a, b):
        x.grad = torch.randn_like(x)
        y.grad = x.grad.clone()
    mk = lambda ps: [{"params": [p], "lr": 1e-3} for p in ps]
    ours, ref = FusedAdam(mk(a), eps=1e-15), torch.optim.Adam(mk(b), eps=1e-15)
    t_o, t_r = timed(ours.step), timed(ref.step)
    by = 59 * P * 28
    out["adam_59_floats_per_gaussian"] = {"ours_ms": t_o, "torch_ms": t_r, "algorithmic_GB": by / 1e9, "ours_GBps": by / t_o / 1e6,
                                          "frac_of_hbm_peak": by / t_o / 1e6 / peak}
    del a, b, ours, ref
    # ---- event loss (2 kernels): 3 renders + 3 GT frames in, 3 seeds out ----
    r = lambda c: torch.rand(c, H, W, device=dev) * 0.9 + 0.05
    image, now, nxt = r(3), r(1), r(1)
    gt = {"intensity": r(3), "event_now": (r(1) * 255).round().div(255).expand(3, H, W).contiguous(),
          "event_next": (r(1) * 255).round().div(255).expand(3, H, W).contiguous()}
    gt8 = {k: (v * 255).round().to(torch.uint8) for k, v in gt.items()}
    c = torch.tensor(0.17, device=dev)
    t_f = timed(lambda: fused.event_loss_and_seeds(image, now, nxt, gt, c))
    t_8 = timed(lambda: fused.event_loss_and_seeds(image, now, nxt, gt8, c))
    img3n, img3x = now.expand(3, H, W).contiguous().requires_grad_(), nxt.expand(3, H, W).contiguous().requires_grad_()
    imgr = image.clone().requires_grad_()
    cp = c.clone().requires_grad_()

    def torch_loss():
        for t in (img3n, img3x, imgr, cp):
            t.grad = None
        pl.event_loss(imgr, img3n, img3x, gt["event_now"], gt["event_next"], gt["intensity"], cp, None).backward()
    t_t = timed(torch_loss, n=10)
    N = W * H
    by = (5 * 4 + 9 * 4) * N * 2 + 5 * 4 * N       # both passes read 5 render + 9 GT floats per pixel, pass 2 writes 5 seeds
    out["event_loss_1280x720"] = {"ours_ms_float_gt": t_f, "ours_ms_u8_gt": t_8, "torch_autograd_ms": t_t, "algorithmic_GB_float_gt": by / 1e9,
                                  "ours_GBps_float_gt": by / t_f / 1e6, "frac_of_hbm_peak": by / t_f / 1e6 / peak}
    # ---- L1 + SSIM (gray), fused stencil kernels vs 5 depthwise convs + autograd ----
    t_o = timed(lambda: fused.l1_ssim_loss_and_seed(image, gt["intensity"], 0.2, True))

    def torch_ssim():
        imgr.grad = None
        pl.l1_ssim_loss(imgr, gt["intensity"], 0.2, True).backward()
    torch.backends.cudnn.allow_tf32 = False
    t_t = timed(torch_ssim, n=10)
    out["l1_ssim_gray_1280x720"] = {"ours_ms": t_o, "torch_autograd_ms": t_t}
    # ---- densification statistics ----
    st = dn.DensificationState(P, dev)
    radii = torch.randint(-5, 30, (P,), device=dev, dtype=torch.int32).clamp_min(0)
    g2 = torch.randn(P, 3, device=dev)
    t_o = timed(lambda: dn.add_densification_stats(st, g2, radii))
    mr, acc, den = torch.zeros(P, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)

    def torch_stats():
        vis = radii > 0
        mr[vis] = torch.max(mr[vis], radii[vis])
        acc[vis] += torch.norm(g2[vis, :2], dim=-1, keepdim=True)
        den[vis] += 1
    t_t = timed(torch_stats, n=10)
    out["densification_stats_1M"] = {"ours_ms": t_o, "torch_ms": t_t}
    # ---- distCUDA2 ----
    for n in (100_000, 1_000_000):
        pts = torch.rand(n, 3, device=dev) * 2.6 - 1.3
        t_o = timed(lambda: knn.distCUDA2(pts), n=5, warm=1)
        out["distCUDA2_%dk_uniform" % (n // 1000)] = {"ours_ms": t_o}
    pts = torch.rand(20_000, 3, device=dev)

    def brute():
        d = torch.cdist(pts, pts)
        d.fill_diagonal_(float("inf"))
        return (d.topk(3, largest=False).values ** 2).mean(1)
    out["distCUDA2_20k_uniform"] = {"ours_ms": timed(lambda: knn.distCUDA2(pts), n=5, warm=1), "torch_bruteforce_ms": timed(brute, n=3, warm=1)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
