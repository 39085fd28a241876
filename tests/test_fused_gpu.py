"""-m gpu: the fused path (raw-parameter preprocess + event-loss kernels) against the
reference-shaped drop-in path (PyTorch activations / eval_sh / loss + autograd over the same
rasterizer) and against the numpy loss oracle."""
import numpy as np
import pytest
import torch

from tests.common import rel_err

pytestmark = pytest.mark.gpu


def _setup(P, W, H, sh_degree, deblur, seed=0):
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    dev = torch.device("cuda:0")
    cfg = dict(P=P, W=W, H=H, sh_degree=sh_degree, deblur=deblur)
    sc = syn.make_scene(cfg, seed=seed)
    pc = pl.GaussianParams(sc["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in sc["cameras"][:8]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in sc["gt"].items()}
    bg = torch.zeros(3, device=dev)
    c = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    return pl, pc, cams, gt, bg, c


@pytest.mark.parametrize("sh_degree,deblur", [(0, False), (3, True), (1, False), (2, True)])
def test_fused_iteration_matches_dropin_autograd(cuda_dev, sh_degree, deblur):
    from event_3dgs_b200 import fused
    pl, pc, cams, gt, bg, c = _setup(30_000, 346, 260, sh_degree, deblur)
    leaves = pc.leaves() + [c]
    loss_a, pkg_a = pl.event_iteration(pc, cams[3], cams[3], cams[4], gt, c, bg, deblur=deblur)
    loss_a.backward()
    ga = [p.grad.clone() for p in leaves]
    vs_a = pkg_a["viewspace_points"].grad.clone()
    for p in leaves:
        p.grad = None
    loss_b, pkg_b = fused.event_iteration_backward(pc, cams[3], cams[3], cams[4], gt, c, bg, deblur=deblur, return_pkg=True)
    gb = [p.grad.clone() for p in leaves]
    assert abs(loss_a.item() - loss_b.item()) <= 1e-5 * abs(loss_a.item())
    # activations are evaluated by different code (ATen vs in-kernel): allow a few borderline radii
    assert (pkg_a["radii"] != pkg_b["radii"]).float().mean().item() < 1e-4
    names = ["xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity", "c"]
    for n, a, b in zip(names, ga, gb):
        assert a.shape == b.shape, n
        fl = 0.0
        if n == "features_rest" and sh_degree == 0:
            assert (a == 0).all() and (b == 0).all()
            continue
        assert rel_err(b.cpu().numpy(), a.cpu().numpy(), fl) < 2e-4, n
    assert rel_err(pkg_b["viewspace_points"].grad.cpu().numpy(), vs_a.cpu().numpy()) < 2e-4
    assert torch.equal(pkg_b["visibility_filter"], pkg_b["radii"] > 0)


def test_event_loss_kernel_vs_numpy_oracle_and_golden(cuda_dev):
    import glob
    import os
    from event_3dgs_b200 import fused
    from oracle import loss_oracle as lo
    G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    for path in sorted(glob.glob(os.path.join(G, "event_loss_*.npz"))):
        d = np.load(path)
        T = lambda k: torch.tensor(d[k]).to(cuda_dev).contiguous()
        deblur = bool(d["deblur"])
        gt = {"event_now": T("gt_now"), "event_next": T("gt_next"), "intensity": T("gt_int"), "blurry": T("gt_blur")}
        c = torch.tensor(float(d["c"]), device=cuda_dev)
        out, seeds = fused.event_loss_and_seeds(T("image"), T("image_now"), T("image_next"), gt, c, deblur)
        out = out.cpu().numpy()
        assert abs(out[0] - float(d["loss"])) <= 1e-5 * max(1e-3, abs(float(d["loss"])))     # reference's own value
        assert abs(out[1] - float(d["d_c"])) <= 1e-4 * max(1e-6, abs(float(d["d_c"])))
        assert rel_err(seeds["image"].cpu().numpy(), d["d_image"]) < 1e-5
        assert rel_err(seeds["image_now"].cpu().numpy(), d["d_image_now"], 1e-12) < 1e-4
        assert rel_err(seeds["image_next"].cpu().numpy(), d["d_image_next"], 1e-12) < 1e-4
        loss, g = lo.event_loss(d["image"], d["image_now"], d["image_next"], d["gt_now"], d["gt_next"], d["gt_int"],
                                float(d["c"]), d["gt_blur"] if deblur else None)
        assert abs(out[0] - loss) <= 1e-5 * max(1e-3, abs(loss))
    # full-size (1280x720) property: seeds scale linearly with `upstream`, loss does not
    rng = np.random.default_rng(0)
    r = lambda: torch.tensor(rng.uniform(0.05, 0.95, (3, 720, 1280)).astype(np.float32)).to(cuda_dev)
    gt = {"event_now": r(), "event_next": r(), "intensity": r(), "blurry": r()}
    imgs = (r(), r(), r())
    c = torch.tensor(0.2, device=cuda_dev)
    o1, s1 = fused.event_loss_and_seeds(*imgs, gt, c, True, upstream=1.0)
    o2, s2 = fused.event_loss_and_seeds(*imgs, gt, c, True, upstream=2.0)
    assert abs(o1[0].item() - o2[0].item()) < 1e-7
    assert rel_err((2 * s1["image_now"]).cpu().numpy(), s2["image_now"].cpu().numpy()) < 1e-6
    assert abs(2 * o1[1].item() - o2[1].item()) <= 1e-6 * abs(o2[1].item())
