"""CPU suite: the python restatements against golden vectors produced by the reference's OWN
utils/loss_utils.py and utils/sh_utils.py (tests/golden/make_golden.py imports them in the build
container)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle as lo
from tests.common import rel_err

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(G, "event_loss_*.npz"))))
def test_loss_oracle_matches_reference_loss_and_autograd(path):
    d = np.load(path)
    blur = d["gt_blur"] if bool(d["deblur"]) else None
    loss, g = lo.event_loss(d["image"], d["image_now"], d["image_next"], d["gt_now"], d["gt_next"], d["gt_int"],
                            float(d["c"]), blur)
    assert abs(loss - float(d["loss"])) <= 1e-5 * max(1e-3, abs(float(d["loss"])))
    assert rel_err(g["image"], d["d_image"]) < 1e-4
    # the event seeds vanish when the GT has no event at all (case c): compare against a floor
    fl = 1e-12
    assert rel_err(g["image_now"], d["d_image_now"], fl) < 1e-4
    assert rel_err(g["image_next"], d["d_image_next"], fl) < 1e-4
    assert abs(g["c"] - float(d["d_c"])) <= 1e-4 * max(1e-6, abs(float(d["d_c"])))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(G, "event_loss_*.npz"))))
def test_pipeline_event_loss_matches_reference(path):
    from event_3dgs_b200 import pipeline as pl
    d = np.load(path)
    t = lambda k: torch.tensor(d[k])
    image, now, nxt = t("image").requires_grad_(), t("image_now").requires_grad_(), t("image_next").requires_grad_()
    c = torch.tensor(float(d["c"]), requires_grad=True)
    loss = pl.event_loss(image, now, nxt, t("gt_now"), t("gt_next"), t("gt_int"), c, t("gt_blur") if bool(d["deblur"]) else None)
    loss.backward()
    assert abs(loss.item() - float(d["loss"])) <= 1e-6 * max(1e-3, abs(float(d["loss"])))
    assert rel_err(image.grad.numpy(), d["d_image"]) < 1e-6
    assert rel_err(now.grad.numpy(), d["d_image_now"], 1e-12) < 1e-5
    assert abs(c.grad.item() - float(d["d_c"])) <= 1e-5 * max(1e-6, abs(float(d["d_c"])))


def test_pipeline_eval_sh_matches_reference():
    from event_3dgs_b200 import pipeline as pl
    d = np.load(os.path.join(G, "eval_sh.npz"))
    sh, dirs = torch.tensor(d["sh"]), torch.tensor(d["dirs"])
    for deg in range(4):
        assert rel_err(pl.eval_sh(deg, sh, dirs).numpy(), d["rgb%d" % deg]) < 1e-6
    assert rel_err(((np.linspace(0, 1, 7) - 0.5) / pl.C0), d["rgb2sh"]) < 1e-6


def test_oracle_sh_colour_matches_reference_eval_sh():
    """computeColorFromSH of the C oracle (forward.cu:20-71) against the reference's python eval_sh."""
    from oracle import raster_oracle as ro
    from tests.common import scene_inputs
    d = np.load(os.path.join(G, "eval_sh.npz"))
    P = d["sh"].shape[0]
    inp = scene_inputs(P, 64, 64, seed=1)
    campos = inp["campos"]
    # place the Gaussians along the golden directions in front of the camera
    means = (campos[None, :] + 3.0 * d["dirs"]).astype(np.float32)
    shs = np.ascontiguousarray(d["sh"].transpose(0, 2, 1))      # [P,16,3] layout of the rasterizer
    for deg in range(4):
        st = ro.rasterize_forward(means, inp["opacities"], inp["viewmatrix"], inp["projmatrix"], campos, inp["bg"], 64, 64,
                                  inp["tanfovx"], inp["tanfovy"], scales=inp["scales"], rotations=inp["rotations"],
                                  shs=shs, sh_degree=deg)
        vis = st["radii"] > 0
        assert vis.sum() > 3
        dirs = means - campos[None, :]
        dirs = dirs / np.linalg.norm(dirs, axis=1, keepdims=True)
        ref = np.maximum(d["rgb%d" % deg] + 0.5, 0.0)       # gaussian_renderer/__init__.py:81
        assert np.allclose(dirs, d["dirs"], atol=1e-5)
        assert rel_err(st["rgb"][vis], ref[vis]) < 1e-4


@pytest.mark.parametrize("name", ["l1_ssim_gray", "l1_ssim_rgb"])
def test_pipeline_l1_ssim_matches_reference(name):
    from event_3dgs_b200 import pipeline as pl
    d = np.load(os.path.join(G, name + ".npz"))
    image = torch.tensor(d["image"]).requires_grad_()
    loss = pl.l1_ssim_loss(image, torch.tensor(d["gt"]), float(d["lam"]), bool(d["gray"]))
    loss.backward()
    assert abs(loss.item() - float(d["loss"])) <= 1e-6 * abs(float(d["loss"]))
    assert rel_err(image.grad.numpy(), d["d_image"]) < 1e-5


@pytest.mark.parametrize("name,cs", [("a", dict(P=1500, W=96, H=64, seed=21, kind="trained", bg=(0.1, 0.3, 0.6))),
                                     ("b", dict(P=800, W=50, H=70, seed=22, kind="init", bg=(0.0, 0.0, 0.0)))])
def test_c_oracle_against_reference_cuda_golden(name, cs):
    """The plain-C oracle against outputs of the UNMODIFIED reference CUDA build captured on a B200
    (tests/golden/make_golden_refcuda.py): geometry, sort order and ranges bit for bit; image 1e-5,
    gradients 1e-4 (the reference accumulates them with unordered fp32 atomics)."""
    from oracle import raster_oracle as ro
    from tests.common import oracle_forward, scene_inputs, smooth_dL
    d = np.load(os.path.join(G, "ref_cuda_%s.npz" % name), allow_pickle=False)
    inp = scene_inputs(cs["P"], cs["W"], cs["H"], seed=cs["seed"], kind=cs["kind"], bg=cs["bg"])
    o = oracle_forward(inp)
    vis = d["radii"] > 0
    assert o["num_rendered"] == int(d["num_rendered"])
    assert (o["radii"] == d["radii"]).all()
    assert (o["tiles_touched"] == d["tiles_touched"].view(np.uint32)).all()
    bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
    P = cs["P"]
    assert (bits(o["depths"])[vis] == bits(d["depths"])[vis]).all()
    assert (bits(o["cov3D"])[vis] == bits(d["cov3D"].reshape(P, 6))[vis]).all()
    assert (bits(o["means2D"])[vis] == bits(d["means2D"].reshape(P, 2))[vis]).all()
    co = d["conic_opacity"].reshape(P, 4)
    assert (bits(o["conic_opacity"])[vis] == bits(co)[vis]).all()
    assert (o["point_list"] == d["point_list"].view(np.uint32)).all()
    assert (o["point_list_keys"] == d["point_list_keys"].view(np.uint64)).all()
    assert (o["ranges"].reshape(-1) == d["ranges"].view(np.uint32)).all()
    assert rel_err(o["out_color"], d["image"]) < 1e-5
    assert rel_err(o["final_T"], d["final_T"]) < 1e-5
    # libm expf vs CUDA expf differ in the last bit: a handful of threshold decisions may flip
    assert (o["n_contrib"] != d["n_contrib"].view(np.uint32)).mean() < 2e-3
    g = ro.rasterize_backward(o, smooth_dL(3, cs["H"], cs["W"], cs["seed"]))
    fl = 1e-3 * float(np.abs(d["dL_dscales"]).max()) * float(np.abs(inp["scales"]).max())
    for k in ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dscales", "dL_drotations", "dL_dcov3D"):
        assert rel_err(g[k].reshape(d[k].shape), d[k], fl if k == "dL_drotations" else 0.0) < 1e-4, k
    conic = d["dL_dconic"].reshape(P, 4)
    assert rel_err(g["dL_dconic"], conic) < 1e-4


def test_synthetic_cameras_follow_the_reference_camera_recipe():
    """event_3dgs_b200.synthetic.make_camera against scene/cameras.py:48-57 evaluated with the reference's own
    utils/graphics_utils.py (tests/golden/cameras.npz): transposed storage, full_proj = view . proj,
    camera_center = inverse(view)[3, :3]."""
    from event_3dgs_b200 import synthetic as syn
    d = np.load(os.path.join(G, "cameras.npz"))
    n = 0
    while "view%d_index" % n in d:
        i, W, H = (int(x) for x in d["view%d_index" % n])
        cam = syn.make_camera(i, 96, W, H)
        for k in ("world_view_transform", "full_proj_transform", "camera_center"):
            want = d["view%d_%s" % (n, k)]
            assert cam[k].shape == want.shape and cam[k].dtype == np.float32
            assert np.abs(cam[k] - want).max() <= 2e-6 * max(1.0, np.abs(want).max()), (n, k)
        n += 1
    assert n == 3


def test_knn_oracle_reproduces_reference_initial_scales():
    """oracle dist2 (brute-force 3-NN in C) against the log-scales the reference's create_from_pcd computed
    (tests/golden/init_pcd.npz, made by running scene/gaussian_model.py:124-147 with an exact float64 3-NN)."""
    from oracle import raster_oracle as ro
    d = np.load(os.path.join(G, "init_pcd.npz"))
    d2 = np.maximum(ro.dist2(d["points"].astype(np.float32)), 1e-7)
    got = np.log(np.sqrt(d2))
    assert np.abs(got[:, None] - d["scaling"]).max() < 2e-6
    assert np.allclose(d["rotation"], np.array([[1.0, 0, 0, 0]])) and np.allclose(d["opacity"], np.log(0.1 / 0.9), atol=1e-6)
