"""CPU suite: the plain-C oracle against (a) an independent PyTorch transcription of the same
algorithm (values and autograd gradients), (b) numpy restatements of the integer stages, and
(c) structural invariants of the reference's data structures (Appendix A of SURVEY.md)."""
import numpy as np
import pytest
import torch

from oracle import raster_oracle as ro
from oracle import torch_rasterizer as tr
from tests.common import oracle_forward, rel_err, scene_inputs, smooth_dL


def test_get_higher_msb_matches_reference_table():
    # 5/9/12/13 for c1/c2/c3/c5 tile counts
    assert [ro.get_higher_msb(n) for n in (16, 374, 3600, 8160)] == [5, 9, 12, 13]
    for n in (1, 2, 3, 255, 256, 257, 65535, 65536):
        assert (1 << ro.get_higher_msb(n)) > n - 1


def _small_scene(P=60, W=48, H=40, seed=3):
    rng = np.random.default_rng(seed)
    inp = scene_inputs(P, W, H, seed=seed, bg=(0.3, 0.1, 0.7))
    inp["means3D"] = rng.uniform(-0.9, 0.9, (P, 3)).astype(np.float32)
    inp["scales"] = rng.uniform(0.05, 0.35, (P, 3)).astype(np.float32)
    inp["opacities"] = rng.uniform(0.05, 0.95, (P, 1)).astype(np.float32)
    return inp


def test_oracle_matches_torch_transcription_forward_and_backward():
    inp = _small_scene()
    W, H = inp["W"], inp["H"]
    st = oracle_forward(inp)
    dL = smooth_dL(3, H, W)
    g = ro.rasterize_backward(st, dL)
    t = lambda x: torch.tensor(x, dtype=torch.float64, requires_grad=True)
    tm, to, ts, trot, tc = t(inp["means3D"]), t(inp["opacities"]), t(inp["scales"]), t(inp["rotations"]), t(inp["colors"])
    img, radii, aux = tr.rasterize(tm, to, ts, trot, tc, torch.tensor(inp["viewmatrix"]), torch.tensor(inp["projmatrix"]),
                                   torch.tensor(inp["bg"]), W, H, inp["tanfovx"], inp["tanfovy"])
    assert (radii.numpy() == st["radii"]).all()
    assert rel_err(st["out_color"], img.detach().numpy()) < 1e-5
    assert rel_err(st["final_T"], aux["final_T"].numpy()) < 1e-5
    (img * torch.tensor(dL, dtype=torch.float64)).sum().backward()
    for tt, k in [(tm, "dL_dmeans3D"), (to, "dL_dopacity"), (ts, "dL_dscales"), (trot, "dL_drotations"), (tc, "dL_dcolors")]:
        assert rel_err(g[k].reshape(tt.shape), tt.grad.numpy()) < 1e-4, k


def test_oracle_single_channel_equals_luma_of_rgb():
    # blending is linear in colour: rendering Y = w.c with bg Y(bg) equals the luma of the RGB render
    inp = _small_scene(seed=5)
    w = np.array([0.4124, 0.35758, 0.1804], np.float32)     # utils/loss_utils.py:25
    st3 = oracle_forward(inp)
    inp1 = dict(inp)
    inp1["colors"] = (inp["colors"] @ w)[:, None].astype(np.float32)
    inp1["bg"] = np.array([inp["bg"] @ w], np.float32)
    st1 = oracle_forward(inp1)
    luma = np.tensordot(w, st3["out_color"], 1)
    assert rel_err(st1["out_color"][0], luma) < 1e-5
    assert (st1["n_contrib"] == st3["n_contrib"]).all()


def test_binning_invariants_and_sort_against_numpy():
    inp = scene_inputs(3000, 96, 80, seed=1)
    st = oracle_forward(inp)
    R = st["num_rendered"]
    assert R == int(st["tiles_touched"].sum()) and R > 0
    # stable sort restated in numpy
    order = np.argsort(st["point_list_keys_unsorted"], kind="stable")
    assert (st["point_list_keys"] == st["point_list_keys_unsorted"][order]).all()
    assert (st["point_list"] == st["point_list_unsorted"][order]).all()
    # keys: high word tile id, low word the raw depth bits of the Gaussian
    tiles = (st["point_list_keys"] >> np.uint64(32)).astype(np.int64)
    depth_bits = st["depths"].view(np.uint32)[st["point_list"]]
    assert ((st["point_list_keys"] & np.uint64(0xffffffff)).astype(np.uint32) == depth_bits).all()
    # ranges partition the sorted list by tile; untouched tiles stay (0,0)
    T = st["gx"] * st["gy"]
    cnt = np.bincount(tiles, minlength=T)
    rng_len = (st["ranges"][:, 1].astype(np.int64) - st["ranges"][:, 0])
    assert (rng_len == cnt).all()
    for t in np.nonzero(cnt)[0][:50]:
        a, b = st["ranges"][t]
        assert (tiles[a:b] == t).all()
        d = st["depths"][st["point_list"][a:b]]
        assert (np.diff(d) >= 0).all()
    assert (st["ranges"][cnt == 0] == 0).all()
    # n_contrib never exceeds the tile's list length
    n = st["n_contrib"].reshape(st["H"], st["W"])
    for ty in range(st["gy"]):
        for tx in range(st["gx"]):
            blk = n[ty * 16:(ty + 1) * 16, tx * 16:(tx + 1) * 16]
            assert blk.max() <= rng_len[ty * st["gx"] + tx]


def test_culling_rules():
    inp = _small_scene(P=8)
    cam_dir = -inp["campos"] / np.linalg.norm(inp["campos"])
    m = inp["means3D"].copy()
    m[0] = inp["campos"] + 0.1 * cam_dir          # z_view = 0.1 <= 0.2: culled (auxiliary.h:154)
    m[1] = inp["campos"] - 1.0 * cam_dir          # behind the camera
    m[2] = inp["campos"] + 0.25 * cam_dir         # just in front of the near limit: kept
    inp["means3D"] = m
    st = oracle_forward(inp)
    assert st["radii"][0] == 0 and st["radii"][1] == 0 and st["radii"][2] > 0
    assert st["tiles_touched"][0] == 0 and st["tiles_touched"][1] == 0
    vis = ro.mark_visible(inp["means3D"], inp["viewmatrix"], inp["projmatrix"])
    assert not vis[0] and not vis[1] and vis[2]


def test_empty_and_degenerate_inputs():
    inp = _small_scene(P=0)
    st = oracle_forward(inp)
    assert st["num_rendered"] == 0 and (st["out_color"] == 0).all()     # rasterize_points.cu:68,81
    # every Gaussian culled: image is the background, R = 0
    inp = _small_scene(P=5)
    inp["means3D"] = (inp["campos"][None, :] * 1.5).repeat(5, 0).astype(np.float32)
    st = oracle_forward(inp)
    assert st["num_rendered"] == 0
    assert np.allclose(st["out_color"], inp["bg"][:, None, None])
    g = ro.rasterize_backward(st, smooth_dL(3, inp["H"], inp["W"]))
    assert all((v == 0).all() for v in g.values())


def test_background_and_early_stop_semantics():
    # a stack of opaque Gaussians in front of one pixel: T stops before 1e-4 and the stopping
    # instance is not blended (forward.cu:346-351)
    inp = _small_scene(P=12, seed=9)
    cam_dir = -inp["campos"] / np.linalg.norm(inp["campos"])
    inp["means3D"] = np.stack([inp["campos"] + (2.0 + 0.1 * i) * cam_dir for i in range(12)]).astype(np.float32)
    inp["scales"] = np.full((12, 3), 0.2, np.float32)
    inp["opacities"] = np.full((12, 1), 0.95, np.float32)
    st = oracle_forward(inp)
    cy, cx = inp["H"] // 2, inp["W"] // 2
    T = st["final_T"].reshape(inp["H"], inp["W"])[cy, cx]
    n = st["n_contrib"].reshape(inp["H"], inp["W"])[cy, cx]
    assert 1e-4 <= T < 1e-2 and 0 < n < 12
    assert T * 0.1 < 1e-4       # the next (alpha ~ 0.95) instance would have crossed the stop threshold


def test_dist2_bruteforce_contract():
    rng = np.random.default_rng(0)
    pts = rng.normal(0, 1, (500, 3)).astype(np.float32)
    pts[7] = pts[3]                                    # duplicates count with distance 0
    d = ro.dist2(pts)
    D = ((pts[:, None, :].astype(np.float64) - pts[None, :, :]) ** 2).sum(-1)
    np.fill_diagonal(D, np.inf)
    ref = np.sort(D, 1)[:, :3].mean(1)
    assert rel_err(d, ref) < 1e-5
    assert d[7] < ref.mean()
