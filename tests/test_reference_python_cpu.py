"""CPU tests (build container only: skipped where /root/reference is absent, e.g. on the GPU box) of SURVEY.md
section 8f rank 4: the reference's OWN Python on top of this repository's drop-in packages.

  * the reference's gaussian_renderer / scene.gaussian_model import unmodified with event_3dgs_b200/dropin on the
    path, and their rasterizer / kNN / ply names resolve to this repository;
  * the synthetic COLMAP-layout dataset (event_3dgs_b200/colmap_dataset.py) is read back by the reference's own
    loader with the cameras, frames and points that were written;
  * what the reference's render() hands to the C boundary equals, bit for bit, what this repository's mirror
    pipeline.render() hands to it (the mirror is what the -m gpu tests and bench.py drive);
  * `python train.py -s <synthetic> --event --gray --deblur --iterations N` exits 0 and learns, with the CPU oracle
    answering below the C boundary (tests/ref_on_cpu.py explains the split).
Each check runs in a subprocess: the reference's top-level package names (utils, scene, arguments) and the CPU
redirections must not leak into the pytest process."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import ref_on_cpu  # noqa: E402

pytestmark = pytest.mark.skipif(not ref_on_cpu.available(), reason="the reference tree is not mounted here")


def _run(code, *argv, timeout=600, with_stderr=False):
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="4")
    p = subprocess.run([sys.executable, "-c", code, *argv], cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-3000:]
    return p.stdout + ("\n" + p.stderr if with_stderr else "")


def _write_dataset(root, n_views=96, W=64, H=48, P=4000, seed=0):
    """Frames rendered from a teacher model by the CPU oracle, so that training has something to learn."""
    from event_3dgs_b200 import colmap_dataset as cd
    from event_3dgs_b200 import synthetic as syn
    from oracle import raster_oracle as ro
    # dense enough that no pixel stays black: the event loss takes log(frame + 1e-8)
    g = syn.make_gaussians(P, W, seed=seed, kind="trained", mean_radius_px=5.0)
    g["opacity"] += 2.0
    a = syn.activate(g)
    cams = [syn.make_camera(i, n_views, W, H) for i in range(n_views)]
    colors = np.clip(a["shs"][:, 0, :] * syn.SH_C0 + 0.5, 0, 1).astype(np.float32)
    frames = []
    for c in cams:
        st = ro.rasterize_forward(a["means3D"], a["opacities"], c["world_view_transform"], c["full_proj_transform"], c["camera_center"],
                                  np.zeros(3, np.float32), W, H, c["tanfovx"], c["tanfovy"], scales=a["scales"], rotations=a["rotations"],
                                  colors_precomp=colors)
        frames.append(np.clip(st["out_color"], 0, 1))

    def frames_for(i):
        blur = (frames[max(i - 1, 0)] + frames[i] + frames[min(i + 1, n_views - 1)]) / 3.0
        return {"images": frames[i], "images_event": frames[i], "images_blurry": blur, "renders": frames[i]}
    names = cd.write_scene(root, cams, frames_for, a["means3D"][::3], colors[::3])
    return cams, frames, names, a["means3D"][::3], colors[::3]


def test_reference_modules_import_on_the_dropin_and_resolve_to_this_repository():
    out = _run("""
import json, sys
from tests import ref_on_cpu
ref_on_cpu.install(core="record")
import gaussian_renderer, scene.gaussian_model as gm, scene.dataset_readers as dr, utils.image_utils, utils.loss_utils
import event_3dgs_b200.rasterizer as rz, event_3dgs_b200.knn as knn, diff_gaussian_rasterization as dgr, plyfile
print(json.dumps(dict(
    rasterizer=gaussian_renderer.GaussianRasterizer is rz.GaussianRasterizer,
    settings=gaussian_renderer.GaussianRasterizationSettings is rz.GaussianRasterizationSettings,
    fields=list(rz.GaussianRasterizationSettings._fields),
    dgr_file=dgr.__file__, ply_file=plyfile.__file__, knn=gm.distCUDA2 is knn.distCUDA2,
    renderer_file=gaussian_renderer.__file__, gm_file=gm.__file__, ply_reader=dr.PlyData is plyfile.PlyData)))
""")
    d = json.loads(out.strip().splitlines()[-1])
    assert d["rasterizer"] and d["settings"] and d["knn"] and d["ply_reader"]
    assert d["fields"] == ["image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
                           "sh_degree", "campos", "prefiltered", "debug"]          # DGR/diff_gaussian_rasterization/__init__.py:157-169
    assert d["dgr_file"].startswith(os.path.join(ROOT, "event_3dgs_b200", "dropin"))
    assert d["ply_file"].startswith(os.path.join(ROOT, "event_3dgs_b200", "dropin"))
    assert d["renderer_file"].startswith(ref_on_cpu.REF) and d["gm_file"].startswith(ref_on_cpu.REF)      # the reference's own files


def test_reference_loader_reads_the_synthetic_colmap_scene(tmp_path):
    root = str(tmp_path / "scene")
    cams, frames, names, pts, cols = _write_dataset(root, n_views=12)
    out = _run("""
import json, sys, numpy as np
from tests import ref_on_cpu
ref_on_cpu.install(core="record")
from scene.dataset_readers import sceneLoadTypeCallbacks
info = sceneLoadTypeCallbacks["Colmap"](sys.argv[1], "images", False, True, False, True, True)
c = info.train_cameras
print(json.dumps(dict(n=[len(info.train_cameras), len(info.event_cameras), len(info.blurry_cameras), len(info.test_cameras)],
    names=[x.image_name for x in c], R=[x.R.tolist() for x in c], T=[x.T.tolist() for x in c], fovx=[x.FovX for x in c], fovy=[x.FovY for x in c],
    size=[c[0].width, c[0].height], img=np.asarray(c[3].image).tolist(), blur=np.asarray(info.blurry_cameras[3].image).tolist(),
    pts=np.asarray(info.point_cloud.points).tolist(), cols=np.asarray(info.point_cloud.colors).tolist(), radius=float(info.nerf_normalization["radius"]))))
""", root)
    d = json.loads(out.strip().splitlines()[-1])
    assert d["n"] == [12, 12, 12, 12] and d["names"] == [n[:-4] for n in names] and d["size"] == [64, 48]
    for i, c in enumerate(cams):
        assert np.allclose(d["R"][i], c["R"], atol=1e-6) and np.allclose(d["T"][i], c["T"], atol=1e-6)      # scene/cameras.py convention
        assert abs(d["fovx"][i] - c["FoVx"]) < 1e-6 and abs(d["fovy"][i] - c["FoVy"]) < 1e-6
    want = np.clip(np.round(np.transpose(frames[3], (1, 2, 0)) * 255.0), 0, 255)
    assert np.array_equal(np.asarray(d["img"]), want)
    assert np.abs(np.asarray(d["blur"]) - want).max() > 0                        # the blurry set really is another set of frames
    assert np.allclose(d["pts"], pts, atol=1e-6)
    assert np.allclose(d["cols"], np.round(cols * 255.0) / 255.0, atol=1e-6)    # (--gray alone would force them to 0.5, dataset_readers.py:195-196)
    assert d["radius"] > 0


def test_reference_render_and_mirror_hand_the_same_arguments_to_the_c_boundary():
    """The -m gpu tests and bench.py drive pipeline.render(), this repository's mirror of gaussian_renderer.render():
    both must put the same tensors in front of e3_rasterize_forward."""
    out = _run("""
import json, math, sys, numpy as np, torch
from tests import ref_on_cpu
ref_on_cpu.install(core="record")
import gaussian_renderer
from scene.gaussian_model import GaussianModel
from scene.cameras import Camera
from event_3dgs_b200 import pipeline as pl, synthetic as syn
g = syn.make_gaussians(500, 96, seed=3, kind="trained", sh_degree=2, gray=False)
cam = syn.make_camera(5, 96, 96, 64)
gm = GaussianModel(3)
P = torch.nn.Parameter
gm._xyz, gm._features_dc, gm._features_rest = P(torch.tensor(g["xyz"])), P(torch.tensor(g["features_dc"])), P(torch.tensor(g["features_rest"]))
gm._scaling, gm._rotation, gm._opacity = P(torch.tensor(g["scaling"])), P(torch.tensor(g["rotation"])), P(torch.tensor(g["opacity"]))
gm.active_sh_degree = 2
rc = Camera(colmap_id=1, R=cam["R"], T=cam["T"], FoVx=cam["FoVx"], FoVy=cam["FoVy"], image=torch.zeros(3, 64, 96), gt_alpha_mask=None,
            image_name="x", uid=0, data_device="cpu")
pipe = type("Pipe", (), dict(convert_SHs_python=False, compute_cov3D_python=False, debug=False))()
bg = torch.tensor([0.1, 0.2, 0.3])
pkg = gaussian_renderer.render(rc, gm, pipe, bg)
ref_call = ref_on_cpu.CALLS[-1][1]
keys_ref = sorted(pkg.keys())
pc = pl.GaussianParams(g, torch.device("cpu"))
pc.active_sh_degree = 2
mine = pl.render(rc, pc, pl.PIPE, bg)          # the same camera object: the matrices themselves are pinned by tests/golden/cameras
my_call = ref_on_cpu.CALLS[-1][1]
same = {}
for k in ref_call:
    a, b = ref_call[k], my_call[k]
    if isinstance(a, torch.Tensor) or isinstance(b, torch.Tensor):
        a = torch.empty(0) if a is None else a
        b = torch.empty(0) if b is None else b
        same[k] = bool(a.shape == b.shape and torch.equal(a.detach(), b.detach()))
    else:
        same[k] = (a is None and b is None) or a == b
print(json.dumps(dict(same=same, keys_ref=keys_ref, keys_mine=sorted(mine.keys()), n_calls=len(ref_on_cpu.CALLS),
                      retains_grad=bool(pkg["viewspace_points"].requires_grad))))
""")
    d = json.loads(out.strip().splitlines()[-1])
    assert all(d["same"].values()), d["same"]
    assert d["keys_ref"] == d["keys_mine"] == ["radii", "render", "viewspace_points", "visibility_filter"]      # gaussian_renderer/__init__.py:101-104
    assert d["n_calls"] == 2 and d["retains_grad"]


def test_reference_train_py_runs_unchanged_and_learns(tmp_path):
    """train.py --event --gray --deblur from the reference tree, unmodified, on the synthetic scene: Scene loading
    through the dataset writer and the plyfile stand-in, create_from_pcd over distCUDA2, three renders per iteration
    through GaussianRasterizer / _RasterizeGaussians, event loss, backward, densification statistics, a densification
    and both optimizers, checkpoint and point-cloud saving. Below the C boundary: the CPU oracle (no GPU here)."""
    root, out_dir = str(tmp_path / "scene"), str(tmp_path / "out")
    n_start = _write_dataset(root, n_views=96)[3].shape[0]
    code = "import sys\nfrom tests import ref_on_cpu\nref_on_cpu.run_train(sys.argv[1:])\n"
    log = _run(code, "-s", root, "-m", out_dir, "--event", "--gray", "--deblur", "--iterations", "130", "--densify_from_iter", "50",
               "--densification_interval", "60", "--save_iterations", "130", "--checkpoint_iterations", "130", "--port", "6131",
               timeout=1500, with_stderr=True)
    assert "Training complete." in log
    import re
    losses = [float(x) for x in re.findall(r"Loss=([0-9.]+)", log)]           # the progress bar's running mean, every 10 iterations
    assert len(losses) >= 10 and losses[-1] < 0.8 * max(losses[:3]), sorted(set(losses), reverse=True)
    assert os.path.isfile(os.path.join(out_dir, "point_cloud", "iteration_130", "point_cloud.ply"))
    assert os.path.isfile(os.path.join(out_dir, "chkpnt130.pth"))
    assert os.path.isfile(os.path.join(out_dir, "cameras.json"))
    # the saved cloud is readable by the stand-in and holds a densified / pruned model
    sys.path.insert(0, os.path.join(ROOT, "event_3dgs_b200", "dropin"))
    try:
        import plyfile
        v = plyfile.PlyData.read(os.path.join(out_dir, "point_cloud", "iteration_130", "point_cloud.ply")).elements[0]
    finally:
        sys.path.pop(0)
    assert len(v) != n_start and np.isfinite(np.asarray(v["opacity"])).all()
    assert len([p.name for p in v.properties if p.name.startswith("f_rest_")]) == 45
