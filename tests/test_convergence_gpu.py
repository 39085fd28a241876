"""GPU: the training loop actually learns. Ground-truth frames are renders of a 'teacher' set of Gaussians
from the scene's cameras (so the event pairs and intensity frames are mutually consistent); a perturbed copy is
trained with event_3dgs_b200.trainer and must move towards the teacher's images, in event mode and in the
non-event (L1 + SSIM) mode."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _teacher_student(dev, P=6000, W=128, H=96, seed=11):
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    sc = syn.make_scene(dict(P=P, W=W, H=H, sh_degree=0, deblur=False), seed=seed, n_views=32)
    teacher = pl.GaussianParams(sc["gaussians"], dev)
    with torch.no_grad():      # independent colour channels, so that the luma and the RGB losses both have work to do
        teacher._features_dc.add_(torch.randn_like(teacher._features_dc) * 0.5)
    cams = [pl.make_camera(c, dev) for c in sc["cameras"]]
    bg = torch.zeros(3, device=dev)
    with torch.no_grad():
        frames = [pl.render(c, teacher, pl.PIPE, bg)["render"].clamp(0, 1).contiguous() for c in cams]
    g = {k: np.array(v, copy=True) for k, v in sc["gaussians"].items() if isinstance(v, np.ndarray)}
    g["active_sh_degree"] = 0
    student = pl.GaussianParams(g, dev)
    gen = torch.Generator(device=dev).manual_seed(seed)
    with torch.no_grad():
        student._features_dc.copy_(teacher._features_dc + torch.randn(teacher._features_dc.shape, device=dev, generator=gen) * 0.6)
        student._xyz.add_(torch.randn(student._xyz.shape, device=dev, generator=gen) * 0.01)
        student._opacity.add_(torch.randn(student._opacity.shape, device=dev, generator=gen) * 0.5)
    return teacher, student, cams, frames, bg


def _psnr(pc, cams, frames, bg, idx):
    from event_3dgs_b200 import pipeline as pl
    with torch.no_grad():
        e = [float(((pl.render(cams[i], pc, pl.PIPE, bg)["render"].clamp(0, 1) - frames[i]) ** 2).mean()) for i in idx]
    return -10.0 * np.log10(np.mean(e))


def test_event_training_recovers_the_teacher(cuda_dev):
    from event_3dgs_b200 import trainer as tr
    teacher, student, cams, frames, bg = _teacher_student(cuda_dev)
    gt_for = lambda i: {"intensity": frames[i], "event_now": frames[i], "event_next": frames[i + 1]}
    opt = tr.optimization_defaults(densify_from_iter=10 ** 9, iterations=10 ** 9)
    t = tr.Trainer(student, cams, cams, gt_for, opt=opt, event=True, extent=2.6, seed=0)
    held = [5, 25]                                   # never sampled (train.py:137-139)
    p0 = _psnr(student, cams, frames, bg, held)
    first = [float(t.step()) for _ in range(30)]
    for _ in range(340):
        t.step()
    last = [float(t.step()) for _ in range(30)]
    p1 = _psnr(student, cams, frames, bg, held)
    # measured over repeated runs (float atomics make the trajectory differ from run to run): the loss falls to
    # 0.36-0.55 of its start, the held-out PSNR rises by 2.0-2.2 dB and then plateaus -- the event term
    # constrains log-intensity differences only, the intensity term carries weight 0.1
    assert np.mean(last) < 0.7 * np.mean(first), (np.mean(first), np.mean(last))
    assert p1 > p0 + 1.2, (p0, p1)
    assert abs(float(t.c.detach()) - 0.17) < 0.1              # the GT event frames were made with the same threshold


def test_frame_training_recovers_the_teacher(cuda_dev):
    from event_3dgs_b200 import trainer as tr
    teacher, student, cams, frames, bg = _teacher_student(cuda_dev, seed=12)
    opt = tr.optimization_defaults(densify_from_iter=10 ** 9, iterations=10 ** 9)
    t = tr.Trainer(student, cams, cams, lambda i: frames[i], opt=opt, event=False, gray=False, extent=2.6, seed=0)
    p0 = _psnr(student, cams, frames, bg, [5, 25])
    first = [float(t.step()) for _ in range(30)]
    for _ in range(340):
        t.step()
    last = [float(t.step()) for _ in range(30)]
    p1 = _psnr(student, cams, frames, bg, [5, 25])
    # measured: loss 0.072 -> 0.015, held-out PSNR 22.7 -> 33.4 dB after 400 iterations
    assert np.mean(last) < 0.4 * np.mean(first), (np.mean(first), np.mean(last))
    assert p1 > p0 + 6.0, (p0, p1)
