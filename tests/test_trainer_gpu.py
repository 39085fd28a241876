"""GPU tests of the training loop (event_3dgs_b200/trainer.py): the same iterations done with the
reference-shaped pieces (drop-in render + torch autograd + torch.optim.Adam) must give the same
parameters, and a run with densification must keep going on the re-sized model."""
import random

import numpy as np
import pytest
import torch

from tests.common import rel_err

pytestmark = pytest.mark.gpu


def scene_on(dev, P=20_000, name="c1"):
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    scene = syn.make_scene(name, P=P)
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in scene["gt"].items()}
    return scene, pc, cams, gt


def test_trainer_steps_match_reference_shaped_loop(cuda_dev):
    """5 iterations: Trainer (fused iteration, FusedAdam, no-op skip of f_rest) against pipeline.event_iteration
    (drop-in rasterizer + autograd) with torch.optim.Adam on the reference's six groups + Adam(c, lr=0.1)."""
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import trainer as tr
    torch.backends.cuda.matmul.allow_tf32 = False
    scene, pc, cams, gt = scene_on(cuda_dev)
    _, pc_ref, _, _ = scene_on(cuda_dev)
    t = tr.Trainer(pc, cams, cams, lambda i: gt, event=True, extent=2.0, seed=3)
    opt = t.opt
    groups = tr.training_setup(pc_ref, opt, 2.0).param_groups
    ref_opt = torch.optim.Adam([{"params": g["params"], "lr": g["lr"], "name": g["name"]} for g in groups], lr=0.0, eps=1e-15)
    c_ref = torch.nn.Parameter(torch.tensor(0.17, device=cuda_dev))
    ref_opt_c = torch.optim.Adam([c_ref], lr=0.1)
    rng = random.Random(3)
    bg = torch.zeros(3, device=cuda_dev)
    for it in range(1, 6):
        loss = t.step()
        ref_opt.param_groups[0]["lr"] = tr.expon_lr(it, opt.position_lr_init * 2.0, opt.position_lr_final * 2.0, 0,
                                                     opt.position_lr_delay_mult, opt.position_lr_max_steps)
        index = tr.pick_index(rng, len(cams), True)
        assert t.last["indices"] == [index]
        l_ref, pkg = pl.event_iteration(pc_ref, cams[index], cams[index], cams[index + 1], gt, c_ref, bg)
        l_ref.backward()
        ref_opt.step()
        ref_opt_c.step()
        ref_opt.zero_grad(set_to_none=True)
        ref_opt_c.zero_grad(set_to_none=True)
        assert abs(float(loss) - float(l_ref.detach())) <= 2e-5 * abs(float(l_ref.detach())), it
    assert t.rest_skipped == 5 and t.optimizer.state[pc._features_rest]["step"] == 5
    # Adam normalises every gradient to +-lr in the first steps, so sign flips of tiny gradients show up:
    # compare against the size of the accumulated update
    for a, b, g in zip(pc.leaves(), pc_ref.leaves(), groups):
        d = (a.detach() - b.detach()).abs()
        frac_off = float((d > 0.05 * 5 * max(g["lr"], 1e-12)).float().mean())
        assert frac_off < 2e-3, (g["name"], frac_off)
    assert abs(float(t.c.detach()) - float(c_ref.detach())) < 1e-3
    assert torch.equal(pc._features_rest.detach(), pc_ref._features_rest.detach())


def test_training_run_with_densification_and_opacity_reset(cuda_dev):
    from event_3dgs_b200 import trainer as tr
    scene, pc, cams, gt = scene_on(cuda_dev, P=8_000)
    opt = tr.optimization_defaults(densify_from_iter=20, densification_interval=25, densify_until_iter=120, iterations=160)
    t = tr.Trainer(pc, cams, cams, lambda i: gt, opt=opt, event=True, extent=2.6, seed=1)
    t.opt.opacity_reset_interval = 60
    P0 = pc._xyz.shape[0]
    losses, sizes = [], [P0]
    for it in range(1, 161):
        losses.append(float(t.step()))
        if "densify" in t.last:
            d = t.last["densify"]
            assert d["rows"] == pc._xyz.shape[0] == d["kept"] + d["cloned"] + 2 * d["children_per_draw"]
            sizes.append(d["rows"])
        if it == 60:
            assert float(torch.sigmoid(pc._opacity).max()) <= 0.0100001        # reset_opacity
    assert len(sizes) == 1 + 4 and len(set(sizes)) > 1                          # iterations 25, 50, 75, 100
    assert all(np.isfinite(losses))
    assert np.mean(losses[-30:]) < np.mean(losses[:30])
    assert t.state.max_radii2D.shape[0] == pc._xyz.shape[0]
    for g in t.optimizer.param_groups:
        p = g["params"][0]
        assert p.shape[0] == pc._xyz.shape[0]
        assert t.optimizer.state[p]["exp_avg"].shape == p.shape
    # the frame (non-event) modes run through the same loop
    scene, pc2, cams, gt = scene_on(cuda_dev, P=8_000)
    t2 = tr.Trainer(pc2, cams, cams, lambda i: gt["intensity"], opt=tr.optimization_defaults(densify_from_iter=5, densification_interval=10),
                    event=False, gray=True, extent=2.6)
    l0 = [float(t2.step()) for _ in range(25)]
    assert all(np.isfinite(l0)) and pc2._xyz.shape[0] != 8_000


def test_trainer_dedupe_views_gives_the_same_losses(cuda_dev):
    """Shared first view (train camera == first event camera rendered once, the Trainer's default): same loss
    values, same parameters up to fp32 summation order after a few steps as with three renders."""
    from event_3dgs_b200 import trainer as tr
    _, pa, cams, gt = scene_on(cuda_dev, P=8_000)
    _, pb, _, _ = scene_on(cuda_dev, P=8_000)
    opt = lambda: tr.optimization_defaults(densify_from_iter=10 ** 9)
    ta = tr.Trainer(pa, cams, cams, lambda i: gt, opt=opt(), event=True, extent=2.6, seed=2, dedupe_views=False)
    tb = tr.Trainer(pb, cams, cams, lambda i: gt, opt=opt(), event=True, extent=2.6, seed=2)
    for it in range(6):
        la, lb = float(ta.step()), float(tb.step())
        assert abs(la - lb) <= 1e-4 * abs(la), it
    assert float((pa._features_dc.detach() - pb._features_dc.detach()).abs().max()) < 2e-3
    assert tb.dedupe_views and not ta.dedupe_views
