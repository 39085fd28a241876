"""CPU suite: the host logic of event_3dgs_b200/trainer.py against values produced by the reference's own
code (tests/golden/schedule.json, written by tests/golden/make_golden.py from utils/general_utils.py's
get_expon_lr_func and the view-sampling expressions of train.py:123-139 under random.seed(0))."""
import json
import os
import random

import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_learning_rate_schedule_and_view_sampling_match_reference():
    from event_3dgs_b200 import trainer as tr
    d = json.load(open(os.path.join(G, "schedule.json")))
    for s, want in zip(d["steps"], d["lr_xyz"]):
        got = tr.expon_lr(s, 0.00016 * 2.6, 0.0000016 * 2.6, 0, 0.01, 30000)
        assert got == pytest.approx(want, rel=1e-12), s
    for s, want in zip(d["steps"], d["lr_delay"]):
        got = tr.expon_lr(s, 0.01, 0.0001, 500, 0.05, 2000)
        assert got == pytest.approx(want, rel=1e-12), s
    assert tr.expon_lr(-1, 1.0, 0.1) == 0.0 and tr.expon_lr(5, 0.0, 0.0) == 0.0
    # utils/general_utils.py:130 seeds Python's RNG with 0; Trainer(seed=0) draws the same indices
    rng = random.Random(0)
    assert [tr.pick_index(rng, 96, True) for _ in range(40)] == d["views_event"]
    rng = random.Random(0)
    assert [tr.pick_index(rng, 96, False) for _ in range(40)] == d["views_frame"]
    assert not set(d["views_event"]) & set(tr.HELD_OUT)


def test_optimization_defaults_are_the_reference_values():
    from event_3dgs_b200 import trainer as tr
    o = tr.optimization_defaults()
    # arguments/__init__.py:76-93
    assert (o.iterations, o.position_lr_init, o.position_lr_final, o.position_lr_delay_mult, o.position_lr_max_steps) == (
        30_000, 0.00016, 0.0000016, 0.01, 30_000)
    assert (o.feature_lr, o.opacity_lr, o.scaling_lr, o.rotation_lr, o.percent_dense, o.lambda_dssim) == (0.0025, 0.05, 0.005, 0.001, 0.01, 0.2)
    assert (o.densification_interval, o.opacity_reset_interval, o.densify_from_iter, o.densify_until_iter, o.densify_grad_threshold) == (
        100, 3000, 500, 15_000, 0.0002)
    assert tr.optimization_defaults(iterations=7).iterations == 7
