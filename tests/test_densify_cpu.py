"""CPU suite: oracle/densify_oracle.py against the golden vectors that tests/golden/make_golden_densify.py
produced by running the reference's own GaussianModel.densify_and_prune / add_densification_stats /
reset_opacity (scene/gaussian_model.py) on CPU tensors."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import densify_oracle as do

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(G, "densify_*.npz")))


def oracle_from_golden(d, device="cpu"):
    t = lambda k: torch.tensor(d[k], device=device)
    m = do.Model({k: t("in_" + k) for k in do.NAMES}, {k: t("in_m_" + k) for k in do.NAMES}, {k: t("in_v_" + k) for k in do.NAMES})
    return m


@pytest.mark.parametrize("path", CASES)
def test_oracle_reproduces_reference_densification(path):
    d = np.load(path)
    assert len(CASES) >= 2
    m = oracle_from_golden(d)
    for radii, grad in zip(d["radii"], d["viewspace_grad"]):
        m.add_densification_stats(torch.tensor(grad), torch.tensor(radii))
    # the statistics are sums of identical float operations: bit-exact
    assert np.array_equal(m.xyz_gradient_accum.numpy(), d["stats_accum"])
    assert np.array_equal(m.max_radii2D.numpy(), d["stats_max_radii2D"])
    if int(d["poke_inf"]):       # the generator zeroed some visit counts afterwards: accum / denom = inf rows
        assert (d["stats_denom"] == 0).sum() >= int(d["poke_inf"]) and ((d["stats_accum"] > 0) & (d["stats_denom"] == 0)).any()
        m.denom = torch.tensor(d["stats_denom"])
    assert np.array_equal(m.denom.numpy(), d["stats_denom"])
    mss = int(d["max_screen_size"]) or None
    n_clone, n_sel = m.densify_and_prune(0.0002, 0.005, float(d["extent"]), mss, torch.tensor(d["z"]))
    assert 2 * n_sel == d["z"].shape[0] and n_clone > 0 and n_sel > 0
    for k in do.NAMES:
        assert m.p[k].shape == d["out_" + k].shape, k
        assert np.array_equal(m.p[k].numpy(), d["out_" + k]), k          # same torch ops on the same CPU
        assert np.array_equal(m.m[k].numpy(), d["out_m_" + k]), k
        assert np.array_equal(m.v[k].numpy(), d["out_v_" + k]), k
    assert not m.xyz_gradient_accum.any() and not m.denom.any() and not m.max_radii2D.any()
    assert m.max_radii2D.shape[0] == d["out_xyz"].shape[0]
    m.reset_opacity()
    assert np.array_equal(m.p["opacity"].numpy(), d["reset_opacity"])
    assert not m.m["opacity"].any() and not d["reset_m"].any()


def test_golden_cases_cover_every_branch():
    """Clone, split, opacity prune, world-size prune of old rows and of split children, never-seen rows."""
    seen = dict(world_prune=False, unseen=False)
    for path in CASES:
        d = np.load(path)
        seen["unseen"] |= bool((d["stats_denom"] == 0).any())
        if int(d["max_screen_size"]):
            big = np.exp(d["in_scaling"]).max(1) > 0.1 * float(d["extent"])
            seen["world_prune"] |= bool(big.any())
        assert d["out_xyz"].shape[0] != d["in_xyz"].shape[0]
    assert all(seen.values()), seen
