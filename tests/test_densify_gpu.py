"""GPU parity of the adaptive density control (csrc/densify.cu through the C ABI) against
 (a) the golden vectors produced by the reference's own GaussianModel methods (CPU run), and
 (b) oracle/densify_oracle.py executed on the GPU at 200k rows (same ATen exp/log/sigmoid),
 plus size-independent properties at 1M rows (BASELINE c3)."""
import glob
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import densify_oracle as do

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", scaling="_scaling", rotation="_rotation")
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)


def make_model(params, m, v, dev, step=2):
    """A model object with the reference's attribute names + FusedAdam with the reference's six groups."""
    from event_3dgs_b200.optim import FusedAdam
    pc = SimpleNamespace(active_sh_degree=0)
    groups = []
    for k in do.NAMES:
        p = torch.nn.Parameter(torch.as_tensor(params[k]).to(dev).contiguous())
        setattr(pc, ATTR[k], p)
        groups.append({"params": [p], "lr": LRS[k], "name": k})
    opt = FusedAdam(groups, lr=0.0, eps=1e-15)
    if m is not None:
        for k in do.NAMES:
            p = getattr(pc, ATTR[k])
            opt.state[p] = {"step": step, "exp_avg": torch.as_tensor(m[k]).to(dev).contiguous(),
                            "exp_avg_sq": torch.as_tensor(v[k]).to(dev).contiguous()}
    return pc, opt


def close(a, b, tol=2e-6):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return bool(np.all(np.abs(a - b) <= tol * np.maximum(1.0, np.abs(b))))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(G, "densify_*.npz"))))
def test_densify_matches_reference_golden(cuda_dev, path):
    from event_3dgs_b200 import densify as dn
    d = np.load(path)
    pc, opt = make_model({k: d["in_" + k] for k in do.NAMES}, {k: d["in_m_" + k] for k in do.NAMES},
                         {k: d["in_v_" + k] for k in do.NAMES}, cuda_dev)
    P = d["in_xyz"].shape[0]
    st = dn.DensificationState(P, cuda_dev)
    for radii, grad in zip(d["radii"], d["viewspace_grad"]):
        dn.add_densification_stats(st, torch.tensor(grad).to(cuda_dev), torch.tensor(radii).to(cuda_dev))
    if int(d["poke_inf"]):       # visit counts zeroed by the generator after the fact (accum / denom = inf rows)
        st.denom.copy_(torch.tensor(d["stats_denom"]))
    assert np.array_equal(st.denom.cpu().numpy(), d["stats_denom"])
    assert np.array_equal(st.max_radii2D.cpu().numpy(), d["stats_max_radii2D"])
    assert close(st.xyz_gradient_accum.cpu().numpy(), d["stats_accum"], 1e-6)
    # the reference's statistics (bit-exact input to the criteria), then its densify_and_prune
    st.xyz_gradient_accum.copy_(torch.tensor(d["stats_accum"]))
    mss = int(d["max_screen_size"]) or None
    counts = dn.densify_and_prune(pc, opt, st, 0.0002, 0.005, float(d["extent"]), mss, normals=torch.tensor(d["z"]))
    assert counts["rows"] == d["out_xyz"].shape[0] and 2 * counts["split_selected"] == d["z"].shape[0]
    for k in do.NAMES:
        p = getattr(pc, ATTR[k])
        assert isinstance(p, torch.nn.Parameter) and p.requires_grad
        assert opt.param_groups[do.NAMES.index(k)]["params"][0] is p
        assert close(p.detach().cpu().numpy(), d["out_" + k]), k
        s = opt.state[p]
        assert np.array_equal(s["exp_avg"].cpu().numpy(), d["out_m_" + k]), k       # pure row moves: bit-exact
        assert np.array_equal(s["exp_avg_sq"].cpu().numpy(), d["out_v_" + k]), k
        assert s["step"] == 2
    assert len(opt.state) == 6
    # rows that are copies (everything but the children's xyz / scaling) are bit-exact
    n_copy = counts["kept"] + counts["cloned"]
    for k in do.NAMES:
        got = getattr(pc, ATTR[k]).detach().cpu().numpy()
        assert np.array_equal(got[:n_copy], d["out_" + k][:n_copy]), k
        if k not in ("xyz", "scaling"):
            assert np.array_equal(got, d["out_" + k]), k
    assert st.max_radii2D.shape[0] == counts["rows"] and not st._buf.any()
    dn.reset_opacity(pc, opt)
    assert close(pc._opacity.detach().cpu().numpy(), d["reset_opacity"], 1e-6)
    assert not opt.state[pc._opacity]["exp_avg"].any() and not opt.state[pc._opacity]["exp_avg_sq"].any()
    # the optimizer still steps on the new tensors
    for k in do.NAMES:
        p = getattr(pc, ATTR[k])
        p.grad = torch.ones_like(p)
    opt.step()
    assert opt.state[pc._xyz]["step"] == 3


def random_case(P, dev, seed, extent):
    g = torch.Generator(device="cpu").manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    params = dict(xyz=r(P, 3), f_dc=r(P, 1, 3) * 0.5, f_rest=r(P, 15, 3) * 0.1, opacity=r(P, 1) * 3 - 2,
                  scaling=float(np.log(0.01 * extent)) - 0.8 + r(P, 3) * 1.2, rotation=r(P, 4))
    m = {k: r(*params[k].shape) * 0.01 for k in do.NAMES}
    v = {k: (r(*params[k].shape) * 0.01) ** 2 for k in do.NAMES}
    accum = (r(P, 1).abs() * 0.0005)
    denom = torch.randint(0, 4, (P, 1), generator=g).float()
    accum[denom == 0] = 0
    return params, m, v, accum, denom


@pytest.mark.parametrize("mss", [None, 20])
def test_densify_matches_oracle_on_gpu_200k(cuda_dev, mss):
    from event_3dgs_b200 import densify as dn
    P, extent = 200_003, 3.0
    params, m, v, accum, denom = random_case(P, cuda_dev, 5, extent)
    pc, opt = make_model(params, m, v, cuda_dev)
    st = dn.DensificationState(P, cuda_dev)
    st.xyz_gradient_accum.copy_(accum)
    st.denom.copy_(denom)
    ora = do.Model({k: params[k].to(cuda_dev) for k in do.NAMES}, {k: m[k].to(cuda_dev) for k in do.NAMES},
                   {k: v[k].to(cuda_dev) for k in do.NAMES})
    ora.xyz_gradient_accum, ora.denom = accum.to(cuda_dev), denom.to(cuda_dev)
    zs = {}

    def draw(n):
        if n not in zs:
            zs[n] = torch.randn(n, 3, generator=torch.Generator().manual_seed(9)).to(cuda_dev)
        return zs[n]

    counts = dn.densify_and_prune(pc, opt, st, 0.0002, 0.005, extent, mss, normals=draw)
    n_clone, n_sel = ora.densify_and_prune(0.0002, 0.005, extent, mss, draw)
    assert counts["split_selected"] == n_sel and counts["rows"] == ora.p["xyz"].shape[0]
    assert counts["kept"] > 0 and counts["cloned"] > 0 and counts["children_per_draw"] > 0
    for k in do.NAMES:
        got = getattr(pc, ATTR[k]).detach()
        assert got.shape == ora.p[k].shape
        if k == "xyz":       # bmm vs the kernel's fma chain
            assert close(got.cpu().numpy(), ora.p[k].cpu().numpy(), 2e-6)
        else:
            assert torch.equal(got, ora.p[k]), k
        assert torch.equal(opt.state[getattr(pc, ATTR[k])]["exp_avg"], ora.m[k]), k
        assert torch.equal(opt.state[getattr(pc, ATTR[k])]["exp_avg_sq"], ora.v[k]), k


def test_densify_properties_at_c3_size_and_edge_cases(cuda_dev):
    """1M rows (BASELINE c3): counts add up, kept originals keep their moments and order, new rows have
    zero moments; nothing selected -> identity; everything pruned -> empty model; no optimizer state."""
    from event_3dgs_b200 import densify as dn
    P, extent = 1_000_000, 3.0
    params, m, v, accum, denom = random_case(P, cuda_dev, 11, extent)
    params["f_rest"] = torch.arange(P, dtype=torch.float32).view(P, 1, 1).expand(P, 15, 3).contiguous()   # row ids travel with the rows
    pc, opt = make_model(params, m, v, cuda_dev)
    st = dn.DensificationState(P, cuda_dev)
    st.xyz_gradient_accum.copy_(accum)
    st.denom.copy_(denom)
    c = dn.densify_and_prune(pc, opt, st, 0.0002, 0.005, extent, 20, generator=torch.Generator(device=cuda_dev).manual_seed(3))
    assert c["rows"] == c["kept"] + c["cloned"] + 2 * c["children_per_draw"] == pc._xyz.shape[0]
    ids = pc._features_rest.detach()[:, 0, 0]
    seg = [0, c["kept"], c["kept"] + c["cloned"], c["kept"] + c["cloned"] + c["children_per_draw"], c["rows"]]
    for a, b in zip(seg[:-1], seg[1:]):
        assert bool((ids[a + 1:b] > ids[a:b - 1]).all())              # source order inside each segment
    assert torch.equal(ids[seg[2]:seg[3]], ids[seg[3]:seg[4]])          # both draws of the same parents
    src = ids[:c["kept"]].long()
    assert torch.equal(opt.state[pc._rotation]["exp_avg"][:c["kept"]], m["rotation"].to(cuda_dev)[src])
    assert not opt.state[pc._rotation]["exp_avg"][c["kept"]:].any()
    assert not opt.state[pc._features_rest]["exp_avg_sq"][c["kept"]:].any()
    assert getattr(pc, "_e3_ws", None) is None
    # nothing selected, nothing pruned: identity
    P2 = pc._xyz.shape[0]
    before = pc._xyz.detach().clone()
    st.reset(P2, cuda_dev)
    c2 = dn.densify_and_prune(pc, opt, st, 0.0002, 0.0, extent, None)
    assert c2 == {"kept": P2, "cloned": 0, "children_per_draw": 0, "split_selected": 0, "rows": P2}
    assert torch.equal(pc._xyz.detach(), before)
    # everything pruned
    c3 = dn.densify_and_prune(pc, opt, st, 0.0002, 2.0, extent, None)
    assert c3["rows"] == 0 and pc._xyz.shape == (0, 3) and pc._features_rest.shape == (0, 15, 3)
    # a model without optimizer state, and bad arguments
    pc, opt = make_model({k: params[k][:1000] for k in do.NAMES}, None, None, cuda_dev)
    st = dn.DensificationState(1000, cuda_dev)
    st.xyz_gradient_accum.fill_(1.0)
    st.denom.fill_(1.0)
    c4 = dn.densify_and_prune(pc, None, st, 0.0002, 0.005, extent, None)
    assert c4["rows"] == pc._xyz.shape[0] > 0
    with pytest.raises(RuntimeError):
        dn.densify_and_prune(pc, None, dn.DensificationState(5, cuda_dev), 0.0002, 0.005, extent, None)
    with pytest.raises(RuntimeError):
        dn.densify_and_prune(pc, None, st, 0.0002, 0.005, extent, None, split_n=3)
