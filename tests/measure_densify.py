"""Times densify_and_prune (csrc/densify.cu) against the reference's sequence of ATen ops (the torch
restatement in oracle/densify_oracle.py, run on the same GPU) at BASELINE c3 size. Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # repo root
from event_3dgs_b200 import densify as dn  # noqa: E402
from oracle import densify_oracle as do  # noqa: E402
from tests.test_densify_gpu import make_model, random_case  # noqa: E402


def main(P=1_000_000, reps=5):
    dev = "cuda:0"
    extent = 3.0
    params, m, v, accum, denom = random_case(P, dev, 11, extent)
    ours, ref = [], []
    rows = None
    for r in range(reps + 1):
        pc, opt = make_model(params, m, v, dev)
        st = dn.DensificationState(P, dev)
        st.xyz_gradient_accum.copy_(accum)
        st.denom.copy_(denom)
        z = torch.randn(2 * P, 3, device=dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        c = dn.densify_and_prune(pc, opt, st, 0.0002, 0.005, extent, 20, normals=lambda n: z[:n])
        torch.cuda.synchronize()
        ours.append(time.perf_counter() - t0)
        rows = c["rows"]
        ora = do.Model({k: params[k].to(dev) for k in do.NAMES}, {k: m[k].to(dev) for k in do.NAMES}, {k: v[k].to(dev) for k in do.NAMES})
        ora.xyz_gradient_accum, ora.denom = accum.to(dev), denom.to(dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ora.densify_and_prune(0.0002, 0.005, extent, 20, lambda n: z[:n])
        torch.cuda.synchronize()
        ref.append(time.perf_counter() - t0)
        del ora, pc, opt
    o, f = float(np.median(ours[1:])), float(np.median(ref[1:]))
    moved = (P + rows) * 59 * 3 * 4
    print(json.dumps({"what": "densify_and_prune, wall clock incl. allocation and the one host sync", "rows_in": P, "rows_out": rows,
                      "ours_ms": o * 1e3, "aten_sequence_ms": f * 1e3, "speedup": f / o, "algorithmic_GB": moved / 1e9,
                      "ours_GBps": moved / o / 1e9}))


if __name__ == "__main__":
    main()
