"""Captures golden vectors from the UNMODIFIED reference CUDA build (oracle/_ref/libref_dgr.so).
Runs on the GPU box only:

    gpurun -- 'python tests/golden/make_golden_refcuda.py'        # writes gpurun_out/ref_cuda_*.npz

The files are then committed under tests/golden/ and pin the CPU oracle (oracle/raster_oracle.c)
in the -m "not gpu" suite: integer outputs bit for bit, image / gradients within the stated
tolerances. Inputs are regenerated from the seed by tests/common.scene_inputs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import gpu_common as gc  # noqa: E402
from tests.common import scene_inputs, smooth_dL  # noqa: E402

CASES = {"a": dict(P=1500, W=96, H=64, seed=21, kind="trained", bg=(0.1, 0.3, 0.6)),
         "b": dict(P=800, W=50, H=70, seed=22, kind="init", bg=(0.0, 0.0, 0.0))}

if __name__ == "__main__":
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    for name, cs in CASES.items():
        inp = scene_inputs(cs["P"], cs["W"], cs["H"], seed=cs["seed"], kind=cs["kind"], bg=cs["bg"])
        ref, R, img, radii = gc.ref_forward(inp)
        dL = smooth_dL(3, cs["H"], cs["W"], cs["seed"])
        g = ref.backward(gc.T(dL))
        n = lambda t: t.detach().cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, "ref_cuda_%s.npz" % name), case=np.array(repr(sorted(cs.items()))),
                            num_rendered=np.int64(R), radii=n(radii), tiles_touched=n(ref.state("tiles_touched")),
                            depths=n(ref.state("depths")), means2D=n(ref.state("means2D")), conic_opacity=n(ref.state("conic_opacity")),
                            cov3D=n(ref.state("cov3D")), point_list=n(ref.state("point_list")), point_list_keys=n(ref.state("point_list_keys")),
                            ranges=n(ref.state("ranges")), n_contrib=n(ref.state("n_contrib")), final_T=n(ref.state("accum_alpha")),
                            image=n(img), **{k: n(v) for k, v in g.items()})
        print("wrote", name, "R =", R)
