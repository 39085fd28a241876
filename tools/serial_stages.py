"""Per-stage GPU time of one event iteration with the three views run back to back on ONE stream
(no_sync=False), so that the stage times add up; the default path overlaps the views on three streams."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from event_3dgs_b200 import _lib, fused, pipeline as pl, synthetic as syn  # noqa: E402


def main(cfg="c3", n=5):
    dev = torch.device("cuda", 0)
    scene = syn.make_scene(cfg)
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in scene["gt"].items()}
    bg = torch.zeros(3, device=dev)
    c = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    deblur = scene["cfg"]["deblur"]
    run = lambda s: fused.event_iteration_backward(pc, cams[2 + s], cams[2 + s], cams[3 + s], gt, c, bg, deblur=deblur, no_sync=False)
    for s in range(3):
        run(s)
    torch.cuda.synchronize()
    _lib.profile_enable(_lib.STAGES)
    _lib.profile_read(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(n):
        run(s)
    e1.record()
    torch.cuda.synchronize()
    st = {k: v[0] / n for k, v in _lib.profile_read(reset=True).items()}
    _lib.profile_enable([])
    print(json.dumps({"config": cfg, "ms_per_iter_serial_with_profiling_events": e0.elapsed_time(e1) / n, "stages_ms_per_iter": st,
                      "sum": sum(st.values())}))


if __name__ == "__main__":
    main(*(sys.argv[1:2]))
