"""A whole training run of BASELINE config c2 (train_colmap_easy-shaped: 300k Gaussians, 346x260, --event,
8000 iterations) through event_3dgs_b200.trainer with the reference's schedule (densification from 500
every 100, SH degree up every 1000, Adam, learnable threshold c). GT frames are renders of a teacher
model, so the run has something to learn. Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(iterations=8000, config="c2"):
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    from event_3dgs_b200 import trainer as tr
    dev = torch.device("cuda", 0)
    sc = syn.make_scene(config)
    teacher = pl.GaussianParams(sc["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in sc["cameras"]]
    bg = torch.zeros(3, device=dev)
    with torch.no_grad():      # 8-bit frames, as a dataset holds them
        frames = [(pl.render(c, teacher, pl.PIPE, bg)["render"].clamp(0, 1) * 255).round().to(torch.uint8).contiguous() for c in cams]
    g = {k: np.array(v, copy=True) for k, v in sc["gaussians"].items() if isinstance(v, np.ndarray)}
    g["active_sh_degree"] = 0
    rng = np.random.default_rng(0)
    keep = rng.random(g["xyz"].shape[0]) < 0.5            # start from half of the teacher's points, perturbed
    for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity"):
        g[k] = g[k][keep]
    g["features_dc"] = g["features_dc"] + rng.normal(0, 0.3, g["features_dc"].shape).astype(np.float32)
    student = pl.GaussianParams(g, dev)
    gt_for = lambda i: {"intensity": frames[i], "event_now": frames[i], "event_next": frames[i + 1]}
    t = tr.Trainer(student, cams, cams, gt_for, event=True, extent=2.6, seed=0)
    rows, losses = [int(student._xyz.shape[0])], []
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in range(1, iterations + 1):
        loss = t.step()
        if it % 500 == 0:
            losses.append(float(loss))
            rows.append(int(student._xyz.shape[0]))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(json.dumps({"config": config, "iterations": iterations, "seconds": dt, "it_per_s": iterations / dt,
                      "rows_every_500": rows, "loss_every_500": losses, "sh_degree_end": int(student.active_sh_degree),
                      "c_end": float(t.c.detach())}))


if __name__ == "__main__":
    main()
