import sys; sys.path.insert(0, '.')
import torch
from tests.common import scene_inputs, smooth_dL
from tests import gpu_common as gc
inp = scene_inputs(3000, 100, 37, seed=1, bg=(0.2,0.5,0.9))
fw = gc.ours_forward(inp); g = gc.ours_backward(fw, inp, smooth_dL(3, 37, 100))
fw = gc.ours_forward(inp, use_sh=True, sh_degree=3); g = gc.ours_backward(fw, inp, smooth_dL(3, 37, 100))
from event_3dgs_b200 import pipeline as pl, synthetic as syn, fused
from event_3dgs_b200.knn import distCUDA2
dev = torch.device("cuda:0")
for deg, deblur in ((0, False), (2, True)):
    sc = syn.make_scene(dict(P=5001, W=70, H=50, sh_degree=deg, deblur=deblur))
    pc = pl.GaussianParams(sc["gaussians"], dev); cams = [pl.make_camera(c, dev) for c in sc["cameras"][:6]]
    gt = {k: torch.as_tensor(v).to(dev) for k, v in sc["gt"].items()}
    c = torch.nn.Parameter(torch.tensor(0.17, device=dev)); bg = torch.zeros(3, device=dev)
    l = fused.event_iteration_backward(pc, cams[2], cams[2], cams[3], gt, c, bg, deblur=deblur)
    l = fused.event_iteration_backward(pc, cams[2], cams[2], cams[3], gt, c, bg, deblur=deblur, dedupe_views=True, gray_event_views=False)
    from event_3dgs_b200.optim import FusedAdam
    opt = FusedAdam([{"params": [p], "lr": 1e-3} for p in pc.leaves()], eps=1e-15); opt.step()
d = distCUDA2(torch.rand(7001, 3, device=dev))
torch.cuda.synchronize(); print("done", float(l), float(d.mean()))
