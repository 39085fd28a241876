#!/usr/bin/env python
"""bench.py -- train iters/sec of the Event-3DGS hot path (3 renders fwd + event loss + backward
to leaf-parameter gradients per event view-pair) on synthetic scenes of the BASELINE.json shapes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0). See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
# NCCL_DEBUG is left as the launcher set it (its INIT lines carry the communicator's rank count); the result
# is the one line of stdout that starts with '{"metric"' or '{"impl"'

import numpy as np  # noqa: E402

METRIC = "train_iters_per_sec"
UNIT = "it/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="c3", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--path", default="auto", choices=["auto", "dropin", "fused"])
    ap.add_argument("--kind", default="trained", choices=["trained", "init"])
    ap.add_argument("--P", type=int, default=None, help="override the number of Gaussians (debugging)")
    ap.add_argument("--pairs", type=int, default=None,
                    help="event view-pairs per optimizer step over all ranks (default: one per rank; c5: 64)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "nccl"],
                    help="multi-GPU gradient exchange: auto = own NVLS all-reduce kernel from 4 GPUs on when the box supports multicast, else NCCL")
    ap.add_argument("--dropin-no-sync", action="store_true",
                    help="with --path dropin: the opt-in forward without the host read-back of num_rendered (E3_DROPIN_NO_SYNC)")
    ap.add_argument("--reference-on-cpu", action="store_true",
                    help="with --impl reference: time the C port on the host cores even if the reference CUDA build is loadable")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=25.0)
    return ap.parse_args()


# --------------------------------------------------------------------------------------------
# shared: algorithmic bytes and workload description
# --------------------------------------------------------------------------------------------
def higher_msb(n):
    msb, step = 16, 16
    while step > 1:
        step //= 2
        msb = msb + step if (n >> msb) else msb - step
    if n >> msb:
        msb += 1
    return msb


def bytes_render(P, R, N, T):
    """B_render of SURVEY.md 8d (reference algorithm, colours precomputed)."""
    passes = -(-(32 + higher_msb(T)) // 8)
    return 508 * P + (24 * passes + 108) * R + 40 * N + 32 * T


def bytes_bwd_blend(P, R, N, T):
    """Algorithmic bytes of ONE launch of the backward blend kernel: per instance the Gaussian id
    (4) + xy (8) + conic/opacity (16) + colour (12); per pixel final_T, n_contrib, dL/dpixel (20);
    per tile its range (8); per Gaussian the 9 accumulated gradients (36)."""
    return 40 * R + 20 * N + 8 * T + 36 * P


def workload(cfg, name, kind, pairs):
    return {"workload": "%s: %d Gaussians, %dx%d, %s, event view-pair = 3 renders fwd + event loss%s + bwd" % (
        name, cfg["P"], cfg["W"], cfg["H"], "RGB SH deg %d" % cfg["sh_degree"] if cfg["sh_degree"] else "gray SH deg 0",
        " + blurry L1" if cfg["deblur"] else ""),
        "gaussians": cfg["P"], "width": cfg["W"], "height": cfg["H"], "sh_degree": cfg["sh_degree"],
        "deblur": bool(cfg["deblur"]), "params": kind, "pairs_per_step": pairs,
        "l2": "inputs of one step (>150 MB of Gaussian state and instance lists at c3) exceed the 126 MB L2; "
              "a different view triplet is rendered every step"}


# --------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md recipe)
# --------------------------------------------------------------------------------------------
class Clocks:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
        "clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                mx = float(f[2])
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[1]))
                    for n, v in zip(names, f[4:8]):
                        if v.lower().startswith("active"):
                            reasons.add(n)
            except ValueError:
                continue
        if not sm:          # region shorter than the sampling period: take everything we saw
            for ts, line in self.rows:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[1]))
                except Exception:
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle (C restatement, OpenMP over all host cores) on a bounded sample
# --------------------------------------------------------------------------------------------
def cpu_iteration(scene, triplet, rows, deblur):
    """One event iteration on the CPU oracle; the blend loops only cover the first `rows` pixel
    rows. Returns (seconds for the per-Gaussian/binning stages, seconds for the blends)."""
    from event_3dgs_b200 import synthetic as syn
    from oracle import loss_oracle as lo
    from oracle import raster_oracle as ro
    a = syn.activate(scene["gaussians"])
    cfg = scene["cfg"]
    W, H = cfg["W"], cfg["H"]
    L = ro.lib()
    sts, t_geo, t_blend = [], 0.0, 0.0
    colors = np.clip(a["shs"][:, 0, :] * 0.28209479177387814 + 0.5, 0, None).astype(np.float32)   # SH deg 0 colours
    bg = np.zeros(3, np.float32)
    for v in triplet:
        cam = scene["cameras"][v]
        t0 = time.perf_counter()
        L.orc_set_pixel_window(0, 0)
        st = ro.rasterize_forward(a["means3D"], a["opacities"], cam["world_view_transform"], cam["full_proj_transform"],
                                  cam["camera_center"], bg, W, H, cam["tanfovx"], cam["tanfovy"], scales=a["scales"],
                                  rotations=a["rotations"], colors_precomp=colors)
        t1 = time.perf_counter()
        L.orc_set_pixel_window(0, rows * W)
        feat = st["colors_precomp"]
        L.orc_render_forward(3, st["gx"], st["gy"], ro._p(st["ranges"]), ro._p(st["point_list"]), W, H, ro._p(st["means2D"]),
                             ro._p(feat), ro._p(st["conic_opacity"]), ro._p(st["final_T"]), ro._p(st["n_contrib"]),
                             ro._p(bg), ro._p(st["out_color"]))
        t2 = time.perf_counter()
        t_geo += t1 - t0
        t_blend += t2 - t1
        sts.append(st)
    t0 = time.perf_counter()
    gt = scene["gt"]
    loss, gr = lo.event_loss(sts[0]["out_color"], sts[1]["out_color"], sts[2]["out_color"], gt["event_now"],
                             gt["event_next"], gt["intensity"], 0.17, gt["blurry"] if deblur else None)
    t_geo += time.perf_counter() - t0
    for st, k in zip(sts, ("image", "image_now", "image_next")):
        t0 = time.perf_counter()
        g = ro.rasterize_backward(st, gr[k].astype(np.float32))      # honours the pixel window
        t_blend += time.perf_counter() - t0
        del g
    L.orc_set_pixel_window(0, 0x7fffffff)
    return t_geo, t_blend


def cpu_arm(scene, steps, warmup, budget_s):
    """Times the oracle; every step is the full per-Gaussian work plus a band of pixel rows of
    the blends, extrapolated to the whole frame (blend cost is proportional to rows)."""
    H = scene["cfg"]["H"]
    from oracle import raster_oracle as ro
    # set explicitly: under torch.distributed.run the process inherits OMP_NUM_THREADS=1; `cores` is the
    # count OpenMP reports back, i.e. the threads really used
    cores = ro.set_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    deblur = scene["cfg"]["deblur"]
    probe_rows = max(16, (H // 16 // 8) * 16)
    tg, tb = cpu_iteration(scene, (3, 3, 4), probe_rows, deblur)          # calibration (also warms caches)
    per_row = tb / probe_rows
    per_step_budget = budget_s / max(1, steps + max(0, warmup - 1))
    rows = int(max(16, min(H, ((per_step_budget - tg) / max(per_row, 1e-9)) // 16 * 16)))
    for _ in range(max(0, warmup - 1)):
        cpu_iteration(scene, (3, 3, 4), rows, deblur)
    est = []
    for s in range(steps):
        v = 2 + (7 * s) % 80
        tg, tb = cpu_iteration(scene, (v, v, v + 1), rows, deblur)
        est.append(tg + tb * (H / float(rows)))
    t_iter = float(np.mean(est))
    sample = "C oracle (OpenMP, %d threads): full projection/binning of all Gaussians for 3 views + event loss, " \
             "blend fwd+bwd over the first %d of %d pixel rows, extrapolated to the frame; %d steps" % (cores, rows, H, steps)
    return 1.0 / t_iter, t_iter, cores, sample


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def main():
    args = parse()
    from event_3dgs_b200 import synthetic as syn
    cfg_name = args.config
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        # rank 0 only. The reference's own implementation of the path: its CUDA sources compiled unmodified for
        # sm_100a (oracle/_ref) under its render()/loss flow when a GPU and the library are there -- the bar the
        # north star names; otherwise its algorithm on the host cores (C port of the oracle)
        if rank != 0:
            return
        line = None
        if not args.reference_on_cpu:
            try:
                line = reference_cuda_arm(args, cfg_name, world)
            except Exception as ex:
                sys.stderr.write("bench.py: reference CUDA build not usable (%r); timing the CPU port instead\n" % (ex,))
        if line is not None:
            print(json.dumps(line))
            return
        scene = syn.make_scene(cfg_name, kind=args.kind, P=args.P)
        v, t_iter, cores, sample = cpu_arm(scene, max(1, args.steps), max(1, args.warmup), 150.0)
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * t_iter, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload(scene["cfg"], cfg_name, args.kind, 1),
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    from event_3dgs_b200 import _lib, pipeline as pl
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()

    scene = syn.make_scene(cfg_name, kind=args.kind, P=args.P)
    cfg = scene["cfg"]
    W, H, P = cfg["W"], cfg["H"], cfg["P"]
    N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    deblur = cfg["deblur"]
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt_host = {k: torch.as_tensor(v).pin_memory() for k, v in scene["gt"].items()}
    gt_dev = {k: v.to(dev) for k, v in gt_host.items()}
    bg = torch.zeros(3, device=dev)
    c_param = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    leaves = pc.leaves() + [c_param]

    use_fused = False
    if args.path in ("auto", "fused"):
        try:
            from event_3dgs_b200 import fused  # noqa: F401
            use_fused = True
        except ImportError:
            if args.path == "fused":
                raise
    path_name = "fused" if use_fused else "dropin"
    if args.dropin_no_sync:
        from event_3dgs_b200 import rasterizer as _rz
        _rz.NO_SYNC = True
        path_name += " (no-sync forward, opt-in)"

    from event_3dgs_b200 import parallel
    pairs = args.pairs or (64 if cfg_name == "c5" else world)
    if pairs % world:
        raise SystemExit("--pairs must be a multiple of the number of ranks")
    my_pairs = parallel.shard_pairs(pairs, rank, world)       # pair b of a step -> rank b mod G

    def view_of(step, b=None):
        b = rank if b is None else b
        return 2 + (7 * (step * pairs + b)) % 80

    exchange_used = {"kind": "none" if world == 1 else "nccl all-reduce (AVG) of the flat gradient buffer"}
    overflowed = []          # steps whose instance count exceeded the workspace (lazy check): reported in the line

    def iteration(step, gt, sync_grads=True, gt_ready=None):
        for p in leaves:
            p.grad = None
        loss = None
        ex = None
        if world >= 4 and use_fused and args.exchange != "nccl" and fused._workspace_for(pc, cams[view_of(step, my_pairs[0])]) is not None:
            # our own in-switch all-reduce kernel over the NVLS multicast alias of the gradient buffer
            # (parallel.FusedGradientExchange, e3_nvls_allreduce); NCCL when the box has no multicast
            ex = parallel.fused_exchange_for(pc, fused.grad_flat_numel(pc, int(pc.active_sh_degree)), dev)
        if use_fused:
            pc._grad_exchange = ex
        for k, b in enumerate(my_pairs):
            v = view_of(step, b)
            if use_fused:
                # lazy_check: num_rendered of a step is read (and checked against the workspace capacity) two steps
                # later; drain_workspace() below verifies that no timed step had exceeded it
                call = lambda: fused.event_iteration_backward(pc, cams[v], cams[v], cams[v + 1], gt, c_param, bg, deblur=deblur,
                                                              gt_ready=gt_ready if k == 0 else None, accumulate=k > 0, lazy_check=True)
                try:
                    l = call()
                except fused.Workspace.Overflow as ex:      # reported two steps late; the buffers have been enlarged
                    overflowed.append(str(ex))
                    l = call()
            else:
                l, pkg = pl.event_iteration(pc, cams[v], cams[v], cams[v + 1], gt, c_param, bg, deblur=deblur)
                l.backward()              # autograd accumulates into .grad over the pairs
            loss = l if loss is None else loss + l
        if ex is not None:
            if sync_grads:
                ex.allreduce(1.0 / pairs)
                exchange_used["kind"] = "e3_nvls_allreduce (multimem.ld_reduce / multimem.st over the NVLS multicast alias), mean over all pairs"
            return loss
        if len(my_pairs) > 1:             # mean over this rank's pairs (the all-reduce averages over ranks)
            if use_fused:
                pc._grad_flat.div_(len(my_pairs))
            else:
                for p in leaves:
                    p.grad.div_(len(my_pairs))
        if world > 1 and sync_grads:
            if use_fused:
                # one collective over the flat gradient buffer (at SH degree 0 it does not even contain
                # the identically-zero features_rest gradient)
                parallel.allreduce_flat(pc._grad_flat, world)
            else:
                parallel.allreduce_gradients(leaves, world)
        return loss

    def timed(fn, steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record()
        for s in range(steps):
            fn(s)
        e1.record()
        torch.cuda.synchronize()
        w1 = time.time()
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, w0, w1

    # ---- warm-up (also grows the caching allocator to its steady state) ----
    for s in range(max(3, args.warmup)):
        iteration(s, gt_dev)
    torch.cuda.synchronize()

    if world >= 4 and use_fused and args.exchange == "auto":
        # (at 2 GPUs NCCL's peer-to-peer path moves 2/3 of the bytes of the in-switch one and is 0.06 ms faster:
        # profiles/r1_exchange_alone_n2.json; from 4 GPUs on the in-switch kernel is used)
        # self-check of the in-switch all-reduce against NCCL on the same step (the two runs differ by the
        # run-to-run spread of the float atomics); any rank failing it sends everyone to NCCL
        args.exchange = "nccl"
        iteration(0, gt_dev)
        ref_flat = pc._grad_flat.clone()
        args.exchange = "auto"
        iteration(0, gt_dev)
        ok = torch.tensor([1.0 if float((pc._grad_flat - ref_flat).abs().max()) <= 2e-3 * float(ref_flat.abs().max()) else 0.0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) < 0.5:
            args.exchange = "nccl"
            exchange_used["kind"] = "nccl all-reduce (AVG) of the flat gradient buffer (the NVLS kernel failed its self-check)"
        del ref_flat
        torch.cuda.synchronize()

    # ---- (1) resident run: inputs already in HBM ----
    clocks = Clocks(local_rank) if rank == 0 else None
    _lib.profile_enable(["render_bwd"])
    _lib.profile_read(reset=True)
    l0 = _lib.lib().e3_kernel_launches()
    ms, w0, w1 = timed(lambda s: iteration(args.warmup + s, gt_dev), args.steps)
    launches = _lib.lib().e3_kernel_launches() - l0
    prof = _lib.profile_read(reset=True)
    _lib.profile_enable([])
    clk = clocks.stop(w0, w1) if clocks else None
    ms_per_step = ms / args.steps
    value = pairs * 1e3 / ms_per_step

    def drain_workspace():
        # a step that produced more instances than the workspace held dropped the excess: say so in the line
        ws = getattr(pc, "_e3_ws", None)
        if use_fused and ws is not None:
            for _ in range(2):
                try:
                    ws.drain()
                    break
                except fused.Workspace.Overflow as ex:
                    overflowed.append(str(ex))
    drain_workspace()

    # ---- (2) end to end through the public API with host buffers ----
    cam_bytes = 3 * (16 + 16 + 3) * 4

    # host side of the e2e run: the GT frames as 8-bit pixels in pinned memory (what the dataset holds
    # on disk); the fused path consumes them as they are, the drop-in path converts them on the device
    gt_host_u8 = {k: torch.as_tensor(np.round(v * 255.0).astype(np.uint8)).pin_memory() for k, v in scene["gt"].items()}
    h2d = sum(v.numel() for k, v in gt_host_u8.items() if deblur or k != "blurry")

    copy_stream = torch.cuda.Stream(dev)
    # the step's result (the loss) comes back through a pinned buffer; it is read on the host one step
    # later, so the host can enqueue step s+1 while step s runs (the reference blocks on loss.item()
    # every iteration, train.py:303). Every step's loss IS copied and read inside the timed region.
    loss_host = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ready = [torch.cuda.Event() for _ in range(2)]
    seen = []

    def e2e_step(s):
        main = torch.cuda.current_stream(dev)
        if use_fused:
            # the copy runs on its own stream and is awaited only by the loss kernel; the frames are fresh
            # allocations kept alive for `main` by record_stream, so the copy of step s+1 overlaps step s
            with torch.cuda.stream(copy_stream):
                gt = {k: v.to(dev, non_blocking=True) for k, v in gt_host_u8.items() if deblur or k != "blurry"}
                ready = torch.cuda.Event()
                ready.record(copy_stream)
            for t in gt.values():
                t.record_stream(main)
            loss = iteration(args.warmup + s, gt, gt_ready=ready)
        else:
            gt = {k: v.to(dev, non_blocking=True).float() / 255.0 for k, v in gt_host_u8.items() if deblur or k != "blurry"}
            loss = iteration(args.warmup + s, gt)
        b = s & 1
        if s > 0:                                    # device -> host read of the previous step's result
            loss_ready[b ^ 1].synchronize()
            seen.append(float(loss_host[b ^ 1]))
        loss_host[b].copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ready[b].record(main)

    def e2e_run(s):
        e2e_step(s)
        if s == args.steps - 1:                      # the last step's loss is read inside the window too
            loss_ready[s & 1].synchronize()
            seen.append(float(loss_host[s & 1]))

    for s in range(2):
        e2e_step(s)
    torch.cuda.synchronize()
    seen.clear()
    ms_e2e, _, _ = timed(e2e_run, args.steps)
    drain_workspace()
    assert len(seen) == args.steps and all(np.isfinite(seen)), "every step's loss must have been read"
    e2e_value = pairs * 1e3 / (ms_e2e / args.steps)

    # ---- (3) per-stage breakdown and workload statistics (untimed pass) ----
    _lib.profile_enable(_lib.STAGES)
    _lib.profile_read(reset=True)
    nprof = 3
    for s in range(nprof):
        iteration(args.warmup + s, gt_dev, sync_grads=False)
    torch.cuda.synchronize()
    stages = {k: v[0] / (nprof * len(my_pairs)) for k, v in _lib.profile_read(reset=True).items()}
    _lib.profile_enable([])
    # measured num_rendered of the three renders of one step
    Rs = []
    with torch.no_grad():
        v = view_of(args.warmup)
        for cam in (cams[v], cams[v], cams[v + 1]):
            pkg = pl.render(cam, pc, pl.PIPE, bg)
            Rs.append(int(last_num_rendered()))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if pairs == world else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload(cfg, cfg_name, args.kind, pairs)}
    line["config"]["path"] = path_name
    line["config"]["gradient_exchange"] = exchange_used["kind"]
    line["config"]["num_rendered"] = Rs
    if world > 1:
        line["comm"] = {"backend": dist.get_backend(), "nranks": dist.get_world_size()}
    line["mpix_per_sec"] = value * 3 * N / 1e6
    line["config"]["unit_note"] = "one iteration = one event view-pair (3 renders fwd + event loss + bwd); value counts view-pairs per second over all ranks"
    line["e2e"] = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d + cam_bytes), "d2h_bytes_per_step": 4,
                   "ms_per_step": ms_e2e / args.steps}
    line["gpu_launches"] = int(launches)
    line["workspace_overflows_in_timed_steps"] = len(overflowed)
    line["clocks"] = clk
    # roofline of the dominant kernel (backward blend), timed live with CUDA events in run (1)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    if "render_bwd" in prof and prof["render_bwd"][1] > 0:
        t_ms, n = prof["render_bwd"]
        per_launch_ms = t_ms / n
        algo = float(np.mean([bytes_bwd_blend(P, r, N, T) for r in Rs]))
        ach = algo / (per_launch_ms * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum and warp instructions per launch come from the committed
        # ncu --set full capture; the file carries the hash of the kernel source it was captured from and is
        # ignored (traffic null) once render.cu has changed since
        tj, traffic_file = load_traffic()
        traffic, traffic_src = None, None
        def kernel_key(prefix):        # template arguments were added over time: <1>, <1, 0>, <1, 0, 0> name the same variant
            return next((k for k in tj if k.startswith(prefix + ">") or k.startswith(prefix + ", 0")), prefix + ">")
        kname = kernel_key("render_bwd_kernel<1")
        if cfg_name == "c3" and kname in tj:
            traffic = tj[kname]["dram_bytes_per_launch"]
            traffic_src = "%s (ncu --set full, c3, %s: the single-channel variant, 2 of the 3 launches per step)" % (traffic_file, kname)
        elif cfg_name == "c3":
            traffic_src = "no ncu capture of the current render.cu is committed (stale captures are refused)"
        line["roofline"] = {"bound": "hbm", "kernel": "render_bwd_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                            "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src, "launch_ms": per_launch_ms, "launches": int(n),
                            "algorithmic_bytes_per_launch": algo,
                            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                            "note": "the blend kernels are issue/latency bound, not HBM bound; see DESIGN.md"}
        # the blend kernels are issue bound: the second roofline is the warp-instruction rate, with the
        # instruction count per launch taken from the same committed ncu capture
        wi = None
        try:
            k3 = kernel_key("render_bwd_kernel<3")
            if cfg_name == "c3":
                wi = (tj[k3]["warp_inst_per_launch"] + 2 * tj[kname]["warp_inst_per_launch"]) / 3
        except Exception:
            wi = None
        if wi and clk and clk.get("sm_mhz"):
            peak_issue = 148 * 4 * clk["sm_mhz"] * 1e6
            line["roofline_issue"] = {"bound": "issue", "kernel": "render_bwd_kernel", "achieved": wi / (per_launch_ms * 1e-3),
                                      "peak": peak_issue, "unit": "warp instructions/s", "frac": wi / (per_launch_ms * 1e-3) / peak_issue,
                                      "note": "warp instructions per launch from %s (ncu), mean over the one C=3 and two C=1 launches of a step; launch time live (three concurrent streams share the SMs, so this is a lower bound of the kernel's own rate), peak = 148 SMs x 4 schedulers x SM clock under load" % traffic_file}
    b_iter = sum(bytes_render(P, r, N, T) for r in Rs) + 108 * N + (24 * N if deblur else 0)
    b_step = b_iter * len(my_pairs)          # one step of this GPU = len(my_pairs) view-pairs
    line["iter_roofline"] = {"algorithmic_bytes_per_pair": b_iter, "pairs_per_gpu_per_step": len(my_pairs),
                             "achieved_gbs": b_step / (ms_per_step * 1e-3) / 1e9,
                             "frac": b_step / (ms_per_step * 1e-3) / 1e9 / peak,
                             "note": "per GPU: B_iter of SURVEY 8d (the reference algorithm's bytes) x the view-pairs this GPU processes per step / step time"}
    line["stages_ms_per_iter"] = stages
    if use_fused:
        # extra (not the headline): the reference's train and event camera lists share their poses
        # (scene/dataset_readers.py:145-160), so the intensity view and the first event view are the
        # same view; rendering it once is loss- and gradient-equivalent (fused.same_view)
        def it_dedupe(s):
            v = view_of(args.warmup + s)
            fused.event_iteration_backward(pc, cams[v], cams[v], cams[v + 1], gt_dev, c_param, bg, deblur=deblur, dedupe_views=True, lazy_check=True)
        for s in range(3):
            it_dedupe(s)
        ms_d, _, _ = timed(it_dedupe, args.steps)
        line["dedupe_identical_views"] = {"value": world * 1e3 / (ms_d / args.steps), "unit": UNIT, "ms_per_step": ms_d / args.steps,
                                          "note": "not the headline: the intensity view and the first event view are the same camera and are rendered once; the backward blend of the shared view carries two seeds, so loss, parameter gradients and the densification statistic equal the three-render iteration (tests); no gradient exchange in this extra"}

    if use_fused and world == 1:
        # extra: the whole training step of train.py:95-332 (iteration + densification statistics + both
        # Adam steps) through event_3dgs_b200.trainer on a copy of the model, densification itself
        # outside the window (it runs every 100th iteration; timed separately in profiles/r1_densify_bench.json)
        from event_3dgs_b200 import trainer as tr
        pc_t = pl.GaussianParams(scene["gaussians"], dev)
        t_opt = tr.optimization_defaults(densify_from_iter=10 ** 9)
        trn = tr.Trainer(pc_t, cams, cams, lambda i: gt_dev, opt=t_opt, event=True, deblur=deblur, extent=2.6)
        for s in range(3):
            trn.step()
        ms_t, _, _ = timed(lambda s: trn.step(), args.steps)
        line["train_loop"] = {"value": 1e3 / (ms_t / args.steps), "unit": UNIT, "ms_per_step": ms_t / args.steps,
                              "what": "iteration + densification statistics + fused Adam (6 groups) + Adam(c), one view-pair per step; Trainer default: the intensity view and the first event view (same camera) are rendered once, results as with three renders"}
        trn3 = tr.Trainer(pl.GaussianParams(scene["gaussians"], dev), cams, cams, lambda i: gt_dev, opt=tr.optimization_defaults(densify_from_iter=10 ** 9),
                          event=True, deblur=deblur, extent=2.6, dedupe_views=False)
        for s in range(3):
            trn3.step()
        ms_t3, _, _ = timed(lambda s: trn3.step(), args.steps)
        line["train_loop"]["three_renders"] = {"value": 1e3 / (ms_t3 / args.steps), "unit": UNIT, "ms_per_step": ms_t3 / args.steps}
        del trn3
        del trn, pc_t

    if use_fused and world > 1:
        # extra: the whole training step on all ranks (one view-pair per rank and step): iteration + densification
        # statistics + gradient exchange + Adam, (a) replicated optimizer behind the all-reduce, (b) exchange and Adam as one
        # kernel per rank (e3_nvls_sharded_adam: in-switch sum of 1/world of the rows -> Adam -> multicast store)
        from event_3dgs_b200 import trainer as tr
        res = {}
        for tag, sharded in (("replicated_adam", False), ("sharded_adam", True)):
            pc_t = pl.GaussianParams(scene["gaussians"], dev)
            trn = tr.Trainer(pc_t, cams, cams, lambda i: gt_dev, opt=tr.optimization_defaults(densify_from_iter=10 ** 9), event=True, deblur=deblur,
                             extent=2.6, pairs_per_step=world, rank=rank, world=world, dedupe_views=False, sharded_optimizer=sharded)
            for s in range(4):
                trn.step()
            ms_t, _, _ = timed(lambda s: trn.step(), args.steps)
            res[tag] = {"value": world * 1e3 / (ms_t / args.steps), "unit": UNIT, "ms_per_step": ms_t / args.steps, "exchange": trn.last_exchange}
            del trn, pc_t
        line["train_loop"] = dict(res, what="three renders per pair, one pair per rank and step, statistics + exchange + Adam on every rank")

    if use_fused and world == 1:
        # extra: the non-event branch of train.py (one render, L1 + SSIM on the luma, :213-223) through
        # fused.frame_iteration_backward (fused SSIM kernel, no-sync workspace)
        def it_frame(s):
            v = view_of(args.warmup + s)
            fused.frame_iteration_backward(pc, cams[v], gt_dev["intensity"], bg, 0.2, True)
        for s in range(3):
            it_frame(s)
        ms_f, _, _ = timed(it_frame, args.steps)
        line["frame_mode_gray"] = {"value": 1e3 / (ms_f / args.steps), "unit": UNIT, "ms_per_step": ms_f / args.steps,
                                   "what": "1 render fwd + 0.8 L1 + 0.2 (1 - SSIM) on the luma + bwd (train.py --gray without --event)"}

    if rank == 0 and world == 1:
        # reference CUDA build (oracle/_ref) on the same tensors, same timing window
        if not args.no_ref_cuda:
            try:
                line["ref_cuda"] = ref_cuda_leg(pc, cams, gt_dev, c_param, bg, deblur, args, view_of, timed, leaves)
                line["ref_cuda"]["speedup"] = value / line["ref_cuda"]["value"]
            except Exception as ex:   # the checker is optional for the number
                line["ref_cuda"] = {"unavailable": repr(ex)[:200]}
        if not args.no_cpu_baseline:
            v, t_iter, cores, sample = cpu_arm(scene, 2, 1, args.cpu_budget_s)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def reference_cuda_arm(args, cfg_name, world):
    """bench.py --impl reference on a GPU box: the UNMODIFIED reference rasterizer (oracle/_ref: forward.cu,
    backward.cu, rasterizer_impl.cu compiled for sm_100a) under the reference's own flow -- PyTorch activations,
    eval_sh, render(), the event loss INCLUDING the ssim_gray evaluation it multiplies by 0 (train.py:177-178),
    loss.backward() -- one view-pair per step on ONE GPU (the reference has no multi-GPU path). None of this
    repository's kernels run here."""
    import torch
    from event_3dgs_b200 import pipeline as pl
    from event_3dgs_b200 import synthetic as syn
    from oracle import ref_cuda
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device")
    if not ref_cuda.available():
        raise RuntimeError("oracle/_ref/libref_dgr.so not built")
    ref_cuda.lib()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    scene = syn.make_scene(cfg_name, kind=args.kind, P=args.P)
    cfg = scene["cfg"]
    deblur = cfg["deblur"]
    pc = pl.GaussianParams(scene["gaussians"], dev)
    cams = [pl.make_camera(c, dev) for c in scene["cameras"]]
    gt_dev = {k: torch.as_tensor(v).to(dev) for k, v in scene["gt"].items()}
    gt_host_u8 = {k: torch.as_tensor(np.round(v * 255.0).astype(np.uint8)).pin_memory() for k, v in scene["gt"].items()
                  if deblur or k != "blurry"}
    bg = torch.zeros(3, device=dev)
    c_param = torch.nn.Parameter(torch.tensor(0.17, device=dev))
    leaves = pc.leaves() + [c_param]
    pairs = args.pairs or (64 if cfg_name == "c5" else world)

    def render_ref(cam, pc_, pipe, bg_, **kw):
        return pl.render(cam, pc_, pipe, bg_, rasterizer_cls=ref_cuda.RefGaussianRasterizer)

    def it(s, gt):
        v = 2 + (7 * s) % 80
        for p in leaves:
            p.grad = None
        loss, _ = pl.event_iteration(pc, cams[v], cams[v], cams[v + 1], gt, c_param, bg, deblur=deblur, render_fn=render_ref,
                                     dead_ssim=True)
        loss.backward()
        return loss

    def timed(fn, steps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record()
        for s in range(steps):
            fn(s)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), w0, time.time()

    W = max(3, args.warmup)
    for s in range(W):
        it(s, gt_dev)
    clocks = Clocks(dev.index or 0)
    ms, w0, w1 = timed(lambda s: it(W + s, gt_dev), args.steps)
    clk = clocks.stop(w0, w1)
    value = 1e3 / (ms / args.steps)

    def e2e_step(s):
        # the frames as they are on disk (8-bit) from pinned host memory, /255 on the device as the reference's
        # loader does on the host (utils/general_utils.py:21-27); loss.item() every iteration (train.py:303)
        gt = {k: v.to(dev, non_blocking=True).float() / 255.0 for k, v in gt_host_u8.items()}
        return float(it(W + s, gt).item())

    for s in range(2):
        e2e_step(s)
    ms_e, _, _ = timed(e2e_step, args.steps)
    e2e_value = 1e3 / (ms_e / args.steps)
    sample = "reference CUDA sources (oracle/_ref, sm_100a) + PyTorch SH / activations / event loss incl. the dead ssim_gray call + " \
             "autograd, one view-pair per step on one GPU; %d timed steps, host work on 1 thread" % args.steps
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload(cfg, cfg_name, args.kind, pairs),
            "reference_class": "cuda (the reference's own kernels compiled for sm_100a; it has no multi-GPU path, so it always runs on one GPU)",
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "reference", "sample": sample},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(sum(v.numel() for v in gt_host_u8.values()) + 3 * 35 * 4),
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e / args.steps},
            "gpu_launches": 0, "clocks": clk}
    return line


def load_traffic():
    """Newest profiles/r*_traffic.json whose recorded sha256 of csrc/render.cu equals the current file's
    (tools/ncu_summary.py stamps it). Returns ({}, None) when none matches: stale counters are not reported."""
    import glob
    import hashlib
    try:
        cur = hashlib.sha256(open(os.path.join(ROOT, "event_3dgs_b200", "csrc", "render.cu"), "rb").read()).hexdigest()
    except OSError:
        return {}, None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic*.json")), reverse=True):
        try:
            tj = json.load(open(path))
        except Exception:
            continue
        if tj.get("_source_sha256", {}).get("render.cu") == cur:
            return tj, os.path.relpath(path, ROOT)
    return {}, None


def last_num_rendered():
    from event_3dgs_b200 import rasterizer
    rasterizer.verify_no_sync(True)
    return rasterizer.LAST_NUM_RENDERED


def ref_cuda_leg(pc, cams, gt, c_param, bg, deblur, args, view_of, timed, leaves):
    """The unmodified reference rasterizer (oracle/_ref) under the reference's own render() /
    loss flow (PyTorch SH + activations + loss), timed exactly like run (1)."""
    import torch
    from event_3dgs_b200 import pipeline as pl
    from oracle import ref_cuda
    if not ref_cuda.available():
        return {"unavailable": "oracle/_ref/libref_dgr.so not built"}

    def render_ref(cam, pc_, pipe, bg_, **kw):
        return pl.render(cam, pc_, pipe, bg_, rasterizer_cls=ref_cuda.RefGaussianRasterizer)

    def it(s):
        v = view_of(args.warmup + s)
        for p in leaves:
            p.grad = None
        loss, _ = pl.event_iteration(pc, cams[v], cams[v], cams[v + 1], gt, c_param, bg, deblur=deblur, render_fn=render_ref, dead_ssim=True)
        loss.backward()

    for s in range(3):
        it(s)
    steps = max(3, min(args.steps, 10))
    ms, _, _ = timed(it, steps)
    out = {"value": 1e3 / (ms / steps), "unit": UNIT, "ms_per_step": ms / steps, "steps": steps,
           "what": "reference diff-gaussian-rasterization sources compiled for sm_100a (oracle/_ref) + PyTorch SH/loss incl. the ssim_gray evaluation the reference multiplies by 0 (train.py:177-178)"}
    # the reference's whole training step: the same iteration + its densification statistics
    # (train.py:319-320) + torch.optim.Adam on its six groups + Adam(c)
    from event_3dgs_b200 import trainer as tr
    pc_r = pl.GaussianParams({k: v.detach().cpu().numpy() for k, v in
                              dict(xyz=pc._xyz, features_dc=pc._features_dc, features_rest=pc._features_rest, scaling=pc._scaling,
                                   rotation=pc._rotation, opacity=pc._opacity).items()}, pc._xyz.device)
    pc_r.active_sh_degree = pc.active_sh_degree
    groups = tr.training_setup(pc_r, tr.optimization_defaults(), 2.6).param_groups
    opt_r = torch.optim.Adam([{"params": g["params"], "lr": g["lr"], "name": g["name"]} for g in groups], lr=0.0, eps=1e-15)
    c_r = torch.nn.Parameter(torch.tensor(0.17, device=pc._xyz.device))
    opt_c = torch.optim.Adam([c_r], lr=0.1)
    P = pc_r._xyz.shape[0]
    max_radii2D = torch.zeros(P, device=pc._xyz.device)
    accum, denom = torch.zeros(P, 1, device=pc._xyz.device), torch.zeros(P, 1, device=pc._xyz.device)

    def train_it(s):
        v = view_of(args.warmup + s)
        loss, pkg = pl.event_iteration(pc_r, cams[v], cams[v], cams[v + 1], gt, c_r, bg, deblur=deblur, render_fn=render_ref, dead_ssim=True)
        loss.backward()
        with torch.no_grad():
            vis, radii, vsp = pkg["visibility_filter"], pkg["radii"], pkg["viewspace_points"]
            max_radii2D[vis] = torch.max(max_radii2D[vis], radii[vis])
            accum[vis] += torch.norm(vsp.grad[vis, :2], dim=-1, keepdim=True)
            denom[vis] += 1
            opt_r.step()
            opt_r.zero_grad(set_to_none=True)
            opt_c.step()
            opt_c.zero_grad(set_to_none=True)

    def frame_it(s):
        v = view_of(args.warmup + s)
        for p in leaves:
            p.grad = None
        pkg = render_ref(cams[v], pc, pl.PIPE, bg)
        pl.l1_ssim_loss(pkg["render"], gt["intensity"].float() if gt["intensity"].dtype != torch.float32 else gt["intensity"], 0.2, True).backward()

    for s in range(3):
        frame_it(s)
    ms_f, _, _ = timed(frame_it, steps)
    out["frame_mode_gray"] = {"value": 1e3 / (ms_f / steps), "unit": UNIT, "ms_per_step": ms_f / steps,
                              "what": "reference rasterizer + PyTorch SH / L1 / 5-conv SSIM (utils/loss_utils.py:368-385) + autograd"}
    for s in range(3):
        train_it(s)
    ms_t, _, _ = timed(train_it, steps)
    out["train_loop"] = {"value": 1e3 / (ms_t / steps), "unit": UNIT, "ms_per_step": ms_t / steps,
                         "what": "the same + densification statistics + torch.optim.Adam (train.py:317-332)"}
    return out


if __name__ == "__main__":
    main()
