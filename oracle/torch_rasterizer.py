"""TEST INFRASTRUCTURE ONLY: dense PyTorch-on-CPU transcription of the reference rasterizer.

An independent second statement of the same algorithm, differentiable by autograd, used to
pin the hand-transcribed backward of oracle/raster_oracle.c ("the two
must agree"). Dense O(pixels x Gaussians): small scenes only.

Follows submodules/diff-gaussian-rasterization/cuda_rasterizer/forward.cu:74-113 (computeCov2D),
:118-152 (computeCov3D), :155-256 (preprocessCUDA), :261-374 (renderCUDA) and
rasterizer_impl.cu:70-138 (tile membership, depth order). The reference's backward treats the
0.99 clamp, the 1/255 and power>0 skips and the 1e-4 stop as constants (backward.cu:495-503);
here they are detached masks / straight-through so autograd reproduces exactly that.
"""

import torch


def _cov3d(scales, mod, rots):
    r, x, y, z = rots.unbind(-1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    # math rotation matrix is the transpose of glm's column-major R; Sigma = Rm S^2 Rm^T
    Rm = R  # rows as listed = math rows of the standard quaternion rotation
    S = torch.diag_embed(scales * mod)
    Mm = Rm @ S
    return Mm @ Mm.transpose(1, 2)


def rasterize(means3D, opacities, scales, rotations, colors, viewmatrix, projmatrix, bg, W, H,
              tanfovx, tanfovy, scale_modifier=1.0):
    """All tensors torch (any float dtype, CPU). viewmatrix/projmatrix in the reference's
    transposed storage. Returns (image [C,H,W], radii [P], aux dict)."""
    dt = means3D.dtype
    P = means3D.shape[0]
    V = viewmatrix.to(dt).T       # math matrix
    PM = projmatrix.to(dt).T
    ones = torch.ones(P, 1, dtype=dt)
    hom = torch.cat([means3D, ones], 1)
    t = (hom @ V.T)[:, :3]
    ph = hom @ PM.T
    p_w = 1.0 / (ph[:, 3] + 0.0000001)
    proj = ph[:, :3] * p_w[:, None]
    in_front = t[:, 2] > 0.2
    focal_x = W / (2.0 * tanfovx)
    focal_y = H / (2.0 * tanfovy)
    Sigma = _cov3d(scales, scale_modifier, rotations)
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tz = t[:, 2]
    txtz, tytz = t[:, 0] / tz, t[:, 1] / tz
    # clamp: value clamps, gradient masked (backward.cu:175-176)
    tx = torch.where((txtz < -limx) | (txtz > limx), (txtz.clamp(-limx, limx) * tz).detach() + 0 * tz, t[:, 0])
    ty = torch.where((tytz < -limy) | (tytz > limy), (tytz.clamp(-limy, limy) * tz).detach() + 0 * tz, t[:, 1])
    # the reference lets d(t.x)/d(t.z) flow through `clamp * t.z`? No: backward.cu:262-264 masks
    # only dL_dtx / dL_dty and keeps the dJ/dtz terms evaluated at the clamped t.
    zeros = torch.zeros_like(tz)
    J = torch.stack([focal_x / tz, zeros, -(focal_x * tx) / (tz * tz),
                     zeros, focal_y / tz, -(focal_y * ty) / (tz * tz)], -1).reshape(-1, 2, 3)
    Wm = V[:3, :3]
    Tm = J @ Wm
    cov = Tm @ Sigma @ Tm.transpose(1, 2)
    a = cov[:, 0, 0] + 0.3
    b = cov[:, 0, 1]
    c = cov[:, 1, 1] + 0.3
    det = a * c - b * b
    det_inv = 1.0 / det
    conic = torch.stack([c * det_inv, -b * det_inv, a * det_inv], -1)
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam)).detach()
    px = ((proj[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((proj[:, 1] + 1.0) * H - 1.0) * 0.5
    gx, gy = (W + 15) // 16, (H + 15) // 16
    pxd, pyd = px.detach(), py.detach()
    rminx = torch.clamp(torch.trunc((pxd - radius) / 16), 0, gx)
    rminy = torch.clamp(torch.trunc((pyd - radius) / 16), 0, gy)
    rmaxx = torch.clamp(torch.trunc((pxd + radius + 15) / 16), 0, gx)
    rmaxy = torch.clamp(torch.trunc((pyd + radius + 15) / 16), 0, gy)
    visible = in_front & (det != 0) & ((rmaxx - rminx) * (rmaxy - rminy) > 0)
    radii = torch.where(visible, radius, torch.zeros_like(radius)).to(torch.int32)
    # depth order, ties by index (stable sort), rasterizer_impl.cu:300-308
    order = torch.sort(torch.where(visible, tz.detach().float(), torch.full_like(tz, float("inf")).float()), stable=True).indices
    order = order[visible[order]]
    ys, xs = torch.meshgrid(torch.arange(H, dtype=dt), torch.arange(W, dtype=dt), indexing="ij")
    pixx, pixy = xs.reshape(-1), ys.reshape(-1)
    tilex, tiley = torch.div(pixx, 16, rounding_mode="floor"), torch.div(pixy, 16, rounding_mode="floor")
    o = order
    member = ((tilex[:, None] >= rminx[o][None]) & (tilex[:, None] < rmaxx[o][None]) &
              (tiley[:, None] >= rminy[o][None]) & (tiley[:, None] < rmaxy[o][None]))
    dx = px[o][None, :] - pixx[:, None]
    dy = py[o][None, :] - pixy[:, None]
    con = conic[o]
    power = -0.5 * (con[None, :, 0] * dx * dx + con[None, :, 2] * dy * dy) - con[None, :, 1] * dx * dy
    G = torch.exp(power)
    alpha_raw = opacities.reshape(-1)[o][None, :] * G
    alpha = alpha_raw + (torch.clamp_max(alpha_raw, 0.99) - alpha_raw).detach()   # straight-through
    valid = member & (power.detach() <= 0) & (alpha.detach() >= 1.0 / 255.0)
    a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
    one_m = 1.0 - a_eff
    T_incl = torch.cumprod(one_m, 1)                       # T after each instance
    T_before = torch.cat([torch.ones_like(T_incl[:, :1]), T_incl[:, :-1]], 1)
    stop = valid & (T_incl.detach() < 0.0001)              # test_T < 1e-4 on a blending instance
    alive = (torch.cumsum(stop.to(torch.int32), 1) == 0)   # strictly before (and excluding) the stop
    w = torch.where(alive, a_eff * T_before, torch.zeros_like(a_eff))
    Cimg = w @ colors[o]
    # final T = product over alive instances
    T_final = torch.prod(torch.where(alive, one_m, torch.ones_like(one_m)), 1)
    img = Cimg + T_final[:, None] * bg.to(dt)[None, :]
    aux = dict(final_T=T_final.detach(), order=order, means2D=torch.stack([px, py], -1), conic=conic)
    return img.T.reshape(-1, H, W), radii, aux
