"""Host-side mirror of the reference's render() and event-loss assembly, on torch tensors.

  * render()           gaussian_renderer/__init__.py:20-104 (same signature and result dict)
  * eval_sh()          utils/sh_utils.py:57-112 (degrees 0-3)
  * GaussianParams     the activations of scene/gaussian_model.py:95-115 over the six leaf tensors
  * event_loss()       train.py:165-203 with utils/loss_utils.py:24-28,234-249,270-271
  * event_iteration()  the three renders + loss of train.py:144-203

This is the drop-in ("unchanged train.py") path: SH evaluation, activations and the loss run as
PyTorch ops exactly like the reference does; only the rasterizer underneath is ours. The fused
path (activations + SH + loss in CUDA) lives in fused.py.
"""
import math
from types import SimpleNamespace

import torch
import torch.nn as nn

from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    """sh [..., C, (deg+1)^2 or more], dirs [..., 3] unit -> [..., C]."""
    assert 0 <= deg <= 3
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = (result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3])
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                      C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] +
                      C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                          C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                          C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                          C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                          C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


class GaussianParams(nn.Module):
    """The six leaf tensors of the reference's GaussianModel and its activated views."""

    def __init__(self, g, device):
        super().__init__()
        t = lambda a: nn.Parameter(torch.as_tensor(a, dtype=torch.float32).to(device).contiguous())
        self._xyz = t(g["xyz"])
        self._features_dc = t(g["features_dc"])
        self._features_rest = t(g["features_rest"])
        self._scaling = t(g["scaling"])
        self._rotation = t(g["rotation"])
        self._opacity = t(g["opacity"])
        self.active_sh_degree = int(g.get("active_sh_degree", 0))
        self.max_sh_degree = 3

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return torch.nn.functional.normalize(self._rotation)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    def leaves(self):
        return [self._xyz, self._features_dc, self._features_rest, self._scaling, self._rotation, self._opacity]


def create_from_pcd(points, colors, device, sh_degree=3):
    """GaussianModel.create_from_pcd (scene/gaussian_model.py:124-147), the call site of distCUDA2: xyz = the
    points, SH DC = RGB2SH(colours) (utils/sh_utils.py:114-115), the rest 0, isotropic log-scale =
    log(sqrt(clamp_min(distCUDA2(points), 1e-7))), rotation (1,0,0,0), opacity = inverse_sigmoid(0.1).
    points / colors: [P,3] arrays or tensors, colours in [0,1]. Returns GaussianParams at SH degree 0."""
    from .knn import distCUDA2
    pts = torch.as_tensor(points, dtype=torch.float32).to(device).contiguous()
    col = torch.as_tensor(colors, dtype=torch.float32).to(device)
    P = pts.shape[0]
    dist2 = torch.clamp_min(distCUDA2(pts), 0.0000001)
    scales = torch.log(torch.sqrt(dist2))[..., None].repeat(1, 3)
    rots = torch.zeros(P, 4, device=device)
    rots[:, 0] = 1
    x = 0.1 * torch.ones(P, 1, dtype=torch.float32, device=device)
    g = {"xyz": pts, "features_dc": ((col - 0.5) / C0).view(P, 1, 3), "features_rest": torch.zeros(P, (sh_degree + 1) ** 2 - 1, 3, device=device),
         "scaling": scales, "rotation": rots, "opacity": torch.log(x / (1 - x)), "active_sh_degree": 0}
    pc = GaussianParams(g, device)
    pc.max_sh_degree = sh_degree
    return pc


def make_camera(cam, device, image=None):
    """numpy camera dict of synthetic.make_camera -> the attributes render() reads (scene/cameras.py)."""
    t = lambda a: torch.as_tensor(a, dtype=torch.float32).to(device)
    return SimpleNamespace(R=cam.get("R"), T=cam.get("T"), FoVx=cam["FoVx"], FoVy=cam["FoVy"], image_width=cam["image_width"],
                           image_height=cam["image_height"], world_view_transform=t(cam["world_view_transform"]),
                           full_proj_transform=t(cam["full_proj_transform"]), camera_center=t(cam["camera_center"]),
                           original_image=None if image is None else t(image))


PIPE = SimpleNamespace(convert_SHs_python=True, compute_cov3D_python=False, debug=False)


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None,
           rasterizer_cls=GaussianRasterizer, settings_cls=GaussianRasterizationSettings):
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = settings_cls(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False, debug=pipe.debug)
    rasterizer = rasterizer_cls(raster_settings=raster_settings)
    means3D = pc.get_xyz
    means2D = screenspace_points
    opacity = pc.get_opacity
    scales = pc.get_scaling
    rotations = pc.get_rotation
    shs = None
    colors_precomp = None
    if override_color is None:
        # the reference forces convert_SHs_python = True (gaussian_renderer/__init__.py:74)
        shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
        dir_pp = (pc.get_xyz - viewpoint_camera.camera_center.repeat(pc.get_features.shape[0], 1))
        dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
        sh2rgb = eval_sh(pc.active_sh_degree, shs_view, dir_pp_normalized)
        colors_precomp = torch.clamp_min(sh2rgb + 0.5, 0.0)
    else:
        colors_precomp = override_color
    rendered_image, radii = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp,
                                       opacities=opacity, scales=scales, rotations=rotations, cov3D_precomp=None)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii}


def rgb_to_LUVscale(image):
    return (0.4124 * image[0, :, :] + 0.35758 * image[1, :, :] + 0.1804 * image[2, :, :]).unsqueeze(0)


def differentialable_event_simu(image, image_next, _unused, C=0.3):
    img1 = rgb_to_LUVscale(image)
    img2 = rgb_to_LUVscale(image_next)
    epsilon = 1e-8
    return (torch.log(img2 + epsilon) - torch.log(img1 + epsilon)) / C


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def event_loss(image, image_now, image_next, gt_now, gt_next, gt_intensity, c, gt_blur=None, dead_ssim=False):
    """train.py:165-203. lambda_dssim is forced to 0 at :177, so the ssim_gray term of :178 contributes
    nothing; dead_ssim=True still EVALUATES it (and lets autograd walk it) exactly as the reference does --
    used where the reference's own cost is timed (bench.py --impl reference)."""
    img_diff = differentialable_event_simu(image_now, image_next, False, c)
    gt_image = differentialable_event_simu(gt_now, gt_next, False, 0.17)
    loss1 = l1_loss(img_diff, gt_image)
    if dead_ssim:
        lambda_dssim = 0
        loss1 = (1.0 - lambda_dssim) * loss1 + lambda_dssim * (1.0 - ssim(img_diff, gt_image))
    loss2 = l1_loss(image, gt_intensity)
    Event_weight = 0.9
    mask = (gt_image != 0).float()
    loss = Event_weight * (loss1 * mask).sum() + (1 - Event_weight) * (loss2 * (1 - mask)).sum()
    loss = loss / (mask.sum() + (1 - mask).sum())
    if gt_blur is not None:
        blur_alpha = 0.5
        loss = (1.0 - blur_alpha) * loss + blur_alpha * l1_loss(image, gt_blur)
    return loss


def event_iteration(pc, cam_train, cam_now, cam_next, gt, c, bg, deblur=False, render_fn=render, pipe=PIPE, dead_ssim=False):
    """One --event training iteration up to (not including) backward: train.py:144-203.
    gt: dict(intensity, event_now, event_next[, blurry]) of [3,H,W] tensors. Returns (loss, first render pkg)."""
    pkg = render_fn(cam_train, pc, pipe, bg)
    image = pkg["render"]
    image_now = render_fn(cam_now, pc, pipe, bg)["render"]
    image_next = render_fn(cam_next, pc, pipe, bg)["render"]
    loss = event_loss(image, image_now, image_next, gt["event_now"], gt["event_next"], gt["intensity"], c,
                      gt["blurry"] if deblur else None, dead_ssim=dead_ssim)
    return loss, pkg


# ---- the non-event losses (utils/loss_utils.py:18-23, 40-48, 359-418), restated for tests/bench ----
def rgb_to_grayscale(image):
    return (0.299 * image[0, :, :] + 0.587 * image[1, :, :] + 0.114 * image[2, :, :]).unsqueeze(0)


def _window(channel, like):
    import math as _m
    g = torch.tensor([_m.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, 11, 11).contiguous().to(like)


def ssim(img1, img2):
    channel = img1.size(-3)
    w = _window(channel, img1)
    F = torch.nn.functional
    mu1, mu2 = F.conv2d(img1, w, padding=5, groups=channel), F.conv2d(img2, w, padding=5, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=channel) - mu1_sq
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=channel) - mu2_sq
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()


def l1_ssim_loss(image, gt, lambda_dssim=0.2, gray=False):
    """train.py:213-223 (gray) / :292-296 (RGB)."""
    if gray:
        a, b = rgb_to_grayscale(image), rgb_to_grayscale(gt)
    else:
        a, b = image, gt
    return (1.0 - lambda_dssim) * torch.abs(a - b).mean() + lambda_dssim * (1.0 - ssim(a, b))
