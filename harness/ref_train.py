"""The reference's training iteration around the UNMODIFIED reference CUDA rasterizer (oracle/_ref) — TEST INFRASTRUCTURE:
the checker of tests/test_gpu_train_step.py and the `--impl reference` arm of bench.py.  Nothing under ms-gs_b200/ imports it.

/root/reference does not travel to the GPU box, so the few torch lines of the reference's Python that surround the
rasterizer are restated here, each citing the lines it follows:

  activations      scene/gaussian_model.py:127-155 (exp, sigmoid, F.normalize, cat) as gaussian_renderer/__init__.py:57-89 calls them
  rasterizer       oracle/_ref/libref_rasterizer.so = cuda_rasterizer/*.cu unmodified, through an autograd Function shaped like
                   diff_gaussian_rasterization/__init__.py:56-195 (uses only grad_out_color; same 11 torch.zeros gradient tensors)
  loss             train.py:208-215 with utils/loss_utils.py:17-64 (l1 + SSIM: five grouped 11x11 conv2d, zero padding 5)
  statistics       train.py:239-250 with scene/gaussian_model.py:662-705 (boolean-mask indexing, as the reference writes it)
  optimiser        scene/gaussian_model.py:229-252: torch.optim.Adam, one param group per tensor, eps 1e-15; train.py:416-418
"""
from math import exp

import torch
import torch.nn.functional as F

from harness import refcuda_bind

LR_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


class _RefRaster(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, shs, opacities, scales, rotations, t, cam, fs, fl, sync):
        tt = dict(t)
        tt.update(means3D=means3D.contiguous(), shs=shs.contiguous(), opacities=opacities.contiguous(),
                  scales=scales.contiguous(), rotations=rotations.contiguous())
        r = refcuda_bind.RefRasterizer(sync=sync).forward(tt, cam, fs, fl)
        ctx.r = r
        ctx.mark_non_differentiable(r.radii, r.pixel_sizes)
        return r.color, r.radii, r.pixel_sizes

    @staticmethod
    def backward(ctx, dcolor, _a, _b):
        g = ctx.r.backward(dcolor)
        ctx.r = None
        return g["mean3D"], g["mean2D"], g["sh"], g["opacity"], g["scale"], g["rot"], None, None, None, None, None


def _window(channel, device):
    # utils/loss_utils.py:24-33
    g = torch.Tensor([exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, 11, 11).contiguous().to(device)


def ssim(img1, img2):
    """utils/loss_utils.py:35-64 (size_average=True)"""
    ch = img1.size(-3)
    w = _window(ch, img1.device)
    mu1 = F.conv2d(img1, w, padding=5, groups=ch)
    mu2 = F.conv2d(img2, w, padding=5, groups=ch)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1_sq
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2_sq
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()


def training_loss(image, gt, lambda_dssim):
    """train.py:209-210"""
    return (1.0 - lambda_dssim) * torch.abs(image - gt).mean() + lambda_dssim * (1.0 - ssim(image, gt))


class RefTrainer:
    """params: {"xyz" (P,3), "f_dc" (P,1,3), "f_rest" (P,15,3), "opacity" (P,1), "scaling" (P,3), "rotation" (P,4)} raw
    leaf tensors (cloned here); state: an msgs_train.MSState-like object (fields are cloned and updated in place)."""

    def __init__(self, params, state, bg, sh_degree=3, lrs=None, lambda_dssim=0.2, filter_small=True, filter_large=True,
                 stats_what=3, base_mask_level=None, sync=False):
        self.p = {k: params[k].detach().clone().requires_grad_(True) for k in LR_NAMES}
        self.opt = torch.optim.Adam([{"params": [self.p[k]], "lr": lrs[k], "name": k} for k in LR_NAMES], lr=0.0, eps=1e-15)
        self.bg, self.sh_degree, self.lam = bg, sh_degree, lambda_dssim
        self.fs, self.fl, self.what, self.base_level, self.sync = filter_small, filter_large, stats_what, base_mask_level, sync
        self.L = state.reso_lvls
        self.target = state.target_reso_lvl.clone()
        self.max_px, self.min_px = state.max_pixel_sizes.clone(), state.min_pixel_sizes.clone()
        self.base = state.base_gaussian_mask.clone()
        self.max_radii2D = state.max_radii2D.clone()
        self.accum, self.denom = state.xyz_gradient_accum.clone(), state.denom.clone()
        self.occ, self.dc = state.occ_multiplier, state.dc_delta

    # scene/gaussian_model.py:662-686
    def _update_pixel_sizes(self, vis, px, lvl):
        mask = torch.logical_and(vis, self.target == lvl)
        if lvl > 0:
            self.max_px[mask] = self.max_px[mask] * 0.95
            self.max_px[mask] = torch.max(self.max_px[mask], px[mask])
        if lvl < self.L - 1:
            self.min_px[mask] = torch.clip(self.min_px[mask] * 1.05, -1)
            self.min_px[mask] = torch.where(
                self.min_px[mask] < 0,
                torch.where(px[mask] > 0, px[mask], self.min_px[mask]),
                torch.where(px[mask] > 0, torch.min(self.min_px[mask], px[mask]), self.min_px[mask]))

    def render(self, v, H, W):
        """gaussian_renderer/__init__.py:18-128"""
        p = self.p
        cam = _Cam(H, W, v.tanfovx, v.tanfovy)
        m2 = torch.zeros_like(p["xyz"], requires_grad=True)
        t = dict(max_pixel_sizes=self.max_px, min_pixel_sizes=self.min_px, base_mask=self.base, occ_multiplier=self.occ,
                 dc_delta=self.dc, bg=self.bg, viewmatrix=v.viewmatrix, projmatrix=v.projmatrix, campos=v.campos,
                 sh_degree=self.sh_degree, scale_modifier=1.0)
        shs = torch.cat((p["f_dc"], p["f_rest"]), dim=1)
        color, radii, px = _RefRaster.apply(p["xyz"], m2, shs, torch.sigmoid(p["opacity"]), torch.exp(p["scaling"]),
                                            F.normalize(p["rotation"]), t, cam, self.fs, self.fl, self.sync)
        return color, radii, px, m2

    def _statistics(self, radii, px, lvl, m2_grad):
        """train.py:239-250 for one rendered unit"""
        vis = radii > 0
        if (self.what & 4) and (self.base_level is None or lvl == self.base_level):
            self.base = torch.logical_or(self.base, vis)                            # gaussian_model.py:703-705
        if self.what & 1:
            self._update_pixel_sizes(vis, px, lvl)                                  # train.py:245
        if self.what & 2:
            self.max_radii2D[vis] = torch.max(self.max_radii2D[vis], radii[vis])    # train.py:249
            self.accum[:, lvl][vis] += torch.norm(m2_grad[vis, :2], dim=-1, keepdim=True)    # :698-701
            self.denom[:, lvl][vis] += 1

    def step(self, views, optimizer=True, defer_statistics=True):
        """Every (view, scale) unit of the step, then ONE optimizer step.  defer_statistics (the batched step's semantics:
        every unit of a step is rendered from the SAME model state): the per-unit statistics updates run, in unit order,
        after the last unit's backward; False interleaves them with the renders exactly as consecutive train.py
        iterations without an optimizer step in between would."""
        total = 0.0
        pending = []
        for v in views:
            for (H, W), gt, lvl in zip(v.sizes, v.gts, v.levels):
                image, radii, px, m2 = self.render(v, H, W)
                loss = training_loss(image, gt, self.lam) * (1 if lvl == 0 else 0.1)          # train.py:208-214
                loss.backward()
                total = total + loss.detach()
                with torch.no_grad():
                    if defer_statistics:
                        pending.append((radii, px, lvl, m2.grad))
                    else:
                        self._statistics(radii, px, lvl, m2.grad)
        with torch.no_grad():
            for a in pending:
                self._statistics(*a)
        if optimizer:
            self.opt.step()                                                                     # train.py:416-418
            self.opt.zero_grad(set_to_none=True)
        return total


class _Cam:
    def __init__(self, H, W, tx, ty):
        self.H, self.W, self.tanfovx, self.tanfovy = int(H), int(W), float(tx), float(ty)
