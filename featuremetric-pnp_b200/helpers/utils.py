"""Drop-in for the reference's `featurePnP/helpers/utils.py` (same names, same call signatures).

The robust losses are gin-registered under the reference's names (helpers/utils.py:15-38 there),
so existing `.gin` files (`sparseFeaturePnP.loss_fn = @geman_mcclure_loss`) bind unchanged.  On
the hot path the loss is evaluated INSIDE the CUDA kernel: `sparseFeaturePnP` only maps the
function object to a loss id (fmpnp_b200.loss_id_of).  The torch bodies below exist because
callers may also invoke the losses directly on tensors; they are host-side API, not the product
path.
"""
import torch

try:  # gin-config is optional at import time; real gin is used when it is installed
    import gin
except ImportError:  # pragma: no cover - exercised in the build container
    from . import _gin_shim as gin

_EPS32 = torch.finfo(torch.float32).eps


def scaled_loss(x, fn, a):
    """loss(x / a^2) * a^2 with matching derivatives (reference utils.py:10-13)."""
    a2 = a ** 2
    rho, d1, d2 = fn(x / a2)
    return rho * a2, d1, d2 / a2


@gin.register
def squared_loss(x):
    return x, torch.ones_like(x), torch.zeros_like(x)


@gin.register
def huber_loss(x):
    inside = x <= 1
    root = torch.sqrt(x)
    inv_root = torch.clamp_min(1 / root, _EPS32)
    rho = torch.where(inside, x, 2 * root - 1)
    d1 = torch.where(inside, torch.ones_like(x), inv_root)
    d2 = torch.where(inside, torch.zeros_like(x), -inv_root / (2 * x))
    return rho, d1, d2


def barron_loss(x, alpha):
    """General robust loss of Barron (CVPR 2019) on an already squared, unit-scale input;
    alpha = 2 squared, alpha = 0 Cauchy, alpha = -2 Geman-McClure (reference utils.py:40-78)."""
    alpha = torch.as_tensor(alpha, dtype=x.dtype, device=x.device)
    eps = torch.tensor(_EPS32, dtype=x.dtype, device=x.device)
    beta = torch.maximum(eps, torch.abs(alpha - 2.0))
    sgn = torch.where(alpha >= 0, torch.ones_like(alpha), -torch.ones_like(alpha))
    a_safe = sgn * torch.maximum(eps, torch.abs(alpha))
    base = x / beta + 1.0
    general = 2 * (beta / a_safe) * (torch.pow(base, 0.5 * alpha) - 1.0)
    general_d1 = torch.pow(base, 0.5 * alpha - 1.0)
    cauchy = 2 * torch.log1p(torch.clamp_max(0.5 * x, 33e37))
    cauchy_d1 = 2 / (x + 2)
    rho = torch.where(alpha == 0, cauchy, torch.where(alpha == 2, x, general))
    d1 = torch.where(alpha == 0, cauchy_d1, torch.where(alpha == 2, torch.ones_like(x), general_d1))
    return rho, d1, torch.zeros_like(x)


@gin.register
def cauchy_loss(x):
    return barron_loss(x, 0.0)


@gin.register
def geman_mcclure_loss(x):
    return barron_loss(x, -2.0)


@gin.configurable
def sobel_normalization(normalized=False):
    """Whether the 3x3 Sobel kernels are scaled by 1/8.  PARITY UNPINNED: the reference calls
    kornia.filters.spatial_gradient with its defaults (utils.py:97,101) and pins kornia 0.2.2, which is neither
    vendored nor installable offline; newer kornia releases default to the normalised kernel, while the only
    in-repo evidence -- the toy notebook's stored cost trace and the vendored helpers/sobel_pytorch.py -- needs the
    un-normalised one.  So the choice is explicit and gin-bindable instead of silent:

        sobel_normalization.normalized = True      # in the .gin file, if the deployment's kornia normalises

    Every place that builds gradient maps (sobel_filter below, build_problem of the device-side adapter, the
    resampled pyramid levels of model.multilevel_optimization) asks this function."""
    return bool(normalized)


def sobel_filter(feature_map, batch=False, normalized=None):
    """3x3 Sobel d/dx, d/dy of a (C, H, W) map [(B, C, H, W) with batch=True]; returns
    (grad_x, grad_y) shaped like the input (reference utils.py:81-104, which delegates to
    kornia.filters.spatial_gradient).  Replicate padding; `normalized=None` takes sobel_normalization()
    (default: the un-normalised kernel, which reproduces the reference notebook's stored known answer,
    SURVEY.md section 4)."""
    if normalized is None:
        normalized = sobel_normalization()
    assert feature_map.dim() > 2
    x = feature_map if batch else feature_map[None]
    b, c, h, w = x.shape
    kx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]], dtype=x.dtype, device=x.device)
    if normalized:
        kx = kx / 8.0
    k = torch.stack([kx, kx.t()])[:, None]
    xp = torch.nn.functional.pad(x.reshape(b * c, 1, h, w), (1, 1, 1, 1), mode="replicate")
    g = torch.nn.functional.conv2d(xp, k).reshape(b, c, 2, h, w)
    gx, gy = g[:, :, 0], g[:, :, 1]
    if batch:
        return gx, gy
    return gx.reshape(feature_map.shape), gy.reshape(feature_map.shape)


def gaussian_kernel2d(kernel_size, sigma):
    """k x k Gaussian window as kornia 0.2.2's get_gaussian_kernel2d((k, k), (sigma, sigma)) builds it
    (float32: exp(-(x - k // 2)^2 / (2 sigma^2)) normalised to sum 1, outer product), returned in float64
    as the reference casts it (model.py:199-200).  kornia is not vendored by the reference: parity unpinned."""
    x = torch.arange(int(kernel_size), dtype=torch.float32) - int(kernel_size) // 2
    g = torch.exp(-(x * x) / float(2 * sigma ** 2))
    g = g / g.sum()
    return torch.outer(g, g).double()


def to_homogeneous(points):
    if isinstance(points, torch.Tensor):
        return torch.cat([points, points.new_ones(points.shape[:-1] + (1,))], dim=-1)
    import numpy as np
    if isinstance(points, np.ndarray):
        return np.concatenate([points, np.ones(points.shape[:-1] + (1,), dtype=points.dtype)], axis=-1)
    raise ValueError


def from_homogeneous(points):
    return points[..., :-1] / points[..., -1:]


def batched_eye_like(x, n):
    return torch.eye(n).to(x)[None].repeat(len(x), 1, 1)


def skew_symmetric(v):
    z = torch.zeros_like(v[..., 0])
    rows = torch.stack([z, -v[..., 2], v[..., 1], v[..., 2], z, -v[..., 0], -v[..., 1], v[..., 0], z], dim=-1)
    return rows.reshape(v.shape[:-1] + (3, 3))


def so3exp_map(w):
    theta = w.norm(p=2, dim=-1, keepdim=True)
    W = skew_symmetric(w / theta)
    theta = theta[..., None]
    res = W * torch.sin(theta) + (W @ W) * (1 - torch.cos(theta))
    res = torch.where(theta < 1e-12, torch.zeros_like(res), res)
    return torch.eye(3).to(W) + res
