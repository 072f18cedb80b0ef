"""kornia.filters stand-in (test infrastructure).

spatial_gradient: un-normalised 3x3 Sobel with replicate padding, output (B, C, 2, H, W) with
index 0 = d/dx, 1 = d/dy.  This is the variant that reproduces the reference notebook's stored
known answer (27497.41105769231 -> 276.125) -- see SURVEY.md section 4; it is the same operator
as the reference's vendored featurePnP/helpers/sobel_pytorch.py:9-59 (which zero-pads; the toy
scene does not distinguish the two paddings).
"""
import torch
import torch.nn.functional as F


def spatial_gradient(x, mode="sobel", order=1, normalized=False):
    assert x.dim() == 4
    b, c, h, w = x.shape
    kx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]]).to(x)
    ky = kx.t().contiguous()
    if normalized:
        kx = kx / 8.0
        ky = ky / 8.0
    k = torch.stack([kx, ky])[:, None]  # 2,1,3,3
    xp = F.pad(x.reshape(b * c, 1, h, w), (1, 1, 1, 1), mode="replicate")
    g = F.conv2d(xp, k)  # b*c, 2, h, w
    return g.reshape(b, c, 2, h, w)


def get_gaussian_kernel2d(kernel_size, sigma):
    def g1(n, s):
        x = torch.arange(n, dtype=torch.float32) - n // 2
        if n % 2 == 0:
            x = x + 0.5
        k = torch.exp(-x ** 2 / (2.0 * s * s))
        return k / k.sum()

    ky = g1(kernel_size[0], sigma[0])
    kx = g1(kernel_size[1], sigma[1])
    return torch.outer(ky, kx)
