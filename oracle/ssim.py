"""ORACLE (test infrastructure, never imported by the product): float64 numpy restatement of the
reference's SSIM (src/entities/losses.py:28-113: 11-tap Gaussian window sigma 1.5 built in float32,
zero padding, C1 = 0.01^2, C2 = 0.03^2, mean over all elements) and of its gradient with respect to
img1.  PINNED: tests/golden/ssim_reference.npz holds values and autograd gradients produced by the
reference's own ``ssim`` (imported from /root/reference by tests/golden/make_golden_ssim.py);
tests/test_oracle_ssim.py checks this restatement against them."""
import numpy as np


def window():
    g = np.array([np.exp(-((x - 5) ** 2) / (2 * 1.5 ** 2)) for x in range(11)]).astype(np.float32)
    return (g / g.sum(dtype=np.float32)).astype(np.float64)


def _conv(img, g):
    """separable 'same' convolution with zero padding over the last two axes; img (C,H,W) float64"""
    C, H, W = img.shape
    p = np.pad(img, ((0, 0), (5, 5), (5, 5)))
    h = sum(g[k] * p[:, :, k:k + W] for k in range(11))
    return sum(g[k] * h[:, k:k + H, :] for k in range(11))


def ssim_and_grad(img1, img2):
    """-> (mean SSIM, d mean / d img1), both float64"""
    x, y = np.asarray(img1, np.float64), np.asarray(img2, np.float64)
    g = window()
    mu1, mu2 = _conv(x, g), _conv(y, g)
    s11, s22, s12 = _conv(x * x, g), _conv(y * y, g), _conv(x * y, g)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * (s12 - mu1 * mu2) + C2
    B1, B2 = mu1 ** 2 + mu2 ** 2 + C1, (s11 - mu1 ** 2) + (s22 - mu2 ** 2) + C2
    S = A1 * A2 / (B1 * B2)
    a = 2 * mu2 * (A2 - A1) / (B1 * B2) - S * 2 * mu1 * (1 / B1 - 1 / B2)
    b = -S / B2
    c = 2 * A1 / (B1 * B2)
    grad = (_conv(a, g) + 2 * x * _conv(b, g) + y * _conv(c, g)) / S.size
    return S.mean(), grad
