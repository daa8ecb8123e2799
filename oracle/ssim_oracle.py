"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, float64) of the reference's SSIM functions; never imported
by the product path.

  ssim / _ssim                     thirdparty/gaussian_splatting/utils/loss_utils.py:41-101
  compute_ssim_components / _ssim  src/utils/dyn_uncertainty/mapping_utils.py:99-204

Pinned against the reference itself: tests/golden/ssim_*.npz were produced by importing the reference's functions
(tests/golden/make_ssim_golden.py); tests/test_loss_oracle.py checks this restatement against them."""
import numpy as np
from scipy.signal import correlate2d

C1, C2 = 0.01 ** 2, 0.03 ** 2


def window_1d(window_size, sigma=1.5):
    # loss_utils.py:41-48: built as a float32 torch.Tensor and normalised in float32
    g = np.array([np.exp(-((x - window_size // 2) ** 2) / float(2 * sigma ** 2)) for x in range(window_size)], np.float32)
    return (g / g.sum(dtype=np.float32)).astype(np.float64)


def _conv(x, w2):
    """depthwise conv2d with zero padding window//2 (loss_utils.py:80-81) on (...,H,W)"""
    flat = x.reshape((-1,) + x.shape[-2:])
    return np.stack([correlate2d(p, w2, mode="same", boundary="fill", fillvalue=0.0) for p in flat]).reshape(x.shape)


def _moments(img1, img2, window_size):
    w = window_1d(window_size)
    w2 = np.outer(w, w)
    a, b = img1.astype(np.float64), img2.astype(np.float64)
    mu1, mu2 = _conv(a, w2), _conv(b, w2)
    return mu1, mu2, _conv(a * a, w2) - mu1 ** 2, _conv(b * b, w2) - mu2 ** 2, _conv(a * b, w2) - mu1 * mu2


def ssim_map(img1, img2, window_size=11):
    mu1, mu2, s1, s2, s12 = _moments(img1, img2, window_size)
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 ** 2 + mu2 ** 2 + C1) * (s1 + s2 + C2))


def ssim(img1, img2, window_size=11, size_average=True):
    m = ssim_map(img1, img2, window_size)
    return m.mean() if size_average else m.mean(axis=(1, 2, 3))


def ssim_grad(img1, img2, window_size=11):
    """d mean(ssim_map) / d img1 -- analytic, through the raw windowed moments (float64)"""
    w = window_1d(window_size)
    w2 = np.outer(w, w)
    a, b = img1.astype(np.float64), img2.astype(np.float64)
    mu1, mu2, s1, s2, s12 = _moments(img1, img2, window_size)
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    B1, B2 = mu1 ** 2 + mu2 ** 2 + C1, s1 + s2 + C2
    S = A1 * A2 / (B1 * B2)
    d_mu1 = 2 * mu2 * (A2 - A1) / (B1 * B2) + 2 * mu1 * S * (1 / B2 - 1 / B1)
    d_s11 = -S / B2
    d_s12 = 2 * A1 / (B1 * B2)
    g = _conv(d_mu1, w2) + 2 * a * _conv(d_s11, w2) + b * _conv(d_s12, w2)   # the window is symmetric
    return g / a.size


def components(img1, img2, window_size=11, eps=float(np.finfo(np.float32).eps)):
    """(luminance, contrast, structure), channel means; input (C,H,W)"""
    mu1, mu2, s1, s2, s12 = _moments(img1, img2, window_size)
    s1, s2 = np.maximum(eps, s1), np.maximum(eps, s2)
    s12 = np.sign(s12) * np.minimum(np.sqrt(s1 * s2), np.abs(s12))
    C3 = C2 / 2
    lum = (2 * mu1 * mu2 + C1) / (mu1 ** 2 + mu2 ** 2 + C1)
    con = np.minimum((2 * np.sqrt(s1) * np.sqrt(s2) + C2) / (s1 + s2 + C2), 0.98)
    stu = np.minimum((s12 + C3) / (np.sqrt(s1) * np.sqrt(s2) + C3), 0.98)
    return lum.mean(axis=-3), con.mean(axis=-3), stu.mean(axis=-3)
