"""Synthetic EM-like uint8 volumes (SURVEY.md 8d recipe): white noise -> Gaussian blur (sigma 2, reflect)
-> + 0.25 x white noise -> mean 128 / std 40 -> clip -> uint8.  Seeds: 42 (domain A), 43 (domain B), 44+."""
import numpy as np


def em_like_u8(shape, seed=42, sigma=2.0):
    from scipy.ndimage import gaussian_filter

    rng = np.random.default_rng(seed)
    v = rng.standard_normal(shape, dtype=np.float32)
    v = gaussian_filter(v, sigma=sigma, mode="reflect")
    v = v / (v.std() + 1e-12)
    v = v + 0.25 * rng.standard_normal(shape, dtype=np.float32)
    v = (v - v.mean()) / (v.std() + 1e-12)
    return np.clip(v * 40.0 + 128.0, 0, 255).astype(np.uint8)
