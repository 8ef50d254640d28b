"""Image-level statistics of SURVEY 8(d): relMSE of a render against a high-spp reference image.

    relMSE(I, R) = mean over pixels and channels of (I - R)^2 / (R^2 + 1e-2)

The acceptance rule for "GPU equals CPU at equal spp": with K independent seeds per implementation,
|mean relMSE_gpu - mean relMSE_cpu| <= 10 % of relMSE_cpu + 3 standard errors of the difference.
"""
import numpy as np


def resolve(film):
    """[H,W,4] {rgbSum, weightSum} -> [H,W,3] rgb (film.h:269-290 GetPixelRGB)."""
    w = film[..., 3:4]
    return np.where(w != 0, film[..., :3] / np.where(w != 0, w, 1), 0.0)


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def compare(rel_a, rel_b):
    """rel_a, rel_b: per-seed relMSE lists of two implementations.  Returns dict with the means, the standard error
    of their difference and whether they agree under the rule above."""
    a, b = np.asarray(rel_a, float), np.asarray(rel_b, float)
    se = float(np.sqrt(a.var(ddof=1) / a.size + b.var(ddof=1) / b.size)) if min(a.size, b.size) > 1 else float("inf")
    diff = float(abs(a.mean() - b.mean()))
    return {"a": float(a.mean()), "b": float(b.mean()), "diff": diff, "stderr": se, "tolerance": 0.10 * float(b.mean()) + 3 * se,
            "agree": bool(diff <= 0.10 * b.mean() + 3 * se)}
