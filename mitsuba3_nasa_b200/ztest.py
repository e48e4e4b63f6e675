"""The statistical image comparison of the reference's render tests (src/render/tests/test_renders.py:159-229): a per-pixel
Z-test of a rendered mean image against a reference image and its per-sample variance image, with the Sidak correction for
the number of pixels and the reference's acceptance rule (99.75 % of the pixels must pass).  Together with `read_exr` (the
reference images are PIZ EXR files) this is the checking side of SURVEY 8f rank 3."""
import math

import numpy as np


def z_test(mean, sample_count, reference, reference_var):
    """test_renders.py:159-176: two-sided p-value per pixel and channel."""
    mean, reference = np.asarray(mean, np.float64), np.asarray(reference, np.float64)
    reference_var = np.maximum(np.asarray(reference_var, np.float64), 1e-4)          # "sanitize the variance images"
    z_stat = np.abs(mean - reference) * np.sqrt(sample_count / reference_var)
    erf = np.vectorize(math.erf, otypes=[np.float64])
    cdf = (1.0 - erf(-z_stat / math.sqrt(2.0))) * 0.5
    return 2.0 * (1.0 - cdf)


def sidak_alpha(significance_level, pixel_count):
    """test_renders.py:219-222."""
    return 1.0 - (1.0 - significance_level) ** (1.0 / pixel_count)


def images_agree(mean, sample_count, reference, reference_var, significance_level=0.01):
    """test_renders.py:202-229.  Returns (accepted, min p-value, alpha).  `mean` etc. are [H, W, C] arrays; the acceptance rule
    counts passing entries over all channels against 0.9975 x pixel count x channel count (the reference divides by its 3
    channels)."""
    mean = np.asarray(mean)
    if mean.ndim == 2:
        mean = mean[..., None]
    reference = np.asarray(reference).reshape(mean.shape)
    reference_var = np.asarray(reference_var).reshape(mean.shape)
    pixel_count = mean.shape[0] * mean.shape[1]
    p_value = z_test(mean, sample_count, reference, reference_var)
    alpha = sidak_alpha(significance_level, pixel_count)
    success = p_value > alpha
    accepted = (np.count_nonzero(success) / mean.shape[2]) >= 0.9975 * pixel_count
    return bool(accepted), float(p_value.min()), float(alpha)
