"""Environment-map emitter, host side (src/emitters/envmap.cpp:108-232): the texel array with its repeated last column and the
luminance warp (Hierarchical2D<Float, 0>, include/mitsuba/core/distr_2d.h:344-480), packed into one float buffer in the
layout `mib_emitter` documents (mi_b200.h).

    build(rgb [H, W0, 3] float32, mono) -> (buffer float32, W, H, level_offsets, level_widths)"""
import numpy as np

from . import _abi as A


def luminance(rgb):
    # include/mitsuba/core/spectrum.h:431-434, float32 left to right
    r, g, b = (rgb[..., i].astype(np.float32) for i in range(3))
    return (r * np.float32(0.212671) + g * np.float32(0.715160) + b * np.float32(0.072169)).astype(np.float32)


def level_index(x, y, width):
    """Hierarchical2D::Level::index (distr_2d.h:766-770): 2 x 2 patches are contiguous."""
    return ((x & 1) | (((x & ~1) | (y & 1)) << 1)) + ((y & ~1) * width)


def build_levels(lum):
    """lum: [H, W] float32 (luminance * sin theta, last column repeated).  -> list of (width, height, data float32 flat)."""
    H, W = lum.shape
    npx, npy = W - 1, H - 1                                  # bilinear patches
    max_level = int(np.ceil(np.log2(max(npx, npy))))         # math::log2i_ceil
    # level 1: patch averages in the blocked layout, zero padded to even sizes
    sizes = []
    lx, ly = npx, npy
    for _ in range(max_level, -1, -1):
        lx += lx & 1
        ly += ly & 1
        sizes.append((lx, ly))
        lx >>= 1
        ly >>= 1
    avg = (np.float32(0.25) * (((lum[:-1, :-1] + lum[:-1, 1:]) + lum[1:, :-1]) + lum[1:, 1:])).astype(np.float32)
    total = float(avg.astype(np.float64).sum())
    scale = np.float32(float(npx * npy) / total)
    levels = [(W, H, (lum.reshape(-1) * scale).astype(np.float32))]
    w1, h1 = sizes[0]
    l1 = np.zeros(w1 * h1, np.float32)
    ys, xs = np.mgrid[0:npy, 0:npx]
    l1[level_index(xs, ys, w1)] = avg
    l1 *= scale
    levels.append((w1, h1, l1))
    cx, cy = npx, npy
    for k in range(1, len(sizes)):
        pw, ph, prev = levels[-1]
        w, h = sizes[k]
        cx, cy = (cx + 1) >> 1, (cy + 1) >> 1
        cur = np.zeros(w * h, np.float32)
        ys, xs = np.mgrid[0:cy, 0:cx]
        i0 = level_index(xs * 2, ys * 2, pw)
        cur[level_index(xs, ys, w)] = ((prev[i0] + prev[i0 + 1]) + prev[i0 + 2]) + prev[i0 + 3]
        levels.append((w, h, cur))
    return levels


def build(rgb, mono):
    rgb = np.ascontiguousarray(rgb, np.float32)
    H, W0 = rgb.shape[:2]
    if W0 < 2 or H < 3:
        raise ValueError("the environment map resolution must be at least 2x3 pixels")
    lum = luminance(rgb)
    texel = np.repeat(lum[..., None], 3, axis=2) if mono else rgb
    data = np.concatenate([texel, texel[:, :1]], axis=1).astype(np.float32)          # last column mirrors the first
    theta_scale = np.float32(1.0) / np.float32(H - 1) * np.float32(np.pi)
    sin_theta = np.sin((np.arange(H, dtype=np.float32) * theta_scale).astype(np.float32)).astype(np.float32)
    lw = (np.maximum(lum, 0) * sin_theta[:, None]).astype(np.float32)
    lw = np.concatenate([lw, lw[:, :1]], axis=1)
    levels = build_levels(lw)
    if len(levels) > A.MIB_MAX_ENV_LEVELS:
        raise ValueError("environment map too large (more than %d warp levels)" % A.MIB_MAX_ENV_LEVELS)
    parts, offs, widths = [data.reshape(-1)], [], []
    o = data.size
    for (w, h, d) in levels:
        offs.append(o)
        widths.append(w)
        parts.append(d)
        o += d.size
    return np.concatenate(parts).astype(np.float32), W0 + 1, H, offs, widths
