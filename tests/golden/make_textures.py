"""Writes tests/golden/textures/sky.exr: a small (48 x 24) procedural environment map -- a blue-to-white gradient, a warm
ground half and a bright sun lobe -- for the envmap tests (the reference reads the same file through its Bitmap class).

    python tests/golden/make_textures.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mitsuba3_nasa_b200 import exr  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "textures")


def sky(w=48, h=24):
    v, u = np.mgrid[0:h, 0:w]
    theta = (v + 0.5) / h * np.pi
    phi = (u + 0.5) / w * 2 * np.pi
    up = np.cos(theta)
    img = np.zeros((h, w, 3), np.float32)
    t = np.clip(up, 0, 1)
    img[..., 0] = 0.9 - 0.6 * t
    img[..., 1] = 0.9 - 0.4 * t
    img[..., 2] = 1.0 - 0.1 * t
    ground = up < 0
    img[ground] = np.array([0.35, 0.3, 0.25], np.float32) * (1.0 + 0.5 * up[ground])[:, None]
    d = np.stack([np.sin(theta) * np.sin(phi), np.cos(theta), -np.sin(theta) * np.cos(phi)], axis=-1)
    sun = np.array([0.4, 0.6, -0.69]); sun /= np.linalg.norm(sun)
    lobe = np.exp(40.0 * ((d * sun).sum(axis=-1) - 1.0))
    img += (lobe[..., None] * np.array([30.0, 26.0, 20.0])).astype(np.float32)
    return img.astype(np.float32)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    exr.write_exr(os.path.join(OUT, "sky.exr"), sky(), compression="zip")
    print(os.path.getsize(os.path.join(OUT, "sky.exr")), "bytes")
