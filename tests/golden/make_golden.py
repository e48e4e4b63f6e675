"""Regenerates tests/golden/ref_blocks.npz from the reference's own headers.

Run in the build container (needs /root/reference): `python tests/golden/make_golden.py`.
It builds oracle/_ref/libmiref.so (reference headers compiled in place by oracle/Makefile,
no reference source is copied) and stores the reference OUTPUTS of every building block of
tests/blocks.py; inputs are regenerated from fixed seeds by the tests.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import blocks  # noqa: E402
import orc  # noqa: E402

N = 1500


def main():
    ref = orc.ref_lib()
    if ref is None:
        raise SystemExit("oracle/_ref/libmiref.so is not available (needs /root/reference)")
    out = {}
    for name, (fn, _tol) in blocks.BLOCKS.items():
        for key, arr in fn(ref, "ref_", N).items():
            out["%s/%s" % (name, key)] = arr
    for name, fn in blocks.EXTRA_GOLDEN.items():
        for key, arr in fn(ref, "ref_", N).items():
            out["%s/%s" % (name, key)] = arr
    path = os.path.join(HERE, "ref_blocks.npz")
    np.savez_compressed(path, n=np.array([N]), **out)
    print("wrote %s (%d arrays, %.1f KiB)" % (path, len(out), os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
