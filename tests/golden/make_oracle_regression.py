"""Regenerates tests/golden/oracle_regression.npz: outputs of the CPU oracle (oracle/liboracle.so) on small fixed-seed cases of
every integrator / spectrum kind.  NOT reference outputs -- the reference cannot run here -- but a
freeze of the oracle as it stood when the GPU path was last verified against it lane by lane, so that later edits of the
oracle (e.g. the refractive branches of volpathaos) cannot silently change the existing paths.
Run: `python tests/golden/make_oracle_regression.py` (only after a deliberate, reviewed change of the oracle's behaviour)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_cases  # noqa: E402


def main():
    out = oracle_cases.run_all()
    path = os.path.join(HERE, "oracle_regression.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%d arrays, %.1f KiB)" % (path, len(out), os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
