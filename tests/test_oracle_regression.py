"""Freeze of the CPU oracle on small fixed-seed cases (tests/golden/oracle_regression.npz, made by
tests/golden/make_oracle_regression.py when the CUDA path was last verified against the oracle lane by lane).  The oracle is
what every GPU parity test compares with; this test makes an accidental change of its existing paths visible on CPU.
Bit-exact: the oracle is compiled with -ffp-contract=off / -fno-fast-math and the cases run on one thread."""
import os

import numpy as np

import oracle_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle_regression.npz")


def test_oracle_outputs_are_unchanged():
    want = np.load(GOLDEN)
    got = oracle_cases.run_all()
    assert sorted(got) == sorted(want.files)
    for k in want.files:
        a, b = np.asarray(got[k]), want[k]
        assert a.shape == b.shape and a.dtype == b.dtype, k
        assert np.array_equal(a, b, equal_nan=True), "%s: %d entries differ, max abs %.3g" % (
            k, int((a != b).sum()), float(np.nanmax(np.abs(a.astype(np.float64) - b))))
