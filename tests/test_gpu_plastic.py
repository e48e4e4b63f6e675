"""GPU parity of the `plastic` and `roughplastic` BSDFs (src/bsdfs/plastic.cpp, src/bsdfs/roughplastic.cpp; the last BSDFs the
fork's participating-media render tests use), walked by the one-thread-one-path kernels through the general BSDF interface of
csrc/dbsdf.cuh.  The oracle reproduces the reference's scalar build on these scenes sample by sample (tests/
test_oracle_vs_scalar_ref.py plastic_*, with the host's numpy tables of mitsuba3_nasa_b200/plastic.py pinned on the reference's
headers in tests/test_oracle_vs_ref.py); here the kernels must reproduce the oracle lane by lane under the llvm_ad_* seeding:
tolerance 1e-4 of the lane's largest component, at most 0.2 % of the lanes flipped."""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S
import scalar_cases
from test_gpu_mesh import _compare

pytestmark = pytest.mark.gpu
CASES = [n for n in scalar_cases.cases() if n.startswith('plastic_')]


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("variant", ["llvm_ad_rgb", "llvm_ad_mono", "llvm_ad_mono_polarized"])
def test_lane_by_lane(name, variant):
    d = scalar_cases.cases()[name]['scene']
    if variant == "llvm_ad_mono_polarized" and d['integrator']['type'] == 'volpathmis':
        d['integrator']['type'] = 'volpath'
    _compare(S.load_dict(d, variant), 5, 32, 10 * 10 * 32)


def test_plastic_under_volpathaos():
    """volpathaos takes such scenes through the interface kernel (docean.cuh); both lobes are reflections, so its refractive
    next-event estimation finds nothing to sample (a transmission-only context)."""
    d = scalar_cases.cases()['plastic_pol']['scene']
    d['integrator']['type'] = 'volpathaos'
    d['medium1']['phase'] = {'type': 'rayleigh'}
    _compare(S.load_dict(d, "llvm_ad_mono_polarized"), 2, 32, 10 * 10 * 32)


def test_prb_integrators_refuse_plastic():
    d = scalar_cases.cases()['plastic_rgb']['scene']
    d['integrator']['type'] = 'prbvolpath'
    from mitsuba3_nasa_b200 import _lib
    with pytest.raises(Exception, match="forward integrators only"):
        _lib.DeviceScene(S.load_dict(d, "llvm_ad_rgb"), 0)
