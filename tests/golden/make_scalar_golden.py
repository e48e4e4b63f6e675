"""Generates tests/golden/scalar_ref.npz: outputs of the UNMODIFIED reference integrators (patched scalar build of
/root/reference, oracle/build_ref_scalar.sh -> oracle/_ref/scalar/ref_driver) for the scenes of tests/scalar_cases.py.

    python tests/golden/make_scalar_golden.py            (needs oracle/_ref/scalar; runs in THIS container only)

For 'samples' cases the file holds, per case, float32 [pixel][sample][4]: what SamplingIntegrator::render_sample wrote
into `aovs` for every sample, pixels in the order and with the seeds of the reference's scalar render loop.  As a check of
the driver's own loop the reference's Integrator::render (one thread) is run on the same scene and must equal the mean of
the dumped samples (rgb variants; in mono variants the fork's film does not hold what render_sample produces).
For 'render1' cases: the float32 [16] / [1] result of SamplingIntegrator::render_1 (one thread)."""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import mitsuba3_nasa_b200 as mi  # noqa: E402
import scalar_cases  # noqa: E402

DRIVER = os.path.join(ROOT, "oracle", "_ref", "scalar", "ref_driver")


def run_case(name, case, tmp):
    d = case['scene']
    case_dir = os.path.join(tmp, name)
    os.makedirs(case_dir)
    xml_path = os.path.join(case_dir, "scene.xml")
    with open(xml_path, "w") as f:
        f.write(mi.dict_to_xml(d, case_dir))          # in-memory grids become .vol files next to the scene
    film = d['sensor']['film']
    W, H = film['width'], film['height']
    out_bin = os.path.join(tmp, name + ".bin")
    if case['kind'] == 'render1':
        subprocess.check_call([DRIVER, case['variant'], xml_path, 'render1', str(case['seed']), str(case['spp']), out_bin])
        return np.fromfile(out_bin, np.float32)
    pix, seeds = scalar_cases.pixel_seeds(W, H, case['seed'])
    px_path = os.path.join(tmp, name + ".px")
    with open(px_path, "w") as f:
        for p, s in zip(pix, seeds):
            f.write("%d %d %d\n" % (p % W, p // W, s))
    subprocess.check_call([DRIVER, case['variant'], xml_path, 'samples', px_path, str(case['spp']), out_bin])
    samples = np.fromfile(out_bin, np.float32).reshape(W * H, case['spp'], 4)
    if case['variant'] == 'scalar_rgb' and film.get('rfilter', {}).get('type', 'gaussian') == 'box':
        # the reference's own render loop on the same scene and seed
        subprocess.check_call([DRIVER, case['variant'], xml_path, 'render', str(case['seed']), str(case['spp']), out_bin])
        img = np.fromfile(out_bin, np.float32).reshape(H, W, -1)[..., :3]
        mean = np.zeros((H * W, 3))
        mean[pix] = samples[..., :3].astype(np.float64).mean(axis=1)
        err = np.abs(mean.reshape(H, W, 3) - img).max() / max(img.max(), 1e-9)
        assert err < 1e-5, (name, "driver loop differs from Integrator::render", err)
        print("   render() of the reference agrees with the mean of the dumped samples: rel %.2g" % err)
    return samples


def run_xml_case(name, case, tmp):
    """Samples of a scene FILE (the fork's own render-test scenes): 48 x 48 film, `spp` samples per pixel."""
    W = H = 48
    pix, seeds = scalar_cases.pixel_seeds(W, H, case['seed'])
    px_path = os.path.join(tmp, name + ".px")
    with open(px_path, "w") as f:
        for p, s in zip(pix, seeds):
            f.write("%d %d %d\n" % (p % W, p // W, s))
    out_bin = os.path.join(tmp, name + ".bin")
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(DRIVER))
    subprocess.check_call([DRIVER, case['variant'], case['file'], 'samples', px_path, str(case['spp']), out_bin], env=env)
    return np.fromfile(out_bin, np.float32).reshape(W * H, case['spp'], 4)


def main():
    if not os.path.exists(DRIVER):
        sys.exit("oracle/_ref/scalar/ref_driver is missing: run oracle/build_ref_scalar.sh first")
    out = {}
    only = set(sys.argv[1:])                       # names given: regenerate those, keep the rest of the file
    npz = os.path.join(HERE, "scalar_ref.npz")
    if only:
        out = dict(np.load(npz))
    with tempfile.TemporaryDirectory() as tmp:
        for name, case in scalar_cases.cases().items():
            if only and name not in only:
                continue
            print(name, case['variant'], case['kind'])
            out[name] = run_case(name, case, tmp)
        for name, case in scalar_cases.xml_cases().items():
            if only and name not in only:
                continue
            print(name, case['variant'], case['file'])
            out[name] = run_xml_case(name, case, tmp)
    np.savez_compressed(os.path.join(HERE, "scalar_ref.npz"), **out)
    print("wrote", os.path.join(HERE, "scalar_ref.npz"), os.path.getsize(os.path.join(HERE, "scalar_ref.npz")), "bytes")


if __name__ == "__main__":
    main()
