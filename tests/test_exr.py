"""EXR writer / reader round trips (uncompressed float32 scanline files; the role of HDRFilm::write for developed films)."""
import struct

import numpy as np
import pytest

from mitsuba3_nasa_b200.exr import write_exr, read_exr


@pytest.mark.parametrize("shape", [(5, 7, 3), (4, 4, 1), (3, 9, 4), (6, 2)])
def test_round_trip(tmp_path, shape):
    img = np.random.default_rng(0).normal(size=shape).astype(np.float32)
    p = tmp_path / "a.exr"
    write_exr(str(p), img)
    got, names = read_exr(str(p))
    c = 1 if len(shape) == 2 else shape[2]
    want = {1: ["Y"], 3: ["B", "G", "R"], 4: ["I", "Q", "U", "V"]}[c]      # file order is alphabetical
    assert names == want
    src = img.reshape(shape[0], shape[1], c)
    order = {1: [0], 3: [2, 1, 0], 4: [0, 1, 2, 3]}[c]
    assert np.array_equal(got, src[:, :, order])


def test_file_layout(tmp_path):
    """Header fields a third-party reader relies on: magic, version, line-offset table pointing at (y, size) chunk heads."""
    img = np.arange(2 * 3 * 3, dtype=np.float32).reshape(2, 3, 3)
    p = tmp_path / "b.exr"
    write_exr(str(p), img)
    buf = p.read_bytes()
    assert buf[:4] == bytes([0x76, 0x2f, 0x31, 0x01]) and struct.unpack_from("<i", buf, 4)[0] == 2
    end = buf.index(b"screenWindowWidth\0float\0") + len(b"screenWindowWidth\0float\0") + 4 + 4 + 1
    offs = struct.unpack_from("<2Q", buf, end)
    for y, o in enumerate(offs):
        assert struct.unpack_from("<ii", buf, o) == (y, 4 * 3 * 3)
    assert len(buf) == offs[1] + 8 + 36


def test_errors(tmp_path):
    with pytest.raises(ValueError):
        write_exr(str(tmp_path / "c.exr"), np.zeros((2, 2, 2), np.float32), channel_names=["A", "A"])
    with pytest.raises(ValueError):
        write_exr(str(tmp_path / "d.exr"), np.zeros(4, np.float32))
