"""EXR writer / reader: round trips of the uncompressed writer (the role of HDRFilm::write for developed films) and the reader's
codecs (RLE, ZIPS, ZIP, PIZ) against the OpenEXR test images that ship with the reference."""
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


# ---------------------------------------------------------------------------------------------------------------
# Compressed files: the reference writes PIZ, so its film.write() outputs and the reference images
# of its test-suite need the codecs.  Ground truth: the pinned OpenEXR dependency of the reference (ext/openexr) ships
# ONE picture stored with every codec (IlmImfTest/comp_*.exr) -- the decoded compressed files must equal the
# uncompressed one bit for bit.  Those files live in the reference checkout (build container only).
# ---------------------------------------------------------------------------------------------------------------
import os
import zlib

IMF_TEST = "/root/reference/ext/openexr/OpenEXR/IlmImfTest"
REF_IMAGES = "/root/reference/resources/data/tests/scenes/participating_media/refs"
needs_reference = pytest.mark.skipif(not os.path.isdir(IMF_TEST), reason="reference checkout absent (build container only)")


@needs_reference
@pytest.mark.parametrize("codec", ["rle", "zips", "zip", "piz"])
def test_codecs_against_the_openexr_test_images(codec):
    want, names = read_exr(os.path.join(IMF_TEST, "comp_none.exr"))
    got, names2 = read_exr(os.path.join(IMF_TEST, "comp_%s.exr" % codec))
    assert names == names2 == ["A", "B", "G", "R"] and want.shape == (675, 587, 4)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))          # HALF data: bit-exact after widening


@needs_reference
@pytest.mark.parametrize("name", ["comp_bad_bounds_piz.exr", "comp_early_eof_piz.exr", "comp_short_decode_piz.exr",
                                  "comp_bad_pos_bounds_piz.exr", "comp_invalid_unknown.exr"])
def test_damaged_files_raise(name):
    """OpenEXR's own fuzz cases: no crash, no hang, a ValueError."""
    with pytest.raises(ValueError):
        read_exr(os.path.join(IMF_TEST, name))


@needs_reference
def test_reads_the_reference_images_of_the_participating_media_tests():
    """FLOAT channels, PIZ (with the writer's raw fallback for blocks that do not shrink).  The rgb and mono references of
    one scene are renders of the same radiance field (the mono file holds a colour-converted luminance, so the scale differs
    by some 10 %): the pictures are strongly correlated pixel by pixel, which a wrong word order or plane layout would destroy."""
    import glob
    files = sorted(glob.glob(os.path.join(REF_IMAGES, "*_ref_rgb.exr")))
    assert len(files) >= 10
    for f in files:
        rgb, names = read_exr(f)
        mono, _ = read_exr(f.replace("_ref_rgb", "_ref_mono"))
        assert names == ["B", "G", "R"] and rgb.shape == mono.shape == (48, 48, 3)
        assert np.isfinite(rgb).all() and (rgb >= 0).all() and rgb.mean() > 0
        lum = 0.212671 * rgb[..., 2] + 0.715160 * rgb[..., 1] + 0.072169 * rgb[..., 0]
        assert np.corrcoef(lum.ravel(), mono[..., 1].ravel())[0, 1] > 0.9, f
        assert 0.7 < mono[..., 1].mean() / lum.mean() < 1.1, f
    # a FLOAT file that PIZ really shrinks (16 KiB for 128 KiB of pixels): values come back in the picture's range
    img, names = read_exr("/root/reference/tutorials/scenes/references/starmoon.exr")
    assert names == ["Y"] and img.shape == (128, 256, 1) and 0.2 < img.min() <= img.max() < 0.35


def _write_zip_exr(path, img, lines):
    """Test helper: ZIPS (lines = 1) / ZIP (lines = 16) writer -- predictor, byte de-interleave, deflate."""
    h, w, c = img.shape
    names = ["B", "G", "R"][:c]
    chlist = b"".join(n.encode() + b"\0" + struct.pack("<iB3xii", 2, 0, 1, 1) for n in names) + b"\0"
    box = struct.pack("<4i", 0, 0, w - 1, h - 1)

    def attr(n, t, d):
        return n.encode() + b"\0" + t.encode() + b"\0" + struct.pack("<i", len(d)) + d
    header = (struct.pack("<ii", 20000630, 2) + attr("channels", "chlist", chlist) +
              attr("compression", "compression", bytes([2 if lines == 1 else 3])) + attr("dataWindow", "box2i", box) +
              attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0") +
              attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + attr("screenWindowCenter", "v2f", struct.pack("<2f", 0, 0)) +
              attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0")
    chunks = []
    for y0 in range(0, h, lines):
        raw = np.ascontiguousarray(img[y0:y0 + lines].transpose(0, 2, 1)).astype("<f4").tobytes()
        a = np.frombuffer(raw, np.uint8)
        t = np.concatenate([a[0::2], a[1::2]]).astype(np.int64)
        t[1:] = (t[1:] - t[:-1] + 128 + 256) & 0xff
        z = zlib.compress(t.astype(np.uint8).tobytes())
        data = z if len(z) < len(raw) else raw
        chunks.append(struct.pack("<ii", y0, len(data)) + data)
    pos = len(header) + 8 * len(chunks)
    with open(path, "wb") as f:
        f.write(header)
        for ch in chunks:
            f.write(struct.pack("<Q", pos))
            pos += len(ch)
        for ch in chunks:
            f.write(ch)


@pytest.mark.parametrize("lines", [1, 16])
def test_zip_round_trip_without_the_reference(tmp_path, lines):
    """Travels to machines without the reference checkout: smooth data (deflate shrinks it) and noise (raw fallback)."""
    rng = np.random.default_rng(5)
    y, x = np.mgrid[0:37, 0:29]
    smooth = np.stack([np.sin(x / 9.0) + y / 40.0, np.floor(x / 8.0), np.zeros_like(x, dtype=float)], axis=-1).astype(np.float32)
    noise = rng.uniform(0, 1, (37, 29, 3)).astype(np.float32)
    for k, img in enumerate((smooth, noise)):
        p = str(tmp_path / ("z%d.exr" % k))
        _write_zip_exr(p, img, lines)
        got, names = read_exr(p)
        assert names == ["B", "G", "R"] and np.array_equal(got, img)


@pytest.mark.parametrize("compression", ["none", "zips", "zip"])
def test_writer_compression_round_trip(tmp_path, compression):
    """write_exr(compression=...) against read_exr, whose ZIP path is pinned on OpenEXR's comp_zip.exr / comp_zips.exr; the
    independent test writer above produces the same chunk bytes for smooth data."""
    rng = np.random.default_rng(9)
    y, x = np.mgrid[0:37, 0:29]
    img = np.stack([np.sin(x / 9.0) + y / 40.0, np.floor(x / 8.0), rng.uniform(0, 1, x.shape)], axis=-1).astype(np.float32)
    p = str(tmp_path / "w.exr")
    write_exr(p, img, compression=compression)
    got, names = read_exr(p)
    assert names == ["B", "G", "R"] and np.array_equal(got[..., ::-1], img)
    if compression != "none":
        raw = str(tmp_path / "raw.exr")
        write_exr(raw, img)
        assert os.path.getsize(p) < os.path.getsize(raw)
    with pytest.raises(ValueError):
        write_exr(p, img, compression="dwaa")


def test_piz_writer_round_trip(tmp_path):
    rng = np.random.default_rng(11)
    y, x = np.mgrid[0:70, 0:45]
    img = np.stack([np.sin(x / 7.0) + y / 30.0, np.floor(x / 8.0), np.zeros_like(x, dtype=float), rng.uniform(0, 1, x.shape)],
                   axis=-1).astype(np.float32)
    for k, im in enumerate((img, np.full((33, 17, 1), 0.25, np.float32), img[:1, :1, :1], img[:, :, :3])):
        p = str(tmp_path / ("p%d.exr" % k))
        write_exr(p, im, compression="piz")
        got, names = read_exr(p)
        order = [names.index(n) for n in {1: "Y", 3: "RGB", 4: "IQUV"}[im.shape[2]]]
        assert np.array_equal(got[..., order], im)
    assert os.path.getsize(str(tmp_path / "p1.exr")) < 33 * 17 * 4 // 2          # a constant image shrinks a lot


@needs_reference
def test_piz_forward_transform_reproduces_the_coefficients_of_an_openexr_file():
    """The part of the PIZ encoder in front of the entropy coder is deterministic: for the first chunk of OpenEXR's own
    comp_piz.exr, the bitmap header and the LUT-mapped wavelet coefficients computed from the decoded pixels equal the ones
    stored in the file (the Huffman code itself depends on tie-breaking and is checked by the round trip)."""
    from mitsuba3_nasa_b200 import exr as E
    path = os.path.join(IMF_TEST, "comp_piz.exr")
    buf = open(path, "rb").read()
    pos = 8
    while buf[pos] != 0:
        _, pos = E._cstr(buf, pos)
        _, pos = E._cstr(buf, pos)
        (size,) = struct.unpack_from("<i", buf, pos)
        pos += 4 + size
    pos += 1
    (o,) = struct.unpack_from("<Q", buf, pos)
    y0, size = struct.unpack_from("<ii", buf, o)
    data = buf[o + 8:o + 8 + size]
    nx, ny, words_per = 587, 32, [1, 1, 1, 1]                                       # four HALF channels, first 32 scanlines
    raw = E._piz_decode(data, nx, ny, words_per)
    head, words = E._piz_forward(np.frombuffer(raw, "<u2").reshape(ny, nx * 4), nx, ny, words_per)
    min_nz, max_nz = struct.unpack_from("<HH", data, 0)
    hl = 4 + (max_nz - min_nz + 1)
    assert head == data[:hl]
    (length,) = struct.unpack_from("<i", data, hl)
    stored = E._huf_uncompress(data[hl + 4:hl + 4 + length], nx * ny * 4)
    assert np.array_equal(words, stored)


@pytest.mark.parametrize("compression", ["none", "zip", "piz"])
def test_half_precision_output(tmp_path, compression):
    """component_format='float16' (hdrfilm's default): HALF channels; values come back as the nearest half floats."""
    rng = np.random.default_rng(13)
    img = rng.uniform(0, 4, (40, 31, 3)).astype(np.float32)
    img[5:30, 3:20] = 0.5
    p = str(tmp_path / "h.exr")
    write_exr(p, img, compression=compression, component_format="float16")
    got, names = read_exr(p)
    assert names == ["B", "G", "R"] and np.array_equal(got[..., ::-1], img.astype(np.float16).astype(np.float32))
    f = str(tmp_path / "f.exr")
    write_exr(f, img, compression=compression)
    assert os.path.getsize(p) < 0.62 * os.path.getsize(f)
    with pytest.raises(ValueError):
        write_exr(p, img, component_format="uint32")
