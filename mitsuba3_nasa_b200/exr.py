"""Minimal OpenEXR scanline writer / reader for developed films (the role of `Bitmap::write` / `HDRFilm::write`,
src/films/hdrfilm.cpp:420-590, for the float32 case): uncompressed, one scanline per chunk, FLOAT channels.

Written from the OpenEXR file-layout specification (magic, version, attribute list, line-offset table, chunks); files
open in any EXR viewer.  The reader understands the files this writer produces (tests, round trips) -- the PIZ-compressed
reference images of the fork's test suite need a real decoder and are out of scope.
"""
import struct

import numpy as np

_MAGIC = 20000630
_FLOAT = 2


def _attr(name, typ, data):
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(data)) + data


def default_channel_names(n):
    return {1: ["Y"], 3: ["R", "G", "B"], 4: ["I", "Q", "U", "V"]}.get(n, ["C%d" % i for i in range(n)])


def write_exr(path, image, channel_names=None):
    """image: [H, W] or [H, W, C] array (stored as float32).  C = 1 -> Y, 3 -> R,G,B, 4 -> Stokes I,Q,U,V by default."""
    img = np.asarray(image, dtype=np.float32)
    if img.ndim == 2:
        img = img[..., None]
    if img.ndim != 3:
        raise ValueError("write_exr: expected an [H, W] or [H, W, C] array")
    h, w, c = img.shape
    names = list(channel_names) if channel_names is not None else default_channel_names(c)
    if len(names) != c or len(set(names)) != c:
        raise ValueError("write_exr: need %d distinct channel names" % c)
    order = sorted(range(c), key=lambda i: names[i])               # the channel list is sorted by name
    chlist = b"".join(names[i].encode() + b"\0" + struct.pack("<iB3xii", _FLOAT, 0, 1, 1) for i in order) + b"\0"
    box = struct.pack("<4i", 0, 0, w - 1, h - 1)
    header = (struct.pack("<ii", _MAGIC, 2) +
              _attr("channels", "chlist", chlist) +
              _attr("compression", "compression", b"\0") +
              _attr("dataWindow", "box2i", box) +
              _attr("displayWindow", "box2i", box) +
              _attr("lineOrder", "lineOrder", b"\0") +
              _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) +
              _attr("screenWindowCenter", "v2f", struct.pack("<2f", 0.0, 0.0)) +
              _attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) +
              b"\0")
    row_bytes = 4 * w * c
    first = len(header) + 8 * h
    offsets = np.arange(h, dtype=np.uint64) * np.uint64(8 + row_bytes) + np.uint64(first)
    planar = np.ascontiguousarray(img[:, :, order].transpose(0, 2, 1))   # per scanline: channel after channel
    with open(path, "wb") as f:
        f.write(header)
        f.write(offsets.astype("<u8").tobytes())
        for y in range(h):
            f.write(struct.pack("<ii", y, row_bytes))
            f.write(planar[y].astype("<f4").tobytes())


def _cstr(buf, pos):
    end = buf.index(b"\0", pos)
    return buf[pos:end].decode(), end + 1


def read_exr(path):
    """Reads an uncompressed FLOAT scanline file (as written by write_exr).  Returns (image [H, W, C], channel names in
    file order)."""
    buf = open(path, "rb").read()
    magic, version = struct.unpack_from("<ii", buf, 0)
    if magic != _MAGIC or (version & 0xff) != 2 or (version & ~0xff):
        raise ValueError("read_exr: not a single-part scanline OpenEXR file")
    pos, attrs = 8, {}
    while buf[pos] != 0:
        name, pos = _cstr(buf, pos)
        typ, pos = _cstr(buf, pos)
        (size,) = struct.unpack_from("<i", buf, pos)
        attrs[name] = (typ, buf[pos + 4:pos + 4 + size])
        pos += 4 + size
    pos += 1
    if attrs["compression"][1] != b"\0":
        raise ValueError("read_exr: only uncompressed files are supported")
    xmin, ymin, xmax, ymax = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = xmax - xmin + 1, ymax - ymin + 1
    names, p, ch = [], 0, attrs["channels"][1]
    while ch[p] != 0:
        n, p = _cstr(ch, p)
        ptype, = struct.unpack_from("<i", ch, p)
        if ptype != _FLOAT:
            raise ValueError("read_exr: only FLOAT channels are supported")
        p += 16
        names.append(n)
    c = len(names)
    offsets = np.frombuffer(buf, dtype="<u8", count=h, offset=pos)
    img = np.empty((h, w, c), dtype=np.float32)
    for k in range(h):
        o = int(offsets[k])
        y, size = struct.unpack_from("<ii", buf, o)
        row = np.frombuffer(buf, dtype="<f4", count=w * c, offset=o + 8).reshape(c, w)
        img[y - ymin] = row.T
    return img, names
