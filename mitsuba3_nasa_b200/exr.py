"""Minimal OpenEXR scanline writer / reader for developed films (the role of `Bitmap::write` / `HDRFilm::write`,
src/films/hdrfilm.cpp:420-590, for the float32 case): FLOAT channels, uncompressed or ZIPS / ZIP / PIZ compressed.

Written from the OpenEXR file-layout specification (magic, version, attribute list, line-offset table, chunks); files
open in any EXR viewer.  The reader understands single-part scanline files with HALF / FLOAT / UINT channels,
uncompressed or RLE / ZIPS / ZIP / PIZ compressed -- PIZ is what the reference writes, so `film.write()` outputs and the reference
images of the fork's test-suite can be read back for comparisons.
"""
import struct

import numpy as np

_MAGIC = 20000630
_FLOAT = 2
_LINES_PER_BLOCK = {0: 1, 1: 1, 2: 1, 3: 16, 4: 32}          # scanlines per chunk: NONE, RLE, ZIPS, ZIP, PIZ


def _attr(name, typ, data):
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(data)) + data


def default_channel_names(n):
    return {1: ["Y"], 3: ["R", "G", "B"], 4: ["I", "Q", "U", "V"]}.get(n, ["C%d" % i for i in range(n)])


def write_exr(path, image, channel_names=None, compression="none", component_format="float32"):
    """image: [H, W] or [H, W, C] array (stored as float32).  C = 1 -> Y, 3 -> R,G,B, 4 -> Stokes I,Q,U,V by default.
    compression: "none" (one scanline per chunk), "zips" (deflate, one scanline), "zip" (deflate, 16 scanlines) or "piz"
    (wavelet + Huffman, 32 scanlines; the reference's default).  component_format: "float32" or "float16" (hdrfilm's
    default, src/films/hdrfilm.cpp:60-64: HALF channels)."""
    if component_format not in ("float32", "float16"):
        raise ValueError("write_exr: component_format must be 'float32' or 'float16'")
    half = component_format == "float16"
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
    chlist = b"".join(names[i].encode() + b"\0" + struct.pack("<iB3xii", 1 if half else _FLOAT, 0, 1, 1) for i in order) + b"\0"
    box = struct.pack("<4i", 0, 0, w - 1, h - 1)
    if compression not in ("none", "zips", "zip", "piz"):
        raise ValueError("write_exr: compression must be 'none', 'zips', 'zip' or 'piz'")
    comp = {"none": 0, "zips": 2, "zip": 3, "piz": 4}[compression]
    header = (struct.pack("<ii", _MAGIC, 2) +
              _attr("channels", "chlist", chlist) +
              _attr("compression", "compression", bytes([comp])) +
              _attr("dataWindow", "box2i", box) +
              _attr("displayWindow", "box2i", box) +
              _attr("lineOrder", "lineOrder", b"\0") +
              _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) +
              _attr("screenWindowCenter", "v2f", struct.pack("<2f", 0.0, 0.0)) +
              _attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) +
              b"\0")
    planar = np.ascontiguousarray(img[:, :, order].transpose(0, 2, 1))   # per scanline: channel after channel
    lines = _LINES_PER_BLOCK[comp]
    chunks = []
    for y0 in range(0, h, lines):
        raw = planar[y0:y0 + lines].astype("<f2" if half else "<f4").tobytes()
        data = raw
        if comp == 4:
            ny = min(lines, h - y0)
            data = _piz_encode(np.frombuffer(raw, "<u2").reshape(ny, -1), w, ny, [1 if half else 2] * c)
            if len(data) >= len(raw):
                data = raw
        elif comp:
            data = _deflate_block(raw)
            if len(data) >= len(raw):
                data = raw                                               # a block that does not shrink is stored as it is
        chunks.append(struct.pack("<ii", y0, len(data)) + data)
    pos = len(header) + 8 * len(chunks)
    with open(path, "wb") as f:
        f.write(header)
        for ch in chunks:
            f.write(struct.pack("<Q", pos))
            pos += len(ch)
        for ch in chunks:
            f.write(ch)


def _deflate_block(raw):
    """ZIP / ZIPS: split the bytes into even and odd halves, delta-code them (+128), deflate."""
    import zlib
    a = np.frombuffer(raw, np.uint8)
    t = np.concatenate([a[0::2], a[1::2]]).astype(np.int64)
    t[1:] = (t[1:] - t[:-1] + 128) & 0xff
    return zlib.compress(t.astype(np.uint8).tobytes())


def _cstr(buf, pos):
    end = buf.index(b"\0", pos)
    return buf[pos:end].decode(), end + 1


# ----------------------------------------------------------------------------------------------------------------
# Block decoders.  Written from the published OpenEXR file layout and the description of its codecs (the pinned
# third-party dependency of the reference is ext/openexr, OpenEXR 2.5: IlmImf/ImfPizCompressor.cpp, ImfHuf.cpp,
# ImfWav.cpp, ImfZip.cpp, ImfRle.cpp); validated against that library's own test images of one picture stored with
# every codec (tests/test_exr.py).  Every decoder returns the block in the uncompressed chunk layout: per scanline,
# channel after channel, little-endian samples.
# ----------------------------------------------------------------------------------------------------------------
_SAMPLE_BYTES = {0: 4, 1: 2, 2: 4}                            # UINT, HALF, FLOAT
_SAMPLE_DTYPE = {0: "<u4", 1: "<f2", 2: "<f4"}


def _unpredict(raw):
    """Inverse of the byte predictor + de-interleave shared by the ZIP and RLE codecs."""
    a = np.frombuffer(raw, dtype=np.uint8).astype(np.int64)
    a[1:] -= 128
    a = (np.cumsum(a) & 0xff).astype(np.uint8)
    half = (len(a) + 1) // 2
    out = np.empty(len(a), np.uint8)
    out[0::2] = a[:half]
    out[1::2] = a[half:]
    return out.tobytes()


def _rle_decode(data, expected):
    out, p = bytearray(), 0
    while p < len(data):
        n = data[p] - 256 if data[p] > 127 else data[p]
        p += 1
        if n < 0:
            out += data[p:p - n]
            p += -n
        else:
            out += bytes([data[p]]) * (n + 1)
            p += 1
    if len(out) != expected:
        raise ValueError("read_exr: corrupt RLE block")
    return bytes(out)


class _Bits:
    """MSB-first bit reader over a bytes object."""
    def __init__(self, data, pos=0):
        self.d, self.p, self.buf, self.n = data, pos, 0, 0

    def get(self, k):
        while self.n < k:
            self.buf = (self.buf << 8) | self.d[self.p]
            self.p += 1
            self.n += 8
        self.n -= k
        return (self.buf >> self.n) & ((1 << k) - 1)


_HUF_ENCSIZE = (1 << 16) + 1       # 65536 values + the run-length symbol
_HUF_TABLE_BITS = 12


def _huf_uncompress(data, n_raw):
    """Canonical Huffman code with 6-bit packed code lengths (runs of zero lengths are run-length coded), the symbol with
    the highest index standing for "repeat the previous output value <next 8 bits> times"."""
    if len(data) < 20:
        if n_raw:
            raise ValueError("read_exr: truncated PIZ block")
        return np.zeros(0, np.uint16)
    im, i_max, _, n_bits, _ = struct.unpack_from("<5I", data, 0)
    if im >= _HUF_ENCSIZE or i_max >= _HUF_ENCSIZE:
        raise ValueError("read_exr: corrupt Huffman table")
    # code lengths
    lengths = np.zeros(_HUF_ENCSIZE, np.int64)
    br = _Bits(data, 20)
    sym = im
    while sym <= i_max:
        l = br.get(6)
        if l == 63:
            sym += br.get(8) + 6
        elif l >= 59:
            sym += l - 59 + 2
        else:
            lengths[sym] = l
            sym += 1
    start = br.p                                     # the code data starts at the next whole byte
    # canonical codes: within one length in symbol order; shorter codes have the numerically larger prefixes
    count = np.bincount(lengths, minlength=59)
    base = [0] * 59
    c = 0
    for l in range(58, 0, -1):
        nc = (c + int(count[l])) >> 1
        base[l] = c
        c = nc
    syms_of = {l: np.nonzero(lengths == l)[0].tolist() for l in range(1, 59) if count[l]}
    # prefix table for the short codes
    T = _HUF_TABLE_BITS
    tab_sym, tab_len = [0] * (1 << T), [0] * (1 << T)
    for l, syms in syms_of.items():
        if l > T:
            continue
        for k, sy in enumerate(syms):
            lo = (base[l] + k) << (T - l)
            for j in range(lo, lo + (1 << (T - l))):
                tab_sym[j], tab_len[j] = sy, l
    long_lengths = sorted(l for l in syms_of if l > T)
    out = [0] * n_raw
    n_out = 0
    d = data
    pos, buf, nb = start, 0, 0
    end_bits = n_bits
    used = 0
    size = len(d)
    while used < end_bits and n_out <= n_raw:
        while nb < 58 and pos < size:                # keep at least 58 bits (longest code) in the window when possible
            buf = ((buf << 8) | d[pos]) & 0xffffffffffffffffffffffff
            pos += 1
            nb += 8
        if nb >= T:
            idx = (buf >> (nb - T)) & ((1 << T) - 1)
        else:
            idx = (buf << (T - nb)) & ((1 << T) - 1)
        l = tab_len[idx]
        if l:
            sy = tab_sym[idx]
        else:
            sy = -1
            for l in long_lengths:
                if nb < l:
                    break
                code = (buf >> (nb - l)) & ((1 << l) - 1)
                k = code - base[l]
                if 0 <= k < len(syms_of[l]):
                    sy = syms_of[l][k]
                    break
            if sy < 0:
                raise ValueError("read_exr: corrupt Huffman data")
        nb -= l
        used += l
        if sy == i_max:                              # run of the previous value
            if nb < 8:
                buf = ((buf << 8) | d[pos]) & 0xffffffffffffffffffffffff
                pos += 1
                nb += 8
            run = (buf >> (nb - 8)) & 0xff
            nb -= 8
            used += 8
            if n_out == 0 or n_out + run > n_raw:
                raise ValueError("read_exr: corrupt Huffman data")
            out[n_out:n_out + run] = [out[n_out - 1]] * run
            n_out += run
        else:
            if n_out >= n_raw:
                raise ValueError("read_exr: corrupt Huffman data")
            out[n_out] = sy
            n_out += 1
    if n_out != n_raw:
        raise ValueError("read_exr: Huffman data ended early")
    return np.array(out, dtype=np.uint16)


def _wdec14(l, h):
    ls, hs = l.astype(np.int16).astype(np.int32), h.astype(np.int16).astype(np.int32)
    ai = ls + (hs & 1) + (hs >> 1)
    return ai.astype(np.uint16), (ai - hs).astype(np.uint16)          # wraps like the C shorts


def _wdec16(l, h):
    m, d = l.astype(np.int32), h.astype(np.int32)
    bb = (m - (d >> 1)) & 0xffff
    aa = (d + bb - 0x8000) & 0xffff
    return aa.astype(np.uint16), bb.astype(np.uint16)


def _wav2_decode(a, max_value):
    """Inverse of the 2-D Haar-like integer wavelet, in place on a [ny, nx] uint16 array (one 16-bit plane of one channel).
    Every level touches disjoint 2x2 cells with stride p2, so a level is four strided array operations."""
    ny, nx = a.shape
    dec = _wdec14 if max_value < (1 << 14) else _wdec16
    n = min(nx, ny)
    p = 1
    while p <= n:
        p <<= 1
    p >>= 1
    p2 = p
    p >>= 1
    while p >= 1:
        cy = (ny - p2) // p2 + 1 if ny >= p2 else 0      # cells with a full 2x2 footprint
        cx = (nx - p2) // p2 + 1 if nx >= p2 else 0
        ys, xs = slice(0, cy * p2, p2), slice(0, cx * p2, p2)
        ys1, xs1 = slice(p, p + cy * p2, p2), slice(p, p + cx * p2, p2)
        if cy and cx:
            i00, i10 = dec(a[ys, xs], a[ys1, xs])
            i01, i11 = dec(a[ys, xs1], a[ys1, xs1])
            a[ys, xs], a[ys, xs1] = dec(i00, i01)
            a[ys1, xs], a[ys1, xs1] = dec(i10, i11)
        if (nx & p) and cy:                              # odd column: 1-D step in y at x = cx * p2
            x = cx * p2
            a[ys, x], a[ys1, x] = dec(a[ys, x], a[ys1, x])
        if (ny & p) and cx:                              # odd line: 1-D step in x at y = cy * p2
            y = cy * p2
            a[y, xs], a[y, xs1] = dec(a[y, xs], a[y, xs1])
        p2 = p
        p >>= 1


def _piz_decode(data, nx, ny, sample_words):
    """PIZ: value-range bitmap -> look-up table, Huffman-coded wavelet coefficients, one plane set per channel."""
    min_nz, max_nz = struct.unpack_from("<HH", data, 0)
    pos = 4
    bitmap = np.zeros(8192, np.uint8)
    if min_nz <= max_nz:
        if max_nz >= 8192:
            raise ValueError("read_exr: corrupt PIZ bitmap")
        bitmap[min_nz:max_nz + 1] = np.frombuffer(data, np.uint8, max_nz - min_nz + 1, pos)
        pos += max_nz - min_nz + 1
    present = np.unpackbits(bitmap, bitorder="little").astype(bool)
    present[0] = True                                     # zero is always part of the table
    lut = np.zeros(65536, np.uint16)
    values = np.nonzero(present)[0]
    lut[:len(values)] = values
    max_value = len(values) - 1
    (length,) = struct.unpack_from("<i", data, pos)
    pos += 4
    if length < 0 or pos + length > len(data):
        raise ValueError("read_exr: truncated PIZ block")
    total = sum(nx * ny * w for w in sample_words)
    words = _huf_uncompress(data[pos:pos + length], total)
    out = np.empty((ny, sum(nx * w for w in sample_words)), np.uint16)
    start, col = 0, 0
    for w in sample_words:
        plane = words[start:start + nx * ny * w].reshape(ny, nx, w).copy()
        for j in range(w):
            sub = np.ascontiguousarray(plane[:, :, j])
            _wav2_decode(sub, max_value)
            plane[:, :, j] = sub
        out[:, col:col + nx * w] = lut[plane.reshape(ny, nx * w)]
        start += nx * ny * w
        col += nx * w
    return out.astype("<u2").tobytes()


def read_exr(path):
    """See _read_exr; damaged files raise ValueError."""
    try:
        return _read_exr(path)
    except (IndexError, struct.error, KeyError) as e:
        raise ValueError("read_exr: corrupt or truncated file (%s: %s)" % (type(e).__name__, e))


def _read_exr(path):
    """Reads a single-part scanline OpenEXR file: HALF / FLOAT / UINT channels, stored uncompressed or with the RLE, ZIPS, ZIP
    or PIZ codec (PIZ is what the reference's `Bitmap::write` produces, i.e. every `film.write()` output and every reference
    image of the fork's test-suite).  Returns (image [H, W, C] float32, channel names in file order)."""
    buf = open(path, "rb").read()
    magic, version = struct.unpack_from("<ii", buf, 0)
    if magic != _MAGIC or (version & 0xff) != 2 or (version & ~0xff & ~0x400):
        raise ValueError("read_exr: not a single-part scanline OpenEXR file")
    pos, attrs = 8, {}
    while buf[pos] != 0:
        name, pos = _cstr(buf, pos)
        typ, pos = _cstr(buf, pos)
        (size,) = struct.unpack_from("<i", buf, pos)
        attrs[name] = (typ, buf[pos + 4:pos + 4 + size])
        pos += 4 + size
    pos += 1
    comp = attrs["compression"][1][0]
    if comp not in _LINES_PER_BLOCK:
        raise ValueError("read_exr: compression %d is not supported (NONE, RLE, ZIPS, ZIP and PIZ are)" % comp)
    xmin, ymin, xmax, ymax = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = xmax - xmin + 1, ymax - ymin + 1
    names, types, p, ch = [], [], 0, attrs["channels"][1]
    while ch[p] != 0:
        n, p = _cstr(ch, p)
        ptype, _lin, xs, ys = struct.unpack_from("<iB3xii", ch, p)
        if ptype not in _SAMPLE_BYTES or xs != 1 or ys != 1:
            raise ValueError("read_exr: unsupported channel '%s' (pixel type %d, sampling %d x %d)" % (n, ptype, xs, ys))
        p += 16
        names.append(n)
        types.append(ptype)
    c = len(names)
    line_bytes = w * sum(_SAMPLE_BYTES[t] for t in types)
    lines = _LINES_PER_BLOCK[comp]
    n_blocks = (h + lines - 1) // lines
    offsets = np.frombuffer(buf, dtype="<u8", count=n_blocks, offset=pos)
    img = np.empty((h, w, c), dtype=np.float32)
    for k in range(n_blocks):
        o = int(offsets[k])
        y, size = struct.unpack_from("<ii", buf, o)
        ny = min(lines, ymax - y + 1)
        raw_size = ny * line_bytes
        data = buf[o + 8:o + 8 + size]
        if len(data) != size:
            raise ValueError("read_exr: truncated file")
        if size == raw_size and comp != 0:
            pass                                           # the writer keeps a block raw when the codec does not shrink it
        elif comp in (2, 3):
            import zlib
            data = _unpredict(zlib.decompress(data))
        elif comp == 1:
            data = _unpredict(_rle_decode(data, raw_size))
        elif comp == 4:
            data = _piz_decode(data, w, ny, [_SAMPLE_BYTES[t] // 2 for t in types])
        if len(data) != raw_size:
            raise ValueError("read_exr: block of scanline %d has %d bytes, expected %d" % (y, len(data), raw_size))
        off = 0
        block = np.frombuffer(data, np.uint8).reshape(ny, line_bytes)
        for i, t in enumerate(types):
            nb = w * _SAMPLE_BYTES[t]
            img[y - ymin:y - ymin + ny, :, i] = np.ascontiguousarray(block[:, off:off + nb]).view(_SAMPLE_DTYPE[t]).astype(np.float32)
            off += nb
    return img, names


# ----------------------------------------------------------------------------------------------------------------
# PIZ encoder (the reference's default for Bitmap::write): forward of _piz_decode.  Checked by decoding its output with the
# decoder above, which is pinned on OpenEXR's own PIZ test image.
# ----------------------------------------------------------------------------------------------------------------
def _wenc14(a, b):
    a_s, b_s = a.astype(np.int16).astype(np.int32), b.astype(np.int16).astype(np.int32)
    return ((a_s + b_s) >> 1).astype(np.uint16), (a_s - b_s).astype(np.uint16)


def _wenc16(a, b):
    ao = (a.astype(np.int32) + 0x8000) & 0xffff
    bi = b.astype(np.int32)
    m = (ao + bi) >> 1
    d = ao - bi
    m = np.where(d < 0, (m + 0x8000) & 0xffff, m)
    return m.astype(np.uint16), (d & 0xffff).astype(np.uint16)


def _wav2_encode(a, max_value):
    ny, nx = a.shape
    enc = _wenc14 if max_value < (1 << 14) else _wenc16
    n = min(nx, ny)
    p, p2 = 1, 2
    while p2 <= n:
        cy, cx = (ny - p2) // p2 + 1, (nx - p2) // p2 + 1
        ys, xs = slice(0, cy * p2, p2), slice(0, cx * p2, p2)
        ys1, xs1 = slice(p, p + cy * p2, p2), slice(p, p + cx * p2, p2)
        i00, i01 = enc(a[ys, xs], a[ys, xs1])
        i10, i11 = enc(a[ys1, xs], a[ys1, xs1])
        a[ys, xs], a[ys1, xs] = enc(i00, i10)
        a[ys, xs1], a[ys1, xs1] = enc(i01, i11)
        if nx & p:
            x = cx * p2
            a[ys, x], a[ys1, x] = enc(a[ys, x], a[ys1, x])
        if ny & p:
            y = cy * p2
            a[y, xs], a[y, xs1] = enc(a[y, xs], a[y, xs1])
        p = p2
        p2 <<= 1


def _huf_lengths(freq):
    """Code lengths of a Huffman code for the symbols with freq > 0 (standard two-queue merge on a heap)."""
    import heapq
    syms = np.nonzero(freq)[0]
    lengths = np.zeros(len(freq), np.int64)
    if len(syms) == 1:
        lengths[syms[0]] = 1
        return lengths
    heap = [(int(freq[s]), int(s), None) for s in syms]
    heapq.heapify(heap)
    parent, counter = {}, len(freq)
    while len(heap) > 1:
        fa, a, _ = heapq.heappop(heap)
        fb, b, _ = heapq.heappop(heap)
        parent[a] = parent[b] = counter
        heapq.heappush(heap, (fa + fb, counter, None))
        counter += 1
    depth = {heap[0][1]: 0}
    for node in sorted(parent, reverse=True):          # children have smaller ids than their parents
        depth[node] = depth[parent[node]] + 1
    for s in syms:
        lengths[s] = depth[int(s)]
    if lengths.max() > 58:
        raise ValueError("write_exr: Huffman code longer than 58 bits")
    return lengths


def _huf_compress(words):
    """Inverse of _huf_uncompress (no run symbols are emitted; the run symbol still gets a code, as the reference does)."""
    n = len(words)
    freq = np.bincount(words, minlength=_HUF_ENCSIZE).astype(np.int64)
    im = int(np.nonzero(freq)[0][0])
    i_max = int(np.nonzero(freq)[0][-1]) + 1           # pseudo-symbol for run-length coding
    freq[i_max] = 1
    lengths = _huf_lengths(freq)
    count = np.bincount(lengths, minlength=59)
    base, c = [0] * 59, 0
    for l in range(58, 0, -1):
        nc = (c + int(count[l])) >> 1
        base[l] = c
        c = nc
    codes = np.zeros(_HUF_ENCSIZE, np.uint64)
    for l in range(1, 59):
        if count[l]:
            s = np.nonzero(lengths == l)[0]
            codes[s] = np.uint64(base[l]) + np.arange(len(s), dtype=np.uint64)
    # packed table: 6 bits per length, runs of zero lengths as 59..62 (2..5) or 63 + 8 bits (6..261)
    tl, tc = [], []
    s = im
    while s <= i_max:
        l = int(lengths[s])
        if l == 0:
            run = 1
            while s + run <= i_max and lengths[s + run] == 0 and run < 261:
                run += 1
            if run >= 6:
                tl += [6, 8]; tc += [63, run - 6]
                s += run
                continue
            if run >= 2:
                tl.append(6); tc.append(59 + run - 2)
                s += run
                continue
        tl.append(6); tc.append(l)
        s += 1
    table = _pack_bits(np.array(tc, np.uint64), np.array(tl, np.int64))
    data_len = lengths[words]
    data = _pack_bits(codes[words], data_len)
    n_bits = int(data_len.sum())
    return struct.pack("<5I", im, i_max, len(table), n_bits, 0) + table + data


def _pack_bits(codes, lengths):
    """MSB-first concatenation of codes[i] in lengths[i] bits."""
    total = int(lengths.sum())
    if total == 0:
        return b""
    off = np.cumsum(lengths) - lengths
    idx = np.repeat(np.arange(len(codes)), lengths)
    within = np.arange(total) - np.repeat(off, lengths)
    shift = (lengths[idx] - 1 - within).astype(np.uint64)
    bits = ((codes[idx] >> shift) & np.uint64(1)).astype(np.uint8)
    return np.packbits(bits).tobytes()


def _piz_encode(block_words, nx, ny, sample_words):
    """block_words: the chunk in its uncompressed layout as uint16 [ny, sum(nx * w)]."""
    head, words = _piz_forward(block_words, nx, ny, sample_words)
    huf = _huf_compress(words)
    return head + struct.pack("<i", len(huf)) + huf


def _piz_forward(block_words, nx, ny, sample_words):
    """Everything before the entropy coder: (bitmap header bytes, LUT-mapped wavelet coefficients).  Deterministic, so it can
    be compared with the coefficients stored in a file written by OpenEXR itself."""
    planes, col = [], 0
    for w in sample_words:
        planes.append(np.ascontiguousarray(block_words[:, col:col + nx * w]).reshape(ny, nx, w))
        col += nx * w
    allw = np.concatenate([p.ravel() for p in planes])
    present = np.zeros(65536, bool)
    present[allw] = True
    present0 = present.copy()
    present0[0] = False                                # zero is implicit
    nz = np.nonzero(present0)[0]
    bitmap = np.packbits(present0, bitorder="little")
    present[0] = True
    fwd = (np.cumsum(present) - 1).astype(np.uint16)   # value -> index among the present values
    max_value = int(present.sum()) - 1
    out = []
    for p_ in planes:
        q = fwd[p_]
        for j in range(q.shape[2]):
            sub = np.ascontiguousarray(q[:, :, j])
            _wav2_encode(sub, max_value)
            q[:, :, j] = sub
        out.append(q.ravel())
    if len(nz):
        lo, hi = int(nz[0]) >> 3, int(nz[-1]) >> 3
        head = struct.pack("<HH", lo, hi) + bitmap[lo:hi + 1].tobytes()
    else:
        head = struct.pack("<HH", 8191, 0)
    return head, np.concatenate(out)
