"""mi.VolumeGrid: the `.vol` (version 3) voxel-grid container read by gridvolume's `filename`.

Mirrors src/render/volumegrid.cpp:29-123 and its Python binding
(src/render/python/volumegrid_v.cpp): header "VOL", u8 version = 3, i32 type = 1 (float32),
i32 nx, ny, nz, i32 channels, 6 x f32 bounding box, then nx*ny*nz*channels little-endian
float32 values with x varying fastest and the channel index innermost, i.e. a C-ordered
[z][y][x][c] array -- the layout the device kernels index directly.
"""
import struct

import numpy as np

_HEADER = struct.Struct("<3sBiiiii6f")


class VolumeGrid:
    def __init__(self, source, compute_max=True):
        self._bbox = (0.0, 0.0, 0.0, 1.0, 1.0, 1.0)
        if isinstance(source, (str, bytes)) or hasattr(source, "__fspath__"):
            self._read(source)
        else:
            a = np.asarray(source, dtype=np.float32)
            if a.ndim == 3:
                a = a[..., None]
            if a.ndim != 4:
                raise ValueError("VolumeGrid: expected an array of shape [z, y, x] or [z, y, x, channels]")
            self._data = np.ascontiguousarray(a)
        if compute_max:
            self._max = float(self._data.max()) if self._data.size else 0.0
            self._max_per_channel = [float(self._data[..., c].max()) for c in range(self._data.shape[3])]
        else:
            self._max = 0.0
            self._max_per_channel = [0.0] * self._data.shape[3]

    def _read(self, path):
        with open(path, "rb") as f:
            raw = f.read(_HEADER.size)
            if len(raw) < 3 or raw[:3] != b"VOL":
                raise RuntimeError("Invalid volume file!")
            if len(raw) < _HEADER.size:
                raise RuntimeError("Invalid volume file!")
            _, version, data_type, nx, ny, nz, ch, *bbox = _HEADER.unpack(raw)
            if version != 3:
                raise RuntimeError("Invalid version, currently only version 3 is supported (found %d)" % version)
            if data_type != 1:
                raise RuntimeError("Wrong type, currently only type == 1 (Float32) data is supported (found type = %d)" % data_type)
            count = nx * ny * nz * ch
            data = np.fromfile(f, dtype="<f4", count=count)
            if data.size != count:
                raise RuntimeError("Invalid volume file! (truncated: %d of %d values)" % (data.size, count))
        self._bbox = tuple(float(b) for b in bbox)
        self._data = np.ascontiguousarray(data.reshape(nz, ny, nx, ch).astype(np.float32))

    def write(self, path):
        nz, ny, nx, ch = self._data.shape
        with open(path, "wb") as f:
            f.write(_HEADER.pack(b"VOL", 3, 1, nx, ny, nz, ch, *self._bbox))
            self._data.astype("<f4").tofile(f)

    def size(self):
        nz, ny, nx, _ = self._data.shape
        return [nx, ny, nz]

    def channel_count(self):
        return self._data.shape[3]

    def max(self):
        return self._max

    def max_per_channel(self):
        return list(self._max_per_channel)

    def bbox(self):
        return self._bbox[:3], self._bbox[3:]

    def set_bbox(self, bmin, bmax):
        self._bbox = tuple(float(x) for x in list(bmin) + list(bmax))

    def __array__(self, dtype=None, copy=None):
        a = self._data if self._data.shape[3] > 1 else self._data
        return a.astype(dtype) if dtype is not None else a


def read_vol(path):
    """-> (data [z, y, x, c] float32, (bbox_min, bbox_max))."""
    g = VolumeGrid(path)
    return np.asarray(g), g.bbox()
