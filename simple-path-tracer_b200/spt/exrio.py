"""A small OpenEXR reader for environment maps (InputParams::get_exr_image, src/core/loader.rs:374-390, which reads the
first RGBA layer through the `exr` crate): single-part scanline files, HALF / FLOAT channels, compression NONE, RLE,
ZIPS and ZIP.  Returns rows top to bottom, image[y][x] = (R, G, B), like the reference's Vec<Vec<Color>>.
Tiled, multi-part, deep, PIZ / PXR24 / B44 / DWA files are rejected with a message that says which feature is missing."""
import struct
import zlib

import numpy as np

COMPRESSION = {0: "none", 1: "rle", 2: "zips", 3: "zip", 4: "piz", 5: "pxr24", 6: "b44", 7: "b44a", 8: "dwaa", 9: "dwab"}
LINES_PER_BLOCK = {0: 1, 1: 1, 2: 1, 3: 16}


class ExrError(ValueError):
    pass


def _cstr(buf, p):
    e = buf.index(b"\0", p)
    return buf[p:e].decode("latin-1"), e + 1


def _unpredict(raw):
    """inverse of OpenEXR's byte predictor + even/odd interleave (ImfZip.cpp / ImfRle.cpp)"""
    t = np.frombuffer(raw, np.uint8).astype(np.int64)
    t[1:] -= 128
    t = (np.cumsum(t) & 0xFF).astype(np.uint8)
    n = len(t)
    out = np.empty(n, np.uint8)
    half = (n + 1) // 2
    out[0::2] = t[:half]
    out[1::2] = t[half:]
    return out.tobytes()


def _unrle(src, expected):
    out = bytearray()
    p = 0
    while p < len(src):
        c = src[p] if src[p] < 128 else src[p] - 256
        p += 1
        if c < 0:
            out += src[p:p - c]; p -= c
        else:
            out += bytes([src[p]]) * (c + 1); p += 1
    if len(out) != expected:
        raise ExrError("exr: RLE block has the wrong size")
    return bytes(out)


def read_exr(path):
    with open(path, "rb") as fh:
        buf = fh.read()
    if len(buf) < 8 or struct.unpack_from("<I", buf, 0)[0] != 20000630:
        raise ExrError(f"'{path}' is not an OpenEXR file")
    version, = struct.unpack_from("<I", buf, 4)
    if version & 0x200:
        raise ExrError("exr: tiled files are not supported")
    if version & 0x1800:
        raise ExrError("exr: deep / multi-part files are not supported")
    p = 8
    attrs = {}
    while buf[p] != 0:
        name, p = _cstr(buf, p)
        ty, p = _cstr(buf, p)
        size, = struct.unpack_from("<i", buf, p); p += 4
        attrs[name] = (ty, buf[p:p + size]); p += size
    p += 1
    for need in ("channels", "compression", "dataWindow"):
        if need not in attrs:
            raise ExrError(f"exr: header without '{need}'")
    comp = attrs["compression"][1][0]
    if comp not in LINES_PER_BLOCK:
        raise ExrError(f"exr: compression '{COMPRESSION.get(comp, comp)}' is not supported (none, rle, zips, zip are)")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    channels = []  # (name, pixel type) in file (alphabetical) order
    cb, q = attrs["channels"][1], 0
    while cb[q] != 0:
        name, q = _cstr(cb, q)
        ptype, _pl, xs, ys = struct.unpack_from("<iB3xii", cb, q); q += 16
        if xs != 1 or ys != 1:
            raise ExrError("exr: sub-sampled channels are not supported")
        if ptype not in (1, 2):
            raise ExrError("exr: UINT channels are not supported")
        channels.append((name, ptype))
    bpp = [2 if t == 1 else 4 for _, t in channels]
    line_bytes = w * sum(bpp)
    lpb = LINES_PER_BLOCK[comp]
    n_blocks = (h + lpb - 1) // lpb
    offsets = struct.unpack_from(f"<{n_blocks}Q", buf, p)
    planes = {name: np.zeros((h, w), np.float32) for name, _ in channels}
    for off in offsets:
        y, size = struct.unpack_from("<ii", buf, off)
        data = buf[off + 8:off + 8 + size]
        rows = min(lpb, y1 - y + 1)
        expected = rows * line_bytes
        if comp != 0 and size < expected:  # a block that did not shrink is stored raw
            data = _unpredict(zlib.decompress(data) if comp in (2, 3) else _unrle(data, expected))
        if len(data) != expected:
            raise ExrError("exr: scanline block has the wrong size")
        q = 0
        for r in range(rows):
            for (name, ptype), b in zip(channels, bpp):
                row = np.frombuffer(data, np.float16 if ptype == 1 else np.float32, w, q)
                planes[name][y - y0 + r] = row.astype(np.float32)
                q += w * b
    def pick(*names):
        for n in names:
            for k in planes:
                if k == n or k.endswith("." + n):
                    return planes[k]
        return None
    r, g, b = pick("R"), pick("G"), pick("B")
    if r is None or g is None or b is None:
        y_ = pick("Y")
        if y_ is None:
            raise ExrError("exr: no R, G, B (or Y) channels")
        r = g = b = y_
    return np.stack([r, g, b], -1).astype(np.float32)
