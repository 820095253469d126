"""Minimal OpenEXR scanline reader (numpy + zlib) for the two writers on this path: our imageio.h (uncompressed half) and the
reference's tinyexr export (ZIP, 16-line blocks, half; util/image.cpp:28-98). -> (H, W, 4) float32 RGBA in FILE row order
(both writers store the film vertically flipped)."""
import struct
import zlib

import numpy as np


def _unzip_block(raw, expect):
    if len(raw) == expect:
        return raw
    d = np.frombuffer(zlib.decompress(raw), np.uint8).astype(np.int32)
    # predictor: d[i] = d[i-1] + d[i] - 128 (mod 256)
    d = (np.cumsum(d - 128) + 128) & 0xFF
    d = d.astype(np.uint8)
    # de-interleave: first half -> even bytes, second half -> odd bytes
    n = len(d)
    out = np.empty(n, np.uint8)
    half = (n + 1) // 2
    out[0::2] = d[:half]
    out[1::2] = d[half:]
    return out.tobytes()


def read_exr(path):
    b = open(path, "rb").read()
    magic, ver = struct.unpack("<II", b[:8])
    assert magic == 20000630 and (ver & 0x200) == 0, "not a scanline OpenEXR file"
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos)
        name = b[pos:e].decode()
        pos = e + 1
        e = b.index(b"\0", pos)
        typ = b[pos:e].decode()
        pos = e + 1
        size = struct.unpack("<i", b[pos:pos + 4])[0]
        attrs[name] = (typ, b[pos + 4:pos + 4 + size])
        pos += 4 + size
    pos += 1
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    comp = attrs["compression"][1][0]
    lines_per_block = {0: 1, 2: 1, 3: 16}[comp]
    # channel list: name\0, pixel type i32, pLinear u8 + 3 pad, xSampling i32, ySampling i32
    chans, cb, p = [], attrs["channels"][1], 0
    while cb[p] != 0:
        e = cb.index(b"\0", p)
        nm = cb[p:e].decode()
        ptype = struct.unpack("<i", cb[e + 1:e + 5])[0]
        chans.append((nm, ptype))
        p = e + 1 + 16
    nblocks = (h + lines_per_block - 1) // lines_per_block
    offs = struct.unpack("<%dQ" % nblocks, b[pos:pos + 8 * nblocks])
    bpp = sum(2 if t == 1 else 4 for _, t in chans)
    img = np.zeros((h, w, 4), np.float32)
    slot = {"R": 0, "G": 1, "B": 2, "A": 3}
    for o in offs:
        yy, n = struct.unpack("<ii", b[o:o + 8])
        rows = min(lines_per_block, y1 + 1 - yy)
        data = b[o + 8:o + 8 + n]
        if comp != 0:
            data = _unzip_block(data, rows * w * bpp)
        q = 0
        for r in range(rows):
            for nm, t in chans:
                if t == 1:
                    v = np.frombuffer(data, np.float16, w, q).astype(np.float32)
                    q += 2 * w
                elif t == 2:
                    v = np.frombuffer(data, np.float32, w, q)
                    q += 4 * w
                else:
                    v = np.frombuffer(data, np.uint32, w, q).astype(np.float32)
                    q += 4 * w
                if nm in slot:
                    img[yy - y0 + r, :, slot[nm]] = v
    return img
