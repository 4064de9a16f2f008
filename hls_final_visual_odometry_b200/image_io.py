"""Frame files for the host side (SURVEY.md row F4, I/O adapters): the reference's raw 8-bit `.dat` frames
(/root/reference/src/demo.cpp:82-110) and KITTI-style PNG files, decoded with zlib + numpy (no imaging library).
Mirrors host/image_io.h; the PNG writer exists for tests and for exporting synthetic sequences."""
from __future__ import annotations

import struct
import zlib

import numpy as np

_SIG = b"\x89PNG\r\n\x1a\n"


def _paeth(a, b, c):
    p = a.astype(np.int32) + b - c
    pa, pb, pc = np.abs(p - a), np.abs(p - b), np.abs(p - c)
    return np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))


def decode_png(data: bytes) -> np.ndarray:
    """8-bit grey image (H, W) from a non-interlaced PNG of depth 8/16 and colour type 0, 2, 4 or 6 (as host/image_io.h)."""
    if data[:8] != _SIG:
        raise ValueError("not a PNG file")
    pos, idat, hdr = 8, [], None
    while pos + 12 <= len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        if pos + 12 + n > len(data):
            raise ValueError("truncated PNG chunk")
        body = data[pos + 8:pos + 8 + n]
        if zlib.crc32(typ + body) != struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0]:
            raise ValueError("PNG chunk CRC mismatch")
        if typ == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat.append(body)
        elif typ == b"IEND":
            break
        pos += 12 + n
    if hdr is None or not idat:
        raise ValueError("PNG without IHDR / IDAT")
    w, h, depth, ctype, comp, flt, interlace = hdr
    if interlace or comp or flt:
        raise ValueError("interlaced PNG files are not supported")
    if depth not in (8, 16) or ctype not in (0, 2, 4, 6):
        raise ValueError("PNG bit depth / colour type not supported")
    ch = {0: 1, 2: 3, 4: 2, 6: 4}[ctype]
    bpp = ch * depth // 8
    stride = bpp * w
    try:
        raw = np.frombuffer(zlib.decompress(b"".join(idat)), dtype=np.uint8)
    except zlib.error as e:
        raise ValueError(f"PNG image data does not inflate: {e}") from None
    if raw.size != (stride + 1) * h:
        raise ValueError("PNG image data does not inflate to the declared size")
    raw = raw.reshape(h, stride + 1)
    out = np.zeros((h, stride), dtype=np.uint8)
    prev = np.zeros(stride, dtype=np.int32)
    for y in range(h):
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:  # filters that look left: bpp interleaved recurrences
            cur = np.zeros(stride, dtype=np.int32)
            for x in range(0, stride, bpp):
                a = cur[x - bpp:x] if x else np.zeros(bpp, dtype=np.int32)
                b = prev[x:x + bpp]
                c = prev[x - bpp:x] if x else np.zeros(bpp, dtype=np.int32)
                pred = a if ft == 1 else ((a + b) >> 1 if ft == 3 else _paeth(a, b, c))
                if ft not in (1, 3, 4):
                    raise ValueError("unknown PNG row filter")
                cur[x:x + bpp] = (line[x:x + bpp] + pred) & 255
        out[y] = cur
        prev = cur
    px = out.reshape(h, w, bpp)
    step = depth // 8
    if ch <= 2:
        return np.ascontiguousarray(px[:, :, 0])
    r, g, b = (px[:, :, k * step].astype(np.uint32) for k in range(3))
    return ((77 * r + 150 * g + 29 * b) >> 8).astype(np.uint8)


def encode_png(img: np.ndarray, filters=None, depth: int = 8, rgb: bool = False) -> bytes:
    """Grey (H, W) u8 image as a PNG; `filters` = one filter type (0..4) per row (default: all 0).  depth 16 stores the
    value in the high byte; rgb=True replicates the grey value into three channels."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    ch = 3 if rgb else 1
    step = depth // 8
    bpp = ch * step
    px = np.zeros((h, w, bpp), dtype=np.uint8)
    for k in range(ch):
        px[:, :, k * step] = img
    rows = px.reshape(h, w * bpp).astype(np.int32)
    filters = [0] * h if filters is None else list(filters)
    out = bytearray()
    zero = np.zeros(w * bpp, dtype=np.int32)
    for y in range(h):
        cur, up = rows[y], rows[y - 1] if y else zero
        a = np.concatenate([np.zeros(bpp, dtype=np.int32), cur[:-bpp]])
        c = np.concatenate([np.zeros(bpp, dtype=np.int32), up[:-bpp]])
        ft = filters[y]
        pred = {0: zero, 1: a, 2: up, 3: (a + up) >> 1, 4: _paeth(a, up, c)}[ft]
        out.append(ft)
        out += ((cur - pred) & 255).astype(np.uint8).tobytes()

    def chunk(typ, body):
        return struct.pack(">I", len(body)) + typ + body + struct.pack(">I", zlib.crc32(typ + body))

    ihdr = struct.pack(">IIBBBBB", w, h, depth, 2 if rgb else 0, 0, 0, 0)
    z = zlib.compress(bytes(out), 6)
    half = len(z) // 2  # two IDAT chunks: readers must concatenate them
    return _SIG + chunk(b"IHDR", ihdr) + chunk(b"IDAT", z[:half]) + chunk(b"IDAT", z[half:]) + chunk(b"IEND", b"")


def read_frame(path: str, width: int = 0, height: int = 0) -> np.ndarray:
    """One frame as (H, W) u8: `.png` is decoded, anything else is a raw width*height frame as the reference demo reads it."""
    data = open(path, "rb").read()
    if path.lower().endswith(".png"):
        img = decode_png(data)
        if (width and img.shape[1] != width) or (height and img.shape[0] != height):
            raise ValueError(f"{path} is {img.shape[1]}x{img.shape[0]}, expected {width}x{height}")
        return img
    if width <= 0 or height <= 0 or len(data) != width * height:
        raise ValueError(f"raw frame {path} does not hold width*height bytes")
    return np.frombuffer(data, dtype=np.uint8).reshape(height, width).copy()
