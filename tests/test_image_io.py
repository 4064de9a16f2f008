"""I/O adapters of SURVEY.md row F4: raw `.dat` frames as /root/reference/src/demo.cpp:82-110 reads them and KITTI-style PNG files,
Python reader (image_io.py) and C++ reader (host/image_io.h through libvo_host.so) against the pixels that were written."""
import ctypes as C
import os

import numpy as np
import pytest

from hls_final_visual_odometry_b200 import build, image_io
from hls_final_visual_odometry_b200.synth import make_sequence


def host_read(path, w=0, h=0):
    lib = C.CDLL(build.build_host() or build.HOST_LIB)
    out = np.zeros(1 << 22, dtype=np.uint8)
    cw, ch = C.c_int32(w), C.c_int32(h)
    err = C.create_string_buffer(256)
    st = lib.vo_host_read_frame(path.encode(), out.ctypes.data_as(C.c_void_p), C.c_int64(out.size), C.byref(cw), C.byref(ch), err, 256)
    if st != 0:
        raise ValueError(err.value.decode() or f"status {st}")
    return out[: cw.value * ch.value].reshape(ch.value, cw.value).copy()


@pytest.fixture(scope="module")
def frame():
    return make_sequence(321, 97, 1, seed=4)[0]


@pytest.mark.parametrize("variant", ["none", "mixed_filters", "paeth", "depth16", "rgb", "rgb16_mixed"])
def test_png_round_trip_both_readers(tmp_path, frame, variant):
    h = frame.shape[0]
    rng = np.random.default_rng(3)
    kw = {"none": {}, "mixed_filters": {"filters": rng.integers(0, 5, h)}, "paeth": {"filters": [4] * h},
          "depth16": {"depth": 16, "filters": rng.integers(0, 5, h)}, "rgb": {"rgb": True, "filters": rng.integers(0, 5, h)},
          "rgb16_mixed": {"rgb": True, "depth": 16, "filters": rng.integers(0, 5, h)}}[variant]
    path = str(tmp_path / "000000.png")
    open(path, "wb").write(image_io.encode_png(frame, **kw))
    want = frame  # grey replicated into R, G, B: (77 + 150 + 29) * v >> 8 == v
    assert (image_io.read_frame(path) == want).all()
    assert (host_read(path) == want).all()
    assert (host_read(path, frame.shape[1], frame.shape[0]) == want).all()


def test_raw_dat_frames(tmp_path, frame):
    path = str(tmp_path / "I1_000000.dat")
    frame.tofile(path)
    hgt, wid = frame.shape
    assert (image_io.read_frame(path, wid, hgt) == frame).all()
    assert (host_read(path, wid, hgt) == frame).all()
    with pytest.raises(ValueError):
        image_io.read_frame(path, wid + 1, hgt)
    with pytest.raises(ValueError):
        host_read(path, wid + 1, hgt)


def test_damaged_and_unsupported_png_are_refused(tmp_path, frame):
    good = image_io.encode_png(frame)
    bad_crc = bytearray(good)
    bad_crc[40] ^= 1
    truncated = good[: len(good) // 2]
    interlaced = bytearray(good)
    interlaced[28] = 1  # IHDR interlace byte (its CRC then fails too: either message is a refusal)
    for name, blob in [("crc.png", bytes(bad_crc)), ("cut.png", truncated), ("lace.png", bytes(interlaced)), ("not.png", b"hello")]:
        p = str(tmp_path / name)
        open(p, "wb").write(blob)
        with pytest.raises(ValueError):
            image_io.read_frame(p)
        with pytest.raises(ValueError):
            host_read(p)
