"""Host-side encoders of the headless output path (hiprt-path-tracer_b200/image_io.py): PNG and Radiance HDR round trips."""
import importlib
import os

import numpy as np

io = importlib.import_module("hiprt-path-tracer_b200.image_io")


def test_png_round_trip(tmp_path):
    rng = np.random.default_rng(1)
    for shape in ((1, 1, 3), (7, 13, 3), (96, 128, 3)):
        img = rng.integers(0, 256, size=shape, dtype=np.uint8)
        path = os.path.join(tmp_path, "a.png")
        io.write_png(path, img)
        assert np.array_equal(io.read_png(path), img)
        io.write_png(path, img, flip_y=True)
        assert np.array_equal(io.read_png(path), img[::-1])


def test_rgbe_round_trip_is_within_one_mantissa_step():
    rng = np.random.default_rng(2)
    img = (rng.random((32, 48, 3), dtype=np.float32) * np.float32(10.0) ** rng.integers(-4, 5, size=(32, 48, 1))).astype(np.float32)
    img[0, 0] = 0.0
    back = io.rgbe_to_float(io.float_to_rgbe(img))
    step = np.ldexp(1.0, np.frexp(img.max(axis=-1))[1] - 8)[..., None]        # one unit of the shared 8-bit mantissa
    assert (np.abs(back - img) <= step + 1e-30).all()
    assert not back[0, 0].any()


def test_hdr_and_pfm_headers(tmp_path):
    img = np.ones((4, 6, 3), np.float32)
    p = os.path.join(tmp_path, "a.hdr")
    io.write_hdr(p, img)
    data = open(p, "rb").read()
    assert data.startswith(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 4 +X 6\n") and len(data) == len(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 4 +X 6\n") + 4 * 6 * 4
    p = os.path.join(tmp_path, "a.pfm")
    io.write_pfm(p, img)
    assert open(p, "rb").read().startswith(b"PF\n6 4\n-1.0\n")


def test_heatmap_ends_and_middle():
    v = np.array([[0.0, 5.0, 10.0, 20.0]])
    h = io.heatmap(v, 0.0, 10.0)
    assert tuple(h[0, 0]) == (0, 0, 255) and tuple(h[0, 1]) == (0, 255, 0) and tuple(h[0, 2]) == (255, 0, 0) and tuple(h[0, 3]) == (255, 0, 0)
