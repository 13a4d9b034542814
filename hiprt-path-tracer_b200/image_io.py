"""Headless image output: the files the reference writes through stb_image_write (src/Image/Image.cpp:427-452,
write_image_png / write_image_hdr) written with the Python standard library -- no window, no GL interop.

  write_png(path, rgb8)       8-bit RGB PNG (zlib-deflated scanlines, filter 0)
  write_hdr(path, rgb32f)     Radiance RGBE, uncompressed scanlines ('#?RADIANCE', FORMAT=32-bit_rle_rgbe)
  write_pfm(path, rgb32f)     raw float32 (lossless: what parity tools diff)
  heatmap(values, lo, hi)     the adaptive-sampling heat map as DATA: the piecewise-linear colour-stop interpolation of
                              src/Shaders/heatmap_int.frag (its stops are uniforms set by the UI; a 5-stop blue-to-red ramp here;
                              the shader itself: unavailable on this box, no graphics stack)
Host-side encoding only; the arithmetic that produces the pixels (tonemap) runs on the device (b200pt_tonemap)."""
import struct
import zlib

import numpy as np


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path, rgb8, flip_y=False):
    img = np.ascontiguousarray(rgb8, dtype=np.uint8)
    if img.ndim != 3 or img.shape[2] != 3:
        raise ValueError("write_png: expected [H, W, 3] uint8")
    if flip_y:
        img = img[::-1]
    h, w, _ = img.shape
    raw = np.empty((h, 1 + w * 3), np.uint8)
    raw[:, 0] = 0
    raw[:, 1:] = img.reshape(h, w * 3)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)))
        f.write(_chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)))
        f.write(_chunk(b"IEND", b""))


def read_png(path):
    """Decoder for the files write_png makes (8-bit RGB, filter 0): used by the tests."""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I", data[pos:pos + 4])[0], data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h = struct.unpack(">II", body[:8])
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * 3)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, 3).copy()


def float_to_rgbe(rgb32f):
    """Greg Ward's RGBE: shared exponent of the largest channel (what stbi_write_hdr stores per pixel)."""
    rgb = np.ascontiguousarray(rgb32f, dtype=np.float32)
    m = rgb.max(axis=-1)
    mant, exp = np.frexp(m)
    scale = np.where(m > 1e-32, mant * 256.0 / np.maximum(m, 1e-38), 0.0).astype(np.float32)
    out = np.zeros(rgb.shape[:-1] + (4,), np.uint8)
    out[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.where(m > 1e-32, exp + 128, 0).astype(np.uint8)
    return out


def rgbe_to_float(rgbe):
    e = rgbe[..., 3].astype(np.int32)
    scale = np.where(e > 0, np.ldexp(1.0, e - 136), 0.0).astype(np.float32)
    return rgbe[..., :3].astype(np.float32) * scale[..., None]


def write_hdr(path, rgb32f, flip_y=False):
    img = np.asarray(rgb32f, dtype=np.float32)
    if flip_y:
        img = img[::-1]
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {h} +X {w}\n".encode())
        f.write(float_to_rgbe(img).tobytes())


def write_pfm(path, rgb32f):
    img = np.ascontiguousarray(rgb32f, dtype="<f4")
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(img[::-1].tobytes())          # PFM stores the bottom row first


def heatmap(values, lo, hi):
    """[H, W] numbers -> [H, W, 3] uint8 on the blue-cyan-green-yellow-red ramp (5 stops, linear), lo maps to blue, hi to red."""
    t = np.clip((np.asarray(values, np.float32) - lo) / max(float(hi - lo), 1e-20), 0.0, 1.0) * 4.0
    stops = np.array([[0, 0, 1], [0, 1, 1], [0, 1, 0], [1, 1, 0], [1, 0, 0]], np.float32)
    i = np.minimum(t.astype(np.int32), 3)
    f = (t - i)[..., None]
    return np.clip((stops[i] * (1 - f) + stops[i + 1] * f) * 255.0, 0, 255).astype(np.uint8)
