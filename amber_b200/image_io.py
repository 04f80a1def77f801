"""Headless image output for the path: the step right after ``Algorithm::Render`` in the
reference (``src/amber/cli/application.cc:98-115``), which writes ``<output>.png`` (Filmic +
Gamma, 8 bit) and ``<output>.exr`` (raw floats) through OpenCV with the columns mirrored
(``src/amber/cli/image.cc:45-71``: ``mat.at(j, Width - i - 1)``; OpenCV's BGR order is its
in-memory convention -- the files hold RGB).  OpenCV is absent here, so this is a dependency-free
PNG writer (zlib from the standard library) and PFM for the raw image.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np


def write_png(path: str, rgb8: np.ndarray) -> None:
    """8-bit RGB (H, W, 3) -> PNG (colour type 2, no interlace)."""
    rgb8 = np.ascontiguousarray(rgb8, dtype=np.uint8)
    h, w, c = rgb8.shape
    assert c == 3
    raw = np.empty((h, 1 + 3 * w), dtype=np.uint8)
    raw[:, 0] = 0  # filter type 0 on every scanline
    raw[:, 1:] = rgb8.reshape(h, 3 * w)

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)))
        f.write(chunk(b"IEND", b""))


def read_png(path: str) -> np.ndarray:
    """Inverse of write_png (filter 0, 8-bit RGB only): for the tests."""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h = struct.unpack(">II", body[:8])
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3).copy()


def write_pfm(path: str, rgb: np.ndarray) -> None:
    """float32 RGB (H, W, 3) -> PFM (little endian; PFM stores rows bottom-up)."""
    rgb = np.ascontiguousarray(rgb, dtype="<f4")
    h, w, c = rgb.shape
    assert c == 3
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(rgb[::-1].tobytes())


def read_pfm(path: str) -> np.ndarray:
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)


def export(device, output: str) -> tuple[str, str]:
    """What ``Application::Run`` does after rendering (``application.cc:98-115``): the tonemapped
    8-bit image and the raw image, both with the columns mirrored as the reference's exporters
    do.  ``device`` is an ``amber_b200.render.Device`` holding a rendered film."""
    ldr, _ = device.resolve_ldr(flip_x=True)
    raw, _ = device.resolve()
    write_png(output + ".png", ldr)
    write_pfm(output + ".pfm", raw[:, ::-1])
    return output + ".png", output + ".pfm"
