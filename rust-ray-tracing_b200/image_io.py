"""Image output — what `RgbImage::save("output.png")` does for the reference (src/main.rs:66).  A minimal PNG
(zlib from the standard library) and binary PPM writer for [H,W,3] uint8 arrays; no third-party imaging library."""
from __future__ import annotations

import struct
import zlib

import numpy as np


def write_ppm(path: str, image: np.ndarray) -> None:
    h, w, _ = image.shape
    with open(path, "wb") as f:
        f.write(f"P6\n{w} {h}\n255\n".encode("ascii"))
        f.write(np.ascontiguousarray(image, dtype=np.uint8).tobytes())


def write_png(path: str, image: np.ndarray) -> None:
    image = np.ascontiguousarray(image, dtype=np.uint8)
    h, w, c = image.shape
    if c != 3:
        raise ValueError("expected an RGB image")
    raw = np.concatenate([np.zeros((h, 1), dtype=np.uint8), image.reshape(h, w * 3)], axis=1).tobytes()

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw, 6)))
        f.write(chunk(b"IEND", b""))
