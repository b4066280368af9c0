"""Mitsuba `.vol` (version 3) grid-volume files.

Reference: src/render/volumegrid.cpp:28-84 (read), :98-127 (write).  Layout: 'V','O','L',
uint8 version=3, int32 type (1 = float32), int32 xres, yres, zres, channels, 6 x float32
bbox, then fp32 data in [z][y][x][channel] order (little endian).
"""
from __future__ import annotations

import struct
import numpy as np


def read_vol(path: str) -> np.ndarray:
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:3] != b"VOL":
        raise RuntimeError(f'Invalid volume file "{path}": bad magic')
    if raw[3] != 3:
        raise RuntimeError(f'Invalid volume file "{path}": only version 3 is supported')
    dtype, xres, yres, zres, channels = struct.unpack_from("<5i", raw, 4)
    if dtype != 1:
        raise RuntimeError(f'Invalid volume file "{path}": only float32 data is supported')
    n = xres * yres * zres * channels
    data = np.frombuffer(raw, dtype="<f4", count=n, offset=48)
    return data.reshape(zres, yres, xres, channels).astype(np.float32, copy=True)


def read_vol_bbox(path: str):
    with open(path, "rb") as f:
        raw = f.read(48)
    return struct.unpack_from("<6f", raw, 24)


def write_vol(path: str, data: np.ndarray, bbox=(0, 0, 0, 1, 1, 1)) -> None:
    data = np.asarray(data, dtype=np.float32)
    if data.ndim == 3:
        data = data[..., None]
    zres, yres, xres, channels = data.shape
    with open(path, "wb") as f:
        f.write(b"VOL" + bytes([3]))
        f.write(struct.pack("<5i", 1, xres, yres, zres, channels))
        f.write(struct.pack("<6f", *bbox))
        f.write(np.ascontiguousarray(data, dtype="<f4").tobytes())
