"""Host-side image input for `envmap` bitmaps (EXR/HDR via OpenCV, .npy/.npz arrays).
Image *output* stays with the host application; the backend returns float tensors."""
from __future__ import annotations

import os
import numpy as np


def read_image_rgb(path: str) -> np.ndarray:
    """Linear RGB float32 [H, W, 3]."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".npy":
        return np.load(path).astype(np.float32)[..., :3]
    if ext == ".npz":
        z = np.load(path)
        return z[z.files[0]].astype(np.float32)[..., :3]
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise RuntimeError(f'could not read image "{path}"')
    img = img.astype(np.float32)
    if img.ndim == 2:
        img = np.repeat(img[..., None], 3, axis=2)
    return np.ascontiguousarray(img[..., 2::-1])     # BGR(A) -> RGB


def srgb_to_linear(v: np.ndarray) -> np.ndarray:
    """Bitmap::convert's inverse sRGB gamma (include/mitsuba/core/math.h srgb_to_linear), fp32."""
    v = v.astype(np.float32)
    lo = v * np.float32(1.0 / 12.92)
    hi = np.power((v + np.float32(0.055)) * np.float32(1.0 / 1.055), np.float32(2.4), dtype=np.float32)
    return np.where(v <= np.float32(0.04045), lo, hi).astype(np.float32)


def texture_from_pixels(px: np.ndarray, raw: bool) -> np.ndarray:
    """What `bitmap` holds after load (src/textures/bitmap.cpp:295-345): Y[A] -> 1 channel, RGB[A] -> 3 channels,
    float32; integer formats are normalised and, unless `raw`, taken from sRGB to linear."""
    px = np.asarray(px)
    if px.ndim == 2:
        px = px[..., None]
    c = px.shape[2]
    px = px[..., :1] if c in (1, 2) else px[..., :3]
    if px.dtype == np.uint8 or px.dtype == np.uint16:
        v = px.astype(np.float32) / np.float32(255.0 if px.dtype == np.uint8 else 65535.0)
        px = v if raw else srgb_to_linear(v)
    return np.ascontiguousarray(px, dtype=np.float32)


def read_texture(path: str, raw: bool) -> np.ndarray:
    """Texture file -> float32 [H, W, 1|3] (PNG/JPG/EXR via OpenCV, .npy arrays as stored)."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".npy":
        return texture_from_pixels(np.load(path), raw)
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise RuntimeError(f'could not read image "{path}"')
    if img.ndim == 3 and img.shape[2] >= 3:
        img = np.concatenate([img[..., 2::-1], img[..., 3:]], axis=2)      # BGR(A) -> RGB(A)
    return texture_from_pixels(img, raw)
