"""Host-side image I/O: `envmap` / `bitmap` inputs (EXR/HDR/PNG/JPG via OpenCV, .npy/.npz arrays) and
`write_bitmap` for results (mi.util.write_bitmap, src/python/python/util.py:528-561), so a render can leave the
process without the reference's host library."""
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
    if img.ndim == 2:
        img = np.repeat(img[..., None], 3, axis=2)
    img = img[..., 2::-1]                            # BGR(A) -> RGB
    if img.dtype == np.uint8 or img.dtype == np.uint16:
        # 8/16-bit files hold sRGB-encoded values in 0..255 / 0..65535: normalise and linearise as the reference's Bitmap
        # conversion does for every non-raw integer image (src/emitters/envmap.cpp:95-107, src/textures/bitmap.cpp:295-345)
        img = srgb_to_linear(img.astype(np.float32) / np.float32(255.0 if img.dtype == np.uint8 else 65535.0))
    return np.ascontiguousarray(img, dtype=np.float32)


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


def linear_to_srgb(v: np.ndarray) -> np.ndarray:
    """Forward sRGB gamma (include/mitsuba/core/math.h linear_to_srgb), fp32."""
    v = np.asarray(v, dtype=np.float32)
    lo = v * np.float32(12.92)
    hi = np.float32(1.055) * np.power(np.maximum(v, 0), np.float32(1.0 / 2.4), dtype=np.float32) - np.float32(0.055)
    return np.where(v <= np.float32(0.0031308), lo, hi).astype(np.float32)


def write_bitmap(filename: str, data) -> None:
    """`mi.util.write_bitmap`: .png/.jpg/.jpeg => 8-bit sRGB (clamped, rounded as Bitmap::convert does: x*255 + 0.5),
    anything else OpenCV can write (.exr, .hdr) => linear float32.  `data` is [H, W, 1|3|4] (or [H, W])."""
    img = np.asarray(data.detach().cpu().numpy() if hasattr(data, "detach") else data, dtype=np.float32)
    if img.ndim == 2:
        img = img[..., None]
    if img.ndim != 3 or img.shape[2] not in (1, 3, 4):
        raise RuntimeError(f"write_bitmap(): expected an [H, W, 1|3|4] image, got {img.shape}")
    ext = os.path.splitext(filename)[1].lower()
    if ext in (".npy",):
        np.save(filename, img); return
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    if ext in (".png", ".jpg", ".jpeg"):
        rgb = linear_to_srgb(np.clip(img[..., :3], 0.0, 1.0)) if img.shape[2] >= 3 else linear_to_srgb(np.clip(img, 0.0, 1.0))
        out = np.concatenate([rgb, np.clip(img[..., 3:], 0.0, 1.0)], axis=2) if img.shape[2] == 4 else rgb
        out = np.floor(out * np.float32(255.0) + np.float32(0.5)).astype(np.uint8)
    else:
        out = img
    if out.shape[2] >= 3:
        out = np.concatenate([out[..., 2::-1], out[..., 3:]], axis=2)          # RGB(A) -> BGR(A)
    if not cv2.imwrite(filename, np.ascontiguousarray(out)):
        raise RuntimeError(f'could not write image "{filename}"')
