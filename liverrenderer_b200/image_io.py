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
