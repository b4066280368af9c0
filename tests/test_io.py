"""Host-side file formats either side of the path (second half): Mitsuba `.vol` grids
(src/render/volumegrid.cpp:28-127) and image output (`mi.util.write_bitmap`, src/python/python/util.py:548-561)."""
import os
import struct
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200.volume_io import read_vol_bbox
from liverrenderer_b200.image_io import linear_to_srgb, srgb_to_linear
from tests import golden_scenes as G


def test_vol_header_layout_and_round_trip(tmp_path):
    """VolumeGrid::write (volumegrid.cpp:98-127): 'VOL', u8 3, i32 type 1, i32 xres yres zres channels, 6 x f32 bbox, data."""
    rng = np.random.default_rng(0); g = rng.random((3, 4, 5, 3)).astype(np.float32)       # [z][y][x][c]
    p = str(tmp_path / "g.vol"); mi.write_vol(p, g, bbox=(-1, -2, -3, 4, 5, 6))
    raw = open(p, "rb").read()
    assert raw[:4] == b"VOL\x03" and len(raw) == 48 + g.size * 4
    assert struct.unpack_from("<5i", raw, 4) == (1, 5, 4, 3, 3)
    assert struct.unpack_from("<6f", raw, 24) == (-1, -2, -3, 4, 5, 6) == tuple(read_vol_bbox(p))
    assert np.array_equal(np.frombuffer(raw, "<f4", offset=48), g.ravel())
    assert np.array_equal(mi.read_vol(p), g)
    mi.write_vol(p, g[..., 0]); assert mi.read_vol(p).shape == (3, 4, 5, 1)
    open(p, "wb").write(b"VOX" + raw[3:])
    with pytest.raises(RuntimeError, match="bad magic"):
        mi.read_vol(p)
    open(p, "wb").write(b"VOL\x02" + raw[4:])
    with pytest.raises(RuntimeError, match="version 3"):
        mi.read_vol(p)


def test_vol_reference_file_round_trip(tmp_path):
    """The reference's own sigmat.vol (fixture made by read_vol from the file) survives write -> read -> load_dict."""
    g = G.golden()["sigmat_vol"]; p = str(tmp_path / "sigmat.vol"); mi.write_vol(p, g)
    assert np.array_equal(mi.read_vol(p), g)
    d = G.scene("test_hetero_sigmat", spp=1); d["medium1"]["sigma_t"] = {"type": "gridvolume", "filename": p}
    sc = mi.load_dict(d)
    assert np.array_equal(np.asarray(sc.grid_arrays[0]).reshape(g.shape), g)


def test_write_bitmap_exr_and_png(tmp_path):
    rng = np.random.default_rng(1); img = (rng.random((6, 7, 3)) * 4).astype(np.float32)
    p = str(tmp_path / "o.exr"); mi.write_bitmap(p, img)
    from liverrenderer_b200.image_io import read_image_rgb
    np.testing.assert_array_equal(read_image_rgb(p), img)                 # float32 EXR is lossless, channel order kept
    q = str(tmp_path / "o.png"); mi.write_bitmap(q, img)
    back = mi.read_texture(q, raw=True)                                   # 8-bit values / 255, no gamma
    want = np.floor(linear_to_srgb(np.clip(img, 0, 1)) * 255 + 0.5) / 255
    np.testing.assert_allclose(back, want, atol=1e-6)
    # sRGB transfer pair: inverse of each other on [0, 1] and the two published anchors
    x = np.linspace(0, 1, 257, dtype=np.float32)
    np.testing.assert_allclose(srgb_to_linear(linear_to_srgb(x)), x, atol=2e-6)
    np.testing.assert_allclose(linear_to_srgb(np.float32([0.0031308, 0.5])), [0.04045, 0.735357], atol=2e-5)
    with pytest.raises(RuntimeError):
        mi.write_bitmap(str(tmp_path / "bad.exr"), np.zeros((4, 4, 2), np.float32))


def test_ldr_environment_image_is_normalised_and_linearised(tmp_path):
    """An 8-bit PNG / 16-bit PNG given to `envmap` is sRGB data in 0..255 (0..65535): read_image_rgb returns linear values in
    0..1 like the `bitmap` texture path (envmap.cpp:95-107 converts through Bitmap), not raw integers 255x too bright."""
    import cv2
    from liverrenderer_b200.image_io import read_image_rgb
    rng = np.random.default_rng(3); px8 = rng.integers(0, 256, size=(6, 9, 3), dtype=np.uint8)
    p8 = str(tmp_path / "env8.png"); cv2.imwrite(p8, px8[..., ::-1])
    got = read_image_rgb(p8)
    assert got.dtype == np.float32 and got.shape == (6, 9, 3) and got.max() <= 1.0
    assert np.allclose(got, srgb_to_linear(px8.astype(np.float32) / 255.0), atol=1e-7)
    assert np.array_equal(np.round(linear_to_srgb(got) * 255.0).astype(np.uint8), px8)           # uint8 round trip
    px16 = rng.integers(0, 65536, size=(4, 5, 3), dtype=np.uint16); p16 = str(tmp_path / "env16.png"); cv2.imwrite(p16, px16[..., ::-1])
    assert np.allclose(read_image_rgb(p16), srgb_to_linear(px16.astype(np.float32) / 65535.0), atol=1e-7)
    f = rng.random((3, 4, 3)).astype(np.float32) * 7.0; np.save(tmp_path / "env.npy", f)          # float data stays as stored
    assert np.array_equal(read_image_rgb(str(tmp_path / "env.npy")), f)
