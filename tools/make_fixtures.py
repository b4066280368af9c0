#!/usr/bin/env python
"""Generate the committed fixtures from the reference tree (run in the build container only;
/root/reference does not exist on the GPU box).

Outputs
  liverrenderer_b200/data/liver_meshes.npz   geometry of scenes/Liver-*/mitsuba3/liver{1,2}.obj (input data of
                                             BASELINE configs C3/C4), parsed with this repo's OBJ loader
  liverrenderer_b200/data/cavity_env_256x128.npz  cavidade_latitude.exr box-downsampled 4x to fp16 (the loader
                                             upsamples it back to 1024x512; synthetic stand-in, stated in bench.py)
  tests/golden/participating_media.npz       the reference's own golden mean/variance images
                                             (resources/data/tests/scenes/participating_media/refs/*_ref_rgb{,_var}.exr)
                                             + their inputs (sigmat.vol, albedo.vol, meshes, museum.exr)
  tests/golden/textures.npz                  decoded pixels of the images src/textures/tests/test_bitmap.py evaluates
                                             (carrot.png, noise_02.png, noise_8x8.png; integer data as stored in the files)
"""
import os
import sys
import numpy as np

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from liverrenderer_b200 import shapes as S  # noqa: E402
from liverrenderer_b200.transforms import ScalarTransform4f  # noqa: E402
from liverrenderer_b200.volume_io import read_vol  # noqa: E402

REF = "/root/reference"


def exr_rgb(path):
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    assert img is not None, path
    return np.ascontiguousarray(img[..., 2::-1].astype(np.float32))


def mesh_arrays(path):
    m = S.load_obj(path, ScalarTransform4f(), face_normals=False)
    d = {"positions": m.positions, "indices": m.indices}
    if m.normals is not None:
        d["normals"] = m.normals
    if m.texcoords is not None:
        d["texcoords"] = m.texcoords
    return d


def make_textures():
    tex = f"{REF}/resources/data/common/textures"; t = {}
    for name in ("carrot", "noise_02", "noise_8x8"):
        img = cv2.imread(f"{tex}/{name}.png", cv2.IMREAD_UNCHANGED)
        if img.ndim == 3 and img.shape[2] >= 3:
            img = np.concatenate([img[..., 2::-1], img[..., 3:]], axis=2)      # BGR(A) -> RGB(A)
        t[name] = np.ascontiguousarray(img)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "textures.npz"), **t)
    print("tests/golden/textures.npz", os.path.getsize(os.path.join(ROOT, "tests", "golden", "textures.npz")) // 1024, "KiB")


def main():
    os.makedirs(os.path.join(ROOT, "liverrenderer_b200", "data"), exist_ok=True)
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)

    # ---- workload geometry (C3 / C4) ----
    out = {}
    for name in ("liver1", "liver2"):
        for k, v in mesh_arrays(f"{REF}/scenes/Liver-SingleMesh/mitsuba3/{name}.obj").items():
            out[f"{name}_{k}"] = v
    np.savez_compressed(os.path.join(ROOT, "liverrenderer_b200", "data", "liver_meshes.npz"), **out)

    env = exr_rgb(f"{REF}/scenes/Liver-SingleMesh/mitsuba3/cavidade_latitude.exr")
    H, W, _ = env.shape
    assert (W, H) == (1024, 512), env.shape
    small = env.reshape(H // 4, 4, W // 4, 4, 3).mean(axis=(1, 3))
    np.savez_compressed(os.path.join(ROOT, "liverrenderer_b200", "data", "cavity_env_256x128.npz"),
                        rgb=np.minimum(small, 60000.0).astype(np.float16))

    # ---- golden fixtures of the reference's own render tests ----
    pm = f"{REF}/resources/data/tests/scenes/participating_media"
    g = {}
    for scene in ("test_hetero_sigmat", "test_hetero_sigmat_albedo", "test_homo_null_envmap", "test_homo_dielectric_envmap"):
        g[f"{scene}_ref"] = exr_rgb(f"{pm}/refs/{scene}_ref_rgb.exr")
        g[f"{scene}_var"] = exr_rgb(f"{pm}/refs/{scene}_ref_rgb_var.exr")
    g["sigmat_vol"] = read_vol(f"{pm}/textures/sigmat.vol")
    g["albedo_vol"] = read_vol(f"{pm}/textures/albedo.vol")
    for mesh in ("background", "unit_cube_centered", "sphere_base", "sphere_tiny"):
        for k, v in mesh_arrays(f"{pm}/meshes/{mesh}.obj").items():
            g[f"mesh_{mesh}_{k}"] = v
    g["museum"] = exr_rgb(f"{REF}/resources/data/common/textures/museum.exr")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "participating_media.npz"), **g)
    make_textures()
    for f in ("liverrenderer_b200/data/liver_meshes.npz", "liverrenderer_b200/data/cavity_env_256x128.npz",
              "tests/golden/participating_media.npz"):
        print(f, os.path.getsize(os.path.join(ROOT, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
