"""CPU tests (no GPU): the C-ABI library loads, exports every symbol include/lr_api.h declares, the ctypes
mirror of the POD structs matches the header's layout, and the product path fails loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "lr_api.h")


def _declared_functions():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lr_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = K.load_library()
    names = _declared_functions()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), f"liblr_b200.so does not export {n}"
    assert lib.lr_api_version() == 2


def test_no_torch_or_cxx_types_in_header():
    src = open(HDR).read()
    assert 'extern "C"' in src
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)          # declarations only, not the comments
    for banned in ("torch", "at::", "std::", "Tensor", "template"):
        assert banned not in src


def test_ctypes_struct_layout_matches_header():
    """Compile a tiny C program against include/lr_api.h that prints sizeof() of every POD struct."""
    structs = ["lr_mesh", "lr_bsdf", "lr_grid", "lr_medium", "lr_emitter", "lr_sensor", "lr_integrator", "lr_scene", "lr_stats", "lr_texture"]
    body = "".join(f'printf("{s} %zu\\n", sizeof({s}));' for s in structs)
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "sz.c"); exe = os.path.join(td, "sz")
        open(c, "w").write(f'#include <stdio.h>\n#include "lr_api.h"\nint main(void){{{body} return 0;}}\n')
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        out = subprocess.check_output([exe]).decode().split()
    sizes = dict(zip(out[0::2], map(int, out[1::2])))
    for s in structs:
        assert C.sizeof(getattr(K, s)) == sizes[s], (s, C.sizeof(getattr(K, s)), sizes[s])


def test_product_fails_loudly_without_gpu():
    """There is no CPU fallback: without a CUDA device lr_create returns LR_ERR_NO_DEVICE and the host raises."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CUDA device|CPU"):
        mi.Context(0)
    with pytest.raises(RuntimeError):
        mi.render(mi.load_dict(mi.cornell_box()), spp=1)


def test_missing_library_raises(tmp_path):
    with pytest.raises(RuntimeError, match="not found"):
        K.load_library(str(tmp_path / "nope.so"))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "liverrenderer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liborc" not in txt and "import orc" not in txt and "tests.orc" not in txt and "from tests" not in txt, f


def test_unsupported_plugins_raise_like_the_reference():
    d = mi.cornell_box(); d["floor"]["bsdf"] = {"type": "plastic"}
    with pytest.raises(NotImplementedError):
        mi.load_dict(d)
    d = mi.cornell_box(); d["floor"]["bsdf"] = {"type": "roughdielectric", "int_ior": 1.5, "ext_ior": 1.5}
    with pytest.raises(RuntimeError, match="indices of refraction"):
        mi.load_dict(d)


def test_ldsampler_host_rounding_and_oracle_image():
    """`ldsampler` sample counts are rounded up to a square power of two like LowDiscrepancySampler::set_sample_count
    (src/samplers/ldsampler.cpp:80-89); the oracle renders with it (CPU, tiny)."""
    from liverrenderer_b200.scene import ld_sample_count
    assert [ld_sample_count(n) for n in (1, 4, 5, 16, 17, 64, 100, 1024, 4096)] == [4, 4, 16, 16, 64, 64, 256, 1024, 4096]
    d = mi.cornell_box(); d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = 16
    d["sensor"]["sampler"] = {"type": "ldsampler", "sample_count": 12}
    sc = mi.load_dict(d)
    assert sc.spp == 16 and sc.integrator.sampler == 1
    from tests import orc
    o = orc.OracleScene(sc); a = o.sample_lanes(0, 16, 0, 16 * 16 * 16)
    d["sensor"]["sampler"] = {"type": "independent", "sample_count": 16}
    b = orc.OracleScene(mi.load_dict(d)).sample_lanes(0, 16, 0, 16 * 16 * 16)
    assert np.isfinite(a).all() and not np.array_equal(a, b)
    assert abs(a[:, :3].mean() - b[:, :3].mean()) / b[:, :3].mean() < 0.15      # same integrand, different point sets
    d["sensor"]["sampler"] = {"type": "stratified"}
    with pytest.raises(NotImplementedError):
        mi.load_dict(d)
