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
    assert lib.lr_api_version() == K.LR_API_VERSION == 4


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


def test_buffer_validation_at_the_python_boundary():
    """backend._ptr: the C side moves a fixed element count of fp32 through every pointer (lr_grid_update: grid floats,
    lr_render_backward: W*H*channels, ...), so dtype, contiguity and size are checked before the pointer is taken."""
    import numpy as np
    import torch
    from liverrenderer_b200.backend import _ptr
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    p, dev, keep = _ptr(a, 24, "x"); assert dev == 0 and keep is a and p.value == a.ctypes.data
    # inputs: converted to a contiguous float32 copy, values kept
    for src in (a.astype(np.float64), a.transpose(2, 1, 0), a[:, :, ::1].astype(np.int32), torch.from_numpy(a).double(), torch.from_numpy(a).permute(2, 0, 1)):
        p, dev, keep = _ptr(src, 24, "x", convert=True)
        k = keep if isinstance(keep, np.ndarray) else keep.numpy()
        assert k.dtype == np.float32 and k.flags.c_contiguous and np.array_equal(k.ravel(), np.asarray(src, dtype=np.float32).ravel() if isinstance(src, np.ndarray) else src.float().contiguous().numpy().ravel())
    # outputs are never converted
    with pytest.raises(TypeError, match="float32"):
        _ptr(a.astype(np.float64), 24, "out")
    with pytest.raises(ValueError, match="contiguous"):
        _ptr(a.transpose(2, 1, 0), 24, "out")
    with pytest.raises(TypeError, match="float32"):
        _ptr(torch.zeros(24, dtype=torch.float16), 24, "out")
    with pytest.raises(ValueError, match="contiguous"):
        _ptr(torch.zeros(4, 6).t(), 24, "out")
    # a short or long buffer is refused whatever its role
    for bad in (a[:1], np.zeros(25, np.float32), torch.zeros(23)):
        with pytest.raises(ValueError, match="expected 24 elements"):
            _ptr(bad, 24, "x", convert=True)
    with pytest.raises(TypeError, match="numpy array or a torch tensor"):
        _ptr(object(), 24, "x")


def test_plugin_stub_is_the_documented_one_and_parses_against_the_reference_headers():
    """integration/b200_volpath.cpp (the reference-side binding of INTEGRATION.md section 2) uses this tree's macros and parses against
    the reference's headers: with SURVEY 8(c)'s recipe (config.h for scalar_rgb, prelude, fwd.h overlay -- oracle/ref_kat/Makefile
    makes them) g++ -fsyntax-only reports no diagnostic in the stub or in lr_api.h.  The fork's own headers do not all compile as
    committed (render/endpoint.h:393, render/mesh.h:678); errors located there are the reference's."""
    import re, subprocess
    stub = os.path.join(ROOT, "integration", "b200_volpath.cpp"); src = open(stub).read()
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read(); i = md.index("```cpp\n") + 7
    assert md[i:md.index("```", i)] == src
    assert "MI_DECLARE_CLASS(B200Volpath)" in src and "MI_EXPORT_PLUGIN(B200Volpath)" in src and "MI_IMPLEMENT_CLASS_VARIANT" not in src
    R = "/root/reference"
    if not os.path.isdir(R):
        pytest.skip("/root/reference is not mounted here")
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle", "ref_kat")])
    B = os.path.join(ROOT, "oracle", "_ref", "build")
    inc = [B + "/overlay", R + "/include", B + "/include", R + "/ext/drjit/include", R + "/ext/drjit/ext/drjit-core/include",
           R + "/ext/drjit/ext/drjit-core/ext/nanothread/include", R + "/ext/drjit/ext/drjit-core/ext/robin_map/include", R + "/ext/tinyformat",
           R + "/ext/drjit/ext/nanobind/include", os.path.join(ROOT, "include")]
    p = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-w", "-include", B + "/prelude.h"] + ["-I" + d for d in inc] + [stub], capture_output=True, text=True)
    where = set(re.findall(r"^(\S+?):\d+:\d+: (?:fatal )?error", p.stderr, flags=re.M))
    assert not any(f.endswith(("b200_volpath.cpp", "lr_api.h")) for f in where), p.stderr[-3000:]
    assert where <= {R + "/include/mitsuba/render/endpoint.h", R + "/include/mitsuba/render/mesh.h"}, where
