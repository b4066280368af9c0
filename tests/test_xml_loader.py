"""`mi.load_file` (src/core/xml.cpp subset): an XML scene written the way the fork's scene files are
(scenes/Liver-SingleMesh/mitsuba3/scene.xml: <default>, $parameters, <ref>, nested <bsdf type="bumpmap">, <transform> blocks
applied in file order, gridvolume / envmap / obj files next to the XML) must give the same POD as the equivalent dictionary."""
import ctypes as C
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K
from liverrenderer_b200.transforms import T
from liverrenderer_b200.xml_loader import xml_to_dict

XML = """<scene version="3.0.0">
    <default name="spp" value="16"/>
    <default name="res" value="32"/>
    <integrator type="volpath">
        <integer name="max_depth" value="12"/>
        <integer name="rr_depth" value="5"/>
    </integrator>
    <sensor type="perspective">
        <float name="fov" value="45"/>
        <transform name="to_world">
            <lookat origin="0, 0.5, 4" target="0, 0, 0" up="0, 1, 0"/>
        </transform>
        <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>
        <film type="hdrfilm">
            <integer name="width" value="$res"/><integer name="height" value="$res"/>
            <rfilter type="box"/>
            <string name="pixel_format" value="rgba"/>
        </film>
    </sensor>
    <bsdf type="bumpmap" id="CapsuleBSDF">
        <float name="scale" value="0.005"/>
        <texture name="texture" type="bitmap">
            <string name="filename" value="tissue.npy"/>
            <boolean name="raw" value="true"/>
        </texture>
        <bsdf type="dielectric">
            <float name="int_ior" value="1.38"/>
            <float name="ext_ior" value="1"/>
        </bsdf>
    </bsdf>
    <medium type="heterogeneous" id="parenchyma">
        <float name="scale" value="2.5"/>
        <volume name="sigma_t" type="gridvolume">
            <string name="filename" value="sigmat.vol"/>
            <transform name="to_world"><translate x="-0.5" y="-0.5" z="-0.5"/><scale value="2"/></transform>
        </volume>
        <rgb name="albedo" value="0.92, 0.45, 0.35"/>
        <phase type="hg"><float name="g" value="0.8"/></phase>
    </medium>
    <shape type="obj">
        <string name="filename" value="quad.obj"/>
        <transform name="to_world">
            <rotate x="1" angle="-90"/>
            <scale x="3" y="1" z="3"/>
            <translate x="0" y="-1" z="0"/>
        </transform>
        <bsdf type="diffuse">
            <texture name="reflectance" type="checkerboard">
                <rgb name="color0" value="0.325, 0.31, 0.25"/>
                <rgb name="color1" value="0.725"/>
                <transform name="to_uv"><scale x="10" y="10"/></transform>
            </texture>
        </bsdf>
    </shape>
    <shape type="cube">
        <ref id="CapsuleBSDF"/>
        <ref name="interior" id="parenchyma"/>
    </shape>
    <emitter type="constant"><rgb name="radiance" value="1.0"/></emitter>
</scene>
"""
OBJ = "v -1 -1 0\nv 1 -1 0\nv 1 1 0\nv -1 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 0 1\nf 1/1/1 2/2/1 3/3/1\nf 1/1/1 3/3/1 4/4/1\n"


def _write_files(d):
    rng = np.random.default_rng(0)
    (d / "scene.xml").write_text(XML); (d / "quad.obj").write_text(OBJ)
    np.save(d / "tissue.npy", rng.random((16, 16, 3)).astype(np.float32))
    mi.write_vol(str(d / "sigmat.vol"), rng.random((4, 5, 6)).astype(np.float32))


def _same_struct(a, b):
    return bytes(a) == bytes(b)


def test_xml_equals_dict(tmp_path):
    _write_files(tmp_path)
    sx = mi.load_file(str(tmp_path / "scene.xml"))
    d = {"type": "scene",
         "integrator": {"type": "volpath", "max_depth": 12, "rr_depth": 5},
         "sensor": {"type": "perspective", "fov": 45.0, "to_world": T().look_at(origin=[0, 0.5, 4], target=[0, 0, 0], up=[0, 1, 0]),
                    "sampler": {"type": "independent", "sample_count": 16},
                    "film": {"type": "hdrfilm", "width": 32, "height": 32, "rfilter": {"type": "box"}, "pixel_format": "rgba"}},
         "CapsuleBSDF": {"type": "bumpmap", "scale": 0.005, "texture": {"type": "bitmap", "filename": "tissue.npy", "raw": True},
                         "bsdf": {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0}},
         "parenchyma": {"type": "heterogeneous", "scale": 2.5,
                        "sigma_t": {"type": "gridvolume", "filename": "sigmat.vol", "to_world": T().scale(2).translate([-0.5, -0.5, -0.5])},
                        "albedo": {"type": "rgb", "value": [0.92, 0.45, 0.35]}, "phase": {"type": "hg", "g": 0.8}},
         # XML order rotate -> scale -> translate == matrix T * S * R
         "floor": {"type": "obj", "filename": "quad.obj", "to_world": T().translate([0, -1, 0]).scale([3, 1, 3]).rotate([1, 0, 0], -90),
                   "bsdf": {"type": "diffuse", "reflectance": {"type": "checkerboard", "color0": [0.325, 0.31, 0.25], "color1": 0.725,
                                                               "to_uv": T().scale([10, 10, 1])}}},
         "organ": {"type": "cube", "bsdf": {"type": "ref", "id": "CapsuleBSDF"}, "interior": {"type": "ref", "id": "parenchyma"}},
         "env": {"type": "constant", "radiance": {"type": "rgb", "value": 1.0}}}
    sd = mi.load_dict(d, base_dir=str(tmp_path))
    assert sx.spp == sd.spp == 16 and sx.film_size == (32, 32) and sx.out_channels == 4
    assert _same_struct(sx.sensor, sd.sensor) and _same_struct(sx.integrator, sd.integrator)
    assert len(sx.bsdfs) == len(sd.bsdfs) and all(_same_struct(a, b) for a, b in zip(sx.bsdfs, sd.bsdfs))
    assert len(sx.media) == len(sd.media) == 1 and _same_struct(sx.media[0], sd.media[0])
    assert len(sx.meshes) == len(sd.meshes) == 2
    for a, b in zip(sx.meshes, sd.meshes):
        assert (a.n_vertices, a.n_faces, a.bsdf, a.interior_medium, a.exterior_medium) == (b.n_vertices, b.n_faces, b.bsdf, b.interior_medium, b.exterior_medium)
        np.testing.assert_array_equal(np.ctypeslib.as_array(a.positions, (a.n_vertices * 3,)), np.ctypeslib.as_array(b.positions, (b.n_vertices * 3,)))
    # the floor ended up at y = -1, spanning +-3 in x and z (rotate, then scale, then translate)
    p = np.ctypeslib.as_array(sx.meshes[0].positions, (4, 3))
    np.testing.assert_allclose(sorted(map(tuple, np.round(p, 5))), sorted([(-3, -1, -3), (3, -1, -3), (3, -1, 3), (-3, -1, 3)]), atol=1e-5)
    gx, gd = sx.grids[0], sd.grids[0]
    np.testing.assert_array_equal(np.asarray(gx.to_local), np.asarray(gd.to_local))
    assert np.array_equal(np.asarray(sx.grid_arrays[0]), np.asarray(sd.grid_arrays[0]))
    bump = sx.bsdfs[sx.meshes[1].bsdf]
    assert bump.kind == K.BSDF_DIELECTRIC and bump.bump_tex == 1 and sx.textures[0].width == 16


def test_default_parameters_and_errors(tmp_path):
    _write_files(tmp_path)
    sc = mi.load_file(str(tmp_path / "scene.xml"), spp=64, res=8)
    assert sc.spp == 64 and sc.film_size == (8, 8)
    d = xml_to_dict(str(tmp_path / "scene.xml"))
    assert d["sensor"]["sampler"]["sample_count"] == 16 and d["CapsuleBSDF"]["type"] == "bumpmap"
    (tmp_path / "bad.xml").write_text(XML.replace('<default name="spp" value="16"/>', ""))
    with pytest.raises(RuntimeError, match=r"undefined parameter"):
        mi.load_file(str(tmp_path / "bad.xml"))
    (tmp_path / "bio.xml").write_text(XML.replace('<integrator type="volpath">', '<integrator type="biovolpath06">'))
    with pytest.raises(NotImplementedError, match="biovolpath06"):
        mi.load_file(str(tmp_path / "bio.xml"))
    # <include>: the shared part of a scene in its own file (resources/data/tests/scenes/participating_media/include/*.xml)
    head, tail = XML.split("<emitter", 1)
    (tmp_path / "inc.xml").write_text('<scene version="3.0.0"><emitter' + tail)
    (tmp_path / "main.xml").write_text(head + '<include filename="inc.xml"/></scene>')
    sc2 = mi.load_file(str(tmp_path / "main.xml"))
    assert len(sc2.emitters) == 1 and len(sc2.meshes) == 2


def test_unknown_top_level_plugin_raises():
    """A plugin the path does not implement must not silently drop out of the render (the reference's PluginManager raises
    on an unknown plugin name, src/core/plugin.cpp): a point light, a ply shape, a plastic BSDF ..."""
    base = {"type": "scene", "integrator": {"type": "path"},
            "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
            "env": {"type": "constant", "radiance": {"type": "rgb", "value": 1.0}}}
    mi.load_dict(dict(base))
    for extra in ({"type": "point", "intensity": 1.0}, {"type": "ply", "filename": "x.ply"}, {"type": "plastic"}):
        with pytest.raises(NotImplementedError, match=extra["type"]):
            mi.load_dict(dict(base, thing=extra))


def test_scene_with_two_sensors_and_wide_filter():
    """Scene::sensors(): every `perspective` of the scene is kept, in order; render(scene, sensor=i) picks one."""
    from liverrenderer_b200 import scenes
    d = scenes.cornell_box_path(res=16, spp=4)
    d["sensor2"] = dict(d["sensor"]); d["sensor2"]["film"] = {"type": "hdrfilm", "width": 24, "height": 12, "rfilter": {"type": "gaussian", "stddev": 1.0}}
    d["sensor2"]["fov"] = 25.0
    sc = mi.load_dict(d)
    assert len(sc.sensors) == 2 and sc.film_size == (16, 16) and sc.sensor_index == 0
    sc.select_sensor(1)
    assert sc.film_size == (24, 12) and abs(sc.sensor.rfilter_radius - 4.0) < 1e-6
    with pytest.raises(RuntimeError, match="out of bounds"):
        sc.select_sensor(2)
    d["sensor2"]["sampler"] = {"type": "ldsampler", "sample_count": 4}
    with pytest.raises(NotImplementedError, match="share the sampler"):
        mi.load_dict(d)
