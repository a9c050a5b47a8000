"""Host logic: SceneIO parser, GenerateAABBForObject, Camera, synthetic generators."""
import os

import numpy as np
import pytest

from opengl_raytracing_b200 import scene as S, synthetic

REF_SCENES = "/root/reference/res/Scene"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_sphere_and_plane_aabbs():
    o = S.new_objects(3)
    o[0]["type"], o[0]["position"], o[0]["radius"] = S.SPHERE, (1, 2, 3), 0.5
    o[1]["type"], o[1]["position"], o[1]["normal"], o[1]["size"] = S.PLANE, (0, -1, -5), (0, 1, 0), (10, 10)
    o[2]["type"], o[2]["position"], o[2]["normal"], o[2]["size"] = S.PLANE, (0, 0, -10), (0, 0, 1), (20, 20)
    S.generate_aabbs(o)
    assert np.allclose(o[0]["bounds"]["min"], (0.5, 1.5, 2.5)) and np.allclose(o[0]["bounds"]["max"], (1.5, 2.5, 3.5))
    # ground: zero-thickness box shifted +0.01 along the normal on BOTH corners (src/SceneIO.h:98-103)
    assert np.allclose(o[1]["bounds"]["min"], (-5, np.float32(-1) + np.float32(0.01), -10))
    assert np.allclose(o[1]["bounds"]["max"], (5, np.float32(-1) + np.float32(0.01), 0))
    assert o[1]["bounds"]["min"][1] == o[1]["bounds"]["max"][1]
    # n = +Z: right = normalize(cross(n,(0,1,0))) = (-1,0,0) -> min.x > max.x
    assert o[2]["bounds"]["min"][0] == 10 and o[2]["bounds"]["max"][0] == -10


def test_parser_stream_semantics():
    text = ("OBJECT SPHERE A 1 2 3 0.5 0 0 0 0 0 1 0.1 0.2 0.3 0.4 0.5 1.5 0.9\n"      # one token short: specular keeps its default
            "\n"
            "OBJECT PLANE G 0 -1 0 0 0 1 0 4 6 0 0.8 0.8 0.8 0 0 1 0 0\n"
            "LIGHT AREA L 0 3 0 0 -1 0 1 1 0.9 5 0.5 16\n"
            "LIGHT WHATEVER P 1 1 1 0 0 0 1 0.8 0.7 8 1.5 0\n")
    sc = S.parse_scene(text)
    assert len(sc.objects) == 2 and len(sc.lights) == 2
    a = sc.objects[0]
    assert a["type"] == S.SPHERE and tuple(a["position"]) == (1, 2, 3) and a["radius"] == np.float32(0.5)
    m = a["material"]
    assert m["type"] == 1 and np.allclose(m["albedo"], (0.1, 0.2, 0.3)) and m["metallic"] == np.float32(0.4)
    assert m["ior"] == np.float32(1.5) and m["transparency"] == np.float32(0.9) and m["specular"] == np.float32(0.5)
    assert m["diffuseStrength"] == 0.0 and m["scatterDistance"] == np.float32(0.1)   # q3 defaults
    assert sc.lights[0]["type"] == S.AREA and sc.lights[1]["type"] == S.POINT
    assert sc.lights[0]["shadowType"] == S.SHADOW_PCF and sc.lights[0]["pcfSamples"] == 4   # q5 defaults
    assert sc.lights[0]["samples"] == 16
    # round trip through the (lossy) save format
    again = S.parse_scene(S.save_scene(sc))
    assert again.objects.tobytes() == sc.objects.tobytes() and again.lights.tobytes() == sc.lights.tobytes()


@pytest.mark.skipif(not os.path.isdir(REF_SCENES), reason="reference tree not present")
def test_default_scene_matches_reference_file():
    sc = S.load_scene(os.path.join(REF_SCENES, "default.scene"))
    mine = synthetic.default_scene()
    assert sc.objects.tobytes() == mine.objects.tobytes()
    assert sc.lights.tobytes() == mine.lights.tobytes()
    assert sc.object_names == mine.object_names


@pytest.mark.skipif(not os.path.isdir(REF_SCENES), reason="reference tree not present")
@pytest.mark.parametrize("name,nobj,nlights", [("SIMPLE", 15, 3), ("performance_test", 15, 8), ("default", 4, 3)])
def test_shipped_scenes_decode_like_golden(name, nobj, nlights):
    sc = S.load_scene(os.path.join(REF_SCENES, name + ".scene"))
    assert len(sc.objects) == nobj and len(sc.lights) == nlights
    g = np.load(os.path.join(GOLDEN, "scenes.npz"))
    assert sc.objects.tobytes() == g[name + "_objects"].tobytes()
    assert sc.lights.tobytes() == g[name + "_lights"].tobytes()


def test_performance_scene_short_lines():
    g = np.load(os.path.join(GOLDEN, "scenes.npz"))
    objs = g["performance_test_objects"].view(S.OBJECT_DTYPE)
    glass = objs[7]                       # "Glass1": one token short: the sentry hits end-of-line, specular keeps 0.5
    m = glass["material"]
    assert m["specular"] == np.float32(0.5) and m["ior"] == 0.0 and m["transparency"] == np.float32(0.95)


def test_camera_defaults_and_update_vectors():
    c = S.Camera()
    assert c.front == (0.0, 0.0, -1.0) and c.fov == 45.0
    c.update_vectors()
    assert np.allclose(c.front, (0, 0, -1), atol=1e-6) and np.allclose(c.right, (1, 0, 0), atol=1e-6)
    assert np.allclose(c.up, (0, 1, 0), atol=1e-6)


def test_synthetic_configs_are_deterministic():
    a, b = synthetic.config(3), synthetic.config(3)
    assert a.objects.tobytes() == b.objects.tobytes() and len(a.objects) == 260 and len(a.lights) == 4
    c5 = synthetic.config(5)
    assert len(c5.objects) == 1024 and (c5.width, c5.height) == (7680, 4320)
    assert synthetic.config(4).num_samples == 64
    assert synthetic.SplitMix64(0).next() == 0xE220A8397B1DCDAF           # published splitmix64 vector
    n1 = synthetic.noise_n1(16)
    r = synthetic.SplitMix64(0xB1E0)
    assert [int(v) for v in n1.ravel()[:8]] == [r.next() & 0xFF for _ in range(8)]


def test_skybox_face_convention():
    sky = synthetic.skybox_s1(8).astype(np.float32)
    # +Y face (zenith) is bluer than -Y; sun lives on +Y/+X/+Z side
    assert sky[2, :, :, 2].mean() > sky[3, :, :, 2].mean()
    d = synthetic._face_dirs(4)
    assert (d[0, :, :, 0] > 0).all() and (d[1, :, :, 0] < 0).all() and (d[5, :, :, 2] < 0).all()


# --------------------------------------------------------------------------- lossless v2 records
@pytest.mark.parametrize("cfg", [1, 2, 3, 4, 5])
def test_v2_format_round_trips_every_field(cfg):
    """save_scene_v2 -> parse_scene gives back the records byte for byte (bounds included: they
    are regenerated), for scenes that use the fields the reference's own format drops
    (diffuseStrength in C4, PCSS / pcfSamples / lightSize in C2-C4)."""
    w = synthetic.config(cfg)
    text = S.save_scene_v2(S.Scene(objects=w.objects, lights=w.lights))
    back = S.parse_scene(text)
    assert back.objects.tobytes() == w.objects.tobytes()
    assert back.lights.tobytes() == w.lights.tobytes()
    # the lossy format does lose them (that is what v2 is for)
    if cfg in (2, 3, 4):
        lossy = S.parse_scene(S.save_scene(S.Scene(objects=w.objects, lights=w.lights)))
        assert lossy.lights.tobytes() != w.lights.tobytes()


def test_v2_format_edge_values_and_mixed_files():
    o = S.new_objects(3)
    o["type"] = [S.SPHERE, S.PLANE, 7]                        # 7: a tag the enums do not name
    o["position"][0] = (1.0000001, -3.4028235e38, 1.1754944e-38)      # 1 ulp above 1, -FLT_MAX, FLT_MIN
    o["position"][1] = (1e-45, -0.0, 16777217.0)                       # denormal, negative zero
    o["radius"] = [-2.5, 0.0, 1.0 / 3.0]
    o["material"]["diffuseStrength"] = [0.8, 0.0, 1e-3]
    o["material"]["subsurfaceScatter"] = [0.25, 0.0, 0.0]
    o["material"]["subsurfaceColor"][0] = (0.1, 0.2, 0.3)
    o["material"]["scatterDistance"] = [0.7, 0.1, 0.1]
    o["material"]["type"] = [0, 1, -3]
    S.generate_aabbs(o)
    l = S.new_lights(3)
    l["type"] = [S.POINT, S.AREA, 9]
    l["shadowType"] = [0, 2, 5]
    l["pcfSamples"] = [0, 16, -1]
    l["lightSize"] = [0.0, 2.5, 1.0]
    l["shadowSoftness"] = [3.0, 0.5, 1.0]
    l["angularRadius"] = [0.0, 0.125, 0.0]
    sc = S.Scene(objects=o, lights=l, object_names=["a", "b", "c"], light_names=["x", "y", "z"])
    text = S.save_scene_v2(sc)
    back = S.parse_scene(text)
    assert back.objects.tobytes() == o.tobytes() and back.lights.tobytes() == l.tobytes()
    assert back.object_names == ["a", "b", "c"] and back.light_names == ["x", "y", "z"]
    # v1 and v2 records in one file, in file order per kind
    mixed = "OBJECT SPHERE First 0 0 -5 1 0 0 0 0 0 0 1 1 1 0 0.5 1 0 0.5\n" + text + "LIGHT POINT Last 0 1 0 0 -1 0 1 1 1 2 0.5 4\n"
    m = S.parse_scene(mixed)
    assert len(m.objects) == 4 and len(m.lights) == 4
    assert m.object_names == ["First", "a", "b", "c"] and m.light_names == ["x", "y", "z", "Last"]
    assert m.objects[1:].tobytes() == o.tobytes()
    # a short v2 line keeps the defaults of the missing tail, like a short v1 line
    short = S.parse_scene("LIGHT2 AREA L 0 3 0 0 1 0 1 1 1 5 0.5 4 2.0 2\n")
    assert int(short.lights["shadowType"][0]) == 2 and float(short.lights["shadowSoftness"][0]) == 2.0
    assert int(short.lights["pcfSamples"][0]) == 4 and float(short.lights["lightSize"][0]) == 1.0
    # non-finite values cannot be extracted by an istream: refused at save time
    o["radius"][0] = np.inf
    with pytest.raises(ValueError):
        S.save_scene_v2(S.Scene(objects=o, lights=l))
