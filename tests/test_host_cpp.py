"""The C++ host layer (host/rt_host.hpp + host/rt_headless): SceneIO / GenerateAABBForObject /
Camera against the Python mirror and the golden decodes of the shipped scenes (CPU), and one
frame through ForwardShadingPipeline::Render against the Python path (GPU, bit-identical)."""
import json
import os
import subprocess

import numpy as np
import pytest

from opengl_raytracing_b200 import build as B, scene as S, synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_SCENES = "/root/reference/res/Scene"


@pytest.fixture(scope="module")
def headless():
    if not os.path.isfile(B.LIB):
        pytest.skip("librt_b200.so not built")
    return B.build_host()


def _dump(headless, scene_path, tmp_path):
    out = str(tmp_path / "scene.bin")
    r = subprocess.run([headless, "--dump-scene", scene_path, out], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    blob = open(out, "rb").read()
    n, m = np.frombuffer(blob[:8], dtype="<i4")
    objs = np.frombuffer(blob[8:8 + 176 * n], dtype=np.uint8)
    lights = np.frombuffer(blob[8 + 176 * n:], dtype=np.uint8)
    assert len(lights) == 96 * m
    names = [ln.split(" ", 1) for ln in r.stdout.splitlines()]
    return objs, lights, names


TRICKY = ("OBJECT SPHERE A 1 2 3 0.5 0 0 0 0 0 1 0.1 0.2 0.3 0.4 0.5 1.5 0.9\n"          # one token short
          "\n"
          "garbage line that is neither\n"
          "OBJECT PLANE G 0 -1 0 0 0 1 0 4 6 0 0.8 0.8 0.8 0 0 1 0 0\n"
          "OBJECT PLANE Tilted 1 2 3 0 0.6 0.3 0.74 7 5 2 0.5 0.25 0.125 1 0.05 1.33 0.5 0.5\n"
          "OBJECT PLANE Wall 0 0 -10 oops 0 0 1 20 20 0 1 1 1 0 0 1 0 0\n"             # failed extraction mid-line
          "OBJECT WHATEVER Odd -1e1 2.5e-1 .5 2\n"                                       # unknown type -> sphere, short
          "LIGHT AREA L 0 3 0 0 -1 0 1 1 0.9 5 0.5 16\n"
          "LIGHT WHATEVER P 1 1 1 0 0 0 1 0.8 0.7 8 1.5 0\n"
          "LIGHT DIRECTIONAL Sun 0 0 0 -0.3 -1 -0.2 1 0.95 0.9 2.5\n")


def test_cpp_sceneio_matches_python_mirror(headless, tmp_path):
    p = tmp_path / "tricky.scene"
    p.write_text(TRICKY)
    objs, lights, names = _dump(headless, str(p), tmp_path)
    sc = S.parse_scene(TRICKY)
    assert objs.tobytes() == sc.objects.tobytes()
    assert lights.tobytes() == sc.lights.tobytes()
    assert [n[1] for n in names if n[0] == "OBJECT"] == sc.object_names
    assert [n[1] for n in names if n[0] == "LIGHT"] == sc.light_names


@pytest.mark.parametrize("cfg", [1, 3])
def test_cpp_sceneio_round_trips_synthetic_scenes(headless, tmp_path, cfg):
    """Python Save -> C++ Load -> C++ Save -> Python Load: the same records at every step
    (the format is lossy only in fields the generators leave at their defaults... except the
    shadow fields, which are compared after the first load)."""
    w = synthetic.config(cfg)
    sc = S.Scene(objects=w.objects, lights=w.lights)
    text = S.save_scene(sc)
    p = tmp_path / "a.scene"
    p.write_text(text)
    objs, lights, _ = _dump(headless, str(p), tmp_path)
    again = S.parse_scene(text)
    assert objs.tobytes() == again.objects.tobytes() and lights.tobytes() == again.lights.tobytes()
    q = tmp_path / "b.scene"
    r = subprocess.run([headless, "--resave", str(p), str(q)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    third = S.parse_scene(q.read_text())
    assert third.objects.tobytes() == again.objects.tobytes() and third.lights.tobytes() == again.lights.tobytes()


@pytest.mark.skipif(not os.path.isdir(REF_SCENES), reason="reference tree not present")
@pytest.mark.parametrize("name", ["default", "SIMPLE", "performance_test"])
def test_cpp_sceneio_decodes_shipped_scenes_like_golden(headless, tmp_path, name):
    objs, lights, _ = _dump(headless, os.path.join(REF_SCENES, name + ".scene"), tmp_path)
    g = np.load(os.path.join(GOLDEN, "scenes.npz"))
    assert objs.tobytes() == g[name + "_objects"].tobytes()
    assert lights.tobytes() == g[name + "_lights"].tobytes()


def test_headless_fails_loudly_without_a_gpu(headless, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = tmp_path / "d.scene"
    p.write_text(S.save_scene(synthetic.default_scene()))
    r = subprocess.run([headless, "--scene", str(p), "--size", "64", "32", "--out", str(tmp_path / "f")],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "rt_create" in r.stderr


@pytest.mark.gpu
def test_headless_frame_is_bit_identical_to_python_path(headless, tmp_path):
    from opengl_raytracing_b200.pipeline import RtContext, render_workload, workload_params
    w = synthetic.config(2, 320, 200)
    # the text format drops the shadow fields: render what the file holds, on both sides
    text = S.save_scene(S.Scene(objects=w.objects, lights=w.lights, object_names=[], light_names=[]))
    loaded = S.parse_scene(text)
    p = tmp_path / "c2.scene"
    p.write_text(text)
    noise = tmp_path / "noise.r8"
    w.noise.astype(np.uint8).tofile(noise)
    sky = tmp_path / "sky.f16"
    skyf = np.ascontiguousarray(w.skybox.astype(np.float16))
    skyf.tofile(sky)
    cam = S.Camera(position=(0.5, 1.0, 2.0), yaw=-100.0, pitch=-8.0, fov=50.0)
    cam.update_vectors()
    out = str(tmp_path / "f")
    r = subprocess.run([headless, "--scene", str(p), "--size", "320", "200", "--frames", "3", "--raw", "--out", out,
                        "--noise", str(noise), str(w.noise.shape[1]), str(w.noise.shape[0]),
                        "--skybox", str(sky), str(skyf.shape[1]), str(skyf.shape[3]), "f16",
                        "--camera", "0.5", "1.0", "2.0", "-100", "-8", "50"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["launches"] >= 3 and info["objects"] == len(w.objects)
    color = np.fromfile(out + "_color.f32", dtype=np.float32).reshape(200, 320, 4)
    position = np.fromfile(out + "_position.f32", dtype=np.float32).reshape(200, 320, 4)
    normal = np.fromfile(out + "_normal.f16", dtype=np.float16).reshape(200, 320, 4)
    # Python path: same scene bytes, same camera, frameCount = 2 (the third frame)
    ctx = RtContext(0)
    w.objects, w.lights, w.camera, w.frame_count = loaded.objects, loaded.lights, cam, 2
    k = render_workload(ctx, w, params=workload_params(w))
    ctx.close()
    assert color.tobytes() == k["color"].tobytes()
    assert position.tobytes() == k["position"].tobytes()
    assert normal.astype(np.float32).tobytes() == k["normal"].tobytes()
    pfm = open(out + "_color.pfm", "rb").read()
    assert pfm.startswith(b"PF\n320 200\n-1.0\n") and len(pfm) == 16 + 320 * 200 * 12


@pytest.mark.parametrize("cfg", [2, 4])
def test_cpp_v2_format_is_lossless_and_matches_python(headless, tmp_path, cfg):
    """Python SaveV2 -> C++ Load == the original records (shadow fields, diffuseStrength and all);
    C++ SaveV2 of that == the Python text, byte for byte; and C++ Load of a v1 file + SaveV2 ->
    Python Load keeps what the v1 file held."""
    w = synthetic.config(cfg)
    text = S.save_scene_v2(S.Scene(objects=w.objects, lights=w.lights))
    p = tmp_path / "a2.scene"
    p.write_text(text)
    objs, lights, _ = _dump(headless, str(p), tmp_path)
    assert objs.tobytes() == w.objects.tobytes() and lights.tobytes() == w.lights.tobytes()
    q = tmp_path / "b2.scene"
    r = subprocess.run([headless, "--resave-v2", str(p), str(q)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    assert q.read_text() == text
    v1 = tmp_path / "v1.scene"
    v1.write_text(S.save_scene(S.Scene(objects=w.objects, lights=w.lights)))
    r = subprocess.run([headless, "--resave-v2", str(v1), str(q)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    a, b = S.parse_scene(v1.read_text()), S.parse_scene(q.read_text())
    assert a.objects.tobytes() == b.objects.tobytes() and a.lights.tobytes() == b.lights.tobytes()


@pytest.mark.gpu
def test_headless_whole_frames_with_post_chain_match_python_frame(headless, tmp_path):
    """rt_headless --post (ForwardShadingPipeline::RenderFrame: ray tracing + bloom + TAA on
    device-resident images, lossless v2 scene file) against the Python Frame driven the same way:
    the presented 8-bit image and the float display image of the third frame, bit for bit."""
    from opengl_raytracing_b200 import abi, post as P
    from opengl_raytracing_b200.frame import Frame
    from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params
    W, H = 256, 144
    w = synthetic.config(2, W, H)
    p = tmp_path / "c2_v2.scene"
    p.write_text(S.save_scene_v2(S.Scene(objects=w.objects, lights=w.lights)))      # keeps the PCSS light
    noise = tmp_path / "noise.r8"
    w.noise.astype(np.uint8).tofile(noise)
    sky = tmp_path / "sky.f16"
    skyf = np.ascontiguousarray(w.skybox.astype(np.float16))
    skyf.tofile(sky)
    out = str(tmp_path / "g")
    r = subprocess.run([headless, "--scene", str(p), "--size", str(W), str(H), "--frames", "3", "--raw", "--post", "--out", out,
                        "--noise", str(noise), str(w.noise.shape[1]), str(w.noise.shape[0]),
                        "--skybox", str(sky), str(skyf.shape[1]), str(skyf.shape[3]), "f16"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    d8 = np.fromfile(out + "_display.rgba8", dtype=np.uint8).reshape(H, W, 4)
    d32 = np.fromfile(out + "_display.f32", dtype=np.float32).reshape(H, W, 4)
    ctx = RtContext(0)
    bind_workload(ctx, w)
    fr = Frame(ctx, W, H)
    for fc in range(3):
        fr.render(workload_params(w, frame_count=fc), P.post_params(W, H, blend=0.1, jitter=P.taa_jitter(fc, W, H)))
    want32 = fr.read_display(fmt=abi.DISPLAY_RGBA32F)
    want8 = fr.read_display(fmt=abi.DISPLAY_RGBA8)
    fr.close()
    ctx.close()
    assert d32.tobytes() == want32.tobytes()
    assert d8.tobytes() == want8.tobytes()
    ppm = open(out + "_display.ppm", "rb").read()
    head = b"P6\n%d %d\n255\n" % (W, H)
    assert ppm.startswith(head) and len(ppm) == len(head) + W * H * 3
    assert ppm[len(head):len(head) + 3] == d8[H - 1, 0, :3].tobytes()            # top row first


@pytest.mark.gpu
def test_headless_kernel_pin_and_tune_give_the_same_image(headless, tmp_path):
    """The C++ host reaches the variant API (rt_set_tuning / rt_tune through ForwardShadingPipeline::PinKernel / Tune):
    every pinned variant and the tuned choice write the same bytes, and the JSON line names the kernel that ran."""
    w = synthetic.config(3, 256, 144)
    p = tmp_path / "c3.scene"
    p.write_text(S.save_scene(S.Scene(objects=w.objects, lights=w.lights, object_names=[], light_names=[])))
    images = {}
    for mode in ("megakernel", "wavefront", "regroup", "tune"):
        out = str(tmp_path / mode)
        args = ["--tune"] if mode == "tune" else ["--kernel", mode]
        r = subprocess.run([headless, "--scene", str(p), "--size", "256", "144", "--frames", "2", "--depth", "4", "--raw", "--out", out] + args,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        info = json.loads(r.stdout.strip().splitlines()[-1])
        if mode == "tune":
            assert "rt_tune:" in r.stderr and info["kernel"] in ("megakernel", "wavefront", "regroup")
        else:
            assert info["kernel"] == mode
        images[mode] = open(out + "_color.f32", "rb").read() + open(out + "_position.f32", "rb").read()
    assert len(set(images.values())) == 1
