"""C-ABI library: loads and exports every symbol include/rt_abi.h, include/rt_post.h, include/rt_frame.h and include/rt_multi.h declare; the ctypes
structures match the header's byte layouts. No compute calls (runs without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from opengl_raytracing_b200 import abi, scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rt_abi.h")
POST_HEADER = os.path.join(ROOT, "include", "rt_post.h")
FRAME_HEADER = os.path.join(ROOT, "include", "rt_frame.h")
MULTI_HEADER = os.path.join(ROOT, "include", "rt_multi.h")


def _declared_functions(header=HEADER):
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_functions_match_binding_list():
    assert _declared_functions() == sorted(abi.EXPORTED_SYMBOLS)
    assert _declared_functions(POST_HEADER) == sorted(abi.POST_EXPORTED_SYMBOLS)
    assert _declared_functions(FRAME_HEADER) == sorted(abi.FRAME_EXPORTED_SYMBOLS)
    assert _declared_functions(MULTI_HEADER) == sorted(abi.MULTI_EXPORTED_SYMBOLS)


def test_library_loads_and_exports_every_symbol():
    from opengl_raytracing_b200 import build
    build.build()
    lib = abi.load_library()
    for name in _declared_functions() + _declared_functions(POST_HEADER) + _declared_functions(FRAME_HEADER) + _declared_functions(MULTI_HEADER):
        assert hasattr(lib, name), name
    assert lib.rt_abi_version() == 3
    assert lib.rt_status_string(5) == b"RT_ERR_UNAVAILABLE"


def test_struct_layouts_against_c_compiler(tmp_path):
    """sizeof/offsetof from gcc on the header == numpy dtypes == ctypes FrameParams."""
    src = tmp_path / "probe.c"
    fields = [("rt_object", f) for f in ("type", "position", "radius", "normal", "size", "material", "bounds")]
    fields += [("rt_material", f) for f in ("albedo", "metallic", "roughness", "diffuseStrength", "ior", "transparency",
                                             "specular", "subsurfaceScatter", "subsurfaceColor", "scatterDistance")]
    fields += [("rt_light", f) for f in ("position", "direction", "color", "intensity", "radius", "samples",
                                          "shadowSoftness", "shadowType", "pcfSamples", "lightSize", "angularRadius")]
    fields += [("rt_frame_params", f) for f in ("row_begin", "camera_pos", "fov_deg", "use_skybox", "noise_scale",
                                                 "frame_count", "num_samples", "max_ray_depth", "flags", "row_band", "row_period", "row_phase")]
    body = "".join(f'printf("{s}.{f} %zu\\n", offsetof({s}, {f}));\n' for s, f in fields)
    body += "".join(f'printf("sizeof.{s} %zu\\n", sizeof({s}));\n' for s in ("rt_object", "rt_material", "rt_light", "rt_aabb", "rt_frame_params"))
    src.write_text(f'#include <stdio.h>\n#include "{HEADER}"\nint main(void){{\n{body}return 0;}}\n')
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c11", str(src), "-o", str(exe)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    assert int(out["sizeof.rt_object"]) == 176 == S.OBJECT_DTYPE.itemsize
    assert int(out["sizeof.rt_material"]) == 80 == S.MATERIAL_DTYPE.itemsize
    assert int(out["sizeof.rt_light"]) == 96 == S.LIGHT_DTYPE.itemsize
    assert int(out["sizeof.rt_aabb"]) == 32 == S.AABB_DTYPE.itemsize
    assert int(out["sizeof.rt_frame_params"]) == C.sizeof(abi.FrameParams)
    # rt_post_params
    psrc = tmp_path / "probe_post.c"
    pf = ("width", "height", "bloom_threshold", "bloom_strength", "blur_passes", "taa_enabled", "taa_blend", "jitter_x", "jitter_y", "flags")
    pbody = "".join(f'printf("{f} %zu\\n", offsetof(rt_post_params, {f}));\n' for f in pf) + 'printf("sizeof %zu\\n", sizeof(rt_post_params));\n'
    psrc.write_text(f'#include <stdio.h>\n#include "{POST_HEADER}"\nint main(void){{\n{pbody}return 0;}}\n')
    pexe = tmp_path / "probe_post"
    subprocess.run(["gcc", "-std=c11", str(psrc), "-o", str(pexe)], check=True)
    pout = dict(l.split() for l in subprocess.run([str(pexe)], capture_output=True, text=True, check=True).stdout.splitlines())
    assert int(pout["sizeof"]) == C.sizeof(abi.PostParams)
    for f in pf:
        assert getattr(abi.PostParams, f).offset == int(pout[f]), f
    for dt, name in ((S.OBJECT_DTYPE, "rt_object"), (S.MATERIAL_DTYPE, "rt_material"), (S.LIGHT_DTYPE, "rt_light")):
        for f in dt.names:
            key = f"{name}.{f}"
            if key in out:
                assert dt.fields[f][1] == int(out[key]), key
    for f in ("row_begin", "camera_pos", "fov_deg", "use_skybox", "noise_scale", "frame_count", "num_samples", "max_ray_depth", "flags"):
        assert getattr(abi.FrameParams, f).offset == int(out[f"rt_frame_params.{f}"]), f
    # the survey's measured std430 offsets
    assert S.OBJECT_DTYPE.fields["material"][1] == 64 and S.OBJECT_DTYPE.fields["bounds"][1] == 144
    assert S.MATERIAL_DTYPE.fields["subsurfaceColor"][1] == 64 and S.LIGHT_DTYPE.fields["shadowSoftness"][1] == 72


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device rt_create must fail loudly (RT_ERR_UNAVAILABLE), never render."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from opengl_raytracing_b200.pipeline import RtContext
    with pytest.raises(abi.RtError) as ei:
        RtContext(0)
    assert ei.value.status == abi.RT_ERR_UNAVAILABLE


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "opengl_raytracing_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, fn), encoding="utf-8", errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), fn
                assert "oracle/" not in text and "oracle_api" not in text and "librt_oracle" not in text, fn


def test_tuning_and_peer_desc_layouts_against_c_compiler(tmp_path):
    src = tmp_path / "probe2.c"
    body = ('printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(rt_tuning), offsetof(rt_tuning, variant_ms), '
            'offsetof(rt_tuning, variant_phase_level), sizeof(rt_peer_reduce_desc), offsetof(rt_peer_reduce_desc, flags), '
            'offsetof(rt_peer_reduce_desc, epoch));')
    src.write_text('#include <stdio.h>\n#include "' + MULTI_HEADER + '"\nint main(void){' + body + 'return 0;}\n')
    exe = tmp_path / "probe2"
    subprocess.run(["gcc", "-std=c11", str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(abi.Tuning), abi.Tuning.variant_ms.offset, abi.Tuning.variant_phase_level.offset,
            C.sizeof(abi.PeerReduceDesc), abi.PeerReduceDesc.flags.offset, abi.PeerReduceDesc.epoch.offset]
    assert got == want, (got, want)
