"""Edge-case workloads (SURVEY.md 8c KAT list): tiny scenes that exercise the quirks the
oracle pins. Used by the CPU tests (port vs _ref / golden) and the GPU tests (kernel vs port)."""
import numpy as np

from opengl_raytracing_b200 import scene as S, synthetic
from opengl_raytracing_b200.synthetic import Workload


def _sphere(o, pos, r, albedo=(0.8, 0.3, 0.2), **mat):
    o["type"], o["position"], o["radius"], o["normal"], o["size"] = S.SPHERE, pos, r, (0, 0, 0), (0, 0)
    o["material"]["albedo"] = albedo
    for k, v in mat.items():
        o["material"][k] = v


def _plane(o, pos, n, size, albedo=(0.8, 0.8, 0.8), **mat):
    o["type"], o["position"], o["radius"], o["normal"], o["size"] = S.PLANE, pos, 0.0, n, size
    o["material"]["albedo"] = albedo
    for k, v in mat.items():
        o["material"][k] = v


def _light(l, typ, pos, d, shadow=S.SHADOW_PCF, pcf=4, intensity=5.0, **kw):
    l["type"], l["position"], l["direction"], l["intensity"], l["shadowType"], l["pcfSamples"] = typ, pos, d, intensity, shadow, pcf
    for k, v in kw.items():
        l[k] = v


def edge_workloads(size=(48, 32)):
    W, H = size
    out = {}

    # 1. straight-down directional light: cross(lightDir,(0,1,0)) = 0 -> NaN tangent -> NaN shadow rays
    o = S.new_objects(2); l = S.new_lights(1)
    _sphere(o[0], (0, 0, -4), 1.0); _plane(o[1], (0, -1, -4), (0, 1, 0), (12, 12))
    _light(l[0], S.DIRECTIONAL, (0, 5, 0), (0, -1, 0))
    out["dir_light_parallel_y"] = Workload("dir_light_parallel_y", S.generate_aabbs(o), l, W, H, 2)

    # 2. pcfSamples = 0 -> 0/0 = NaN shadow factor -> NaN colour on every lit pixel
    o = S.new_objects(1); l = S.new_lights(1)
    _sphere(o[0], (0, 0, -4), 1.0)
    _light(l[0], S.POINT, (2, 3, -1), (0, -1, 0), pcf=0)
    out["pcf_zero_samples"] = Workload("pcf_zero_samples", S.generate_aabbs(o), l, W, H, 1)

    # 3. ior = 0 glass (eta = inf -> TIR branch, F = 1) and ior = -1 (division by zero in r0)
    o = S.new_objects(3); l = S.new_lights(1)
    _sphere(o[0], (-1.2, 0, -4), 1.0, ior=0.0, transparency=0.95)
    _sphere(o[1], (1.2, 0, -4), 1.0, ior=-1.0, transparency=0.5)
    _plane(o[2], (0, -1, -4), (0, 1, 0), (12, 12))
    _light(l[0], S.POINT, (0, 4, -2), (0, -1, 0))
    out["ior_zero_and_minus_one"] = Workload("ior_zero_and_minus_one", S.generate_aabbs(o), l, W, H, 3,
                                             use_skybox=True, skybox=synthetic.skybox_s1(32))

    # 4. camera inside a sphere: near root only -> the enclosing sphere is never hit
    o = S.new_objects(2); l = S.new_lights(1)
    _sphere(o[0], (0, 0, 0), 3.0); _sphere(o[1], (0, 0, -6), 1.0)
    _light(l[0], S.POINT, (0, 5, -6), (0, -1, 0), shadow=S.SHADOW_NONE)
    out["camera_inside_sphere"] = Workload("camera_inside_sphere", S.generate_aabbs(o), l, W, H, 1)

    # 5. walls with min > max AABBs on some axes, all three plane orientations, roughness 0
    o = S.new_objects(4); l = S.new_lights(2)
    _plane(o[0], (0, -2, -6), (0, 1, 0), (12, 12), roughness=0.0)
    _plane(o[1], (0, 0, -10), (0, 0, 1), (12, 8), roughness=0.0)
    _plane(o[2], (-5, 0, -6), (1, 0, 0), (12, 8))
    _plane(o[3], (5, 0, -6), (-1, 0, 0), (12, 8))
    _light(l[0], S.POINT, (0, 1, -5), (0, -1, 0))
    _light(l[1], S.AREA, (0, 3, -6), (0, 1, 0), shadow=S.SHADOW_PCSS, intensity=30.0)
    out["room_planes_minmax_swapped"] = Workload("room_planes_minmax_swapped", S.generate_aabbs(o), l, W, H, 3)

    # 6. empty scene with skybox, and empty scene without (all-miss)
    out["empty_sky"] = Workload("empty_sky", S.new_objects(0), S.new_lights(0), W, H, 3, use_skybox=True,
                                skybox=synthetic.skybox_s1(32), noise=synthetic.noise_n1(64, 7))
    out["empty_black"] = Workload("empty_black", S.new_objects(0), S.new_lights(0), W, H, 3)

    # 7. no lights; unknown object / light / shadow types
    o = S.new_objects(2); l = S.new_lights(2)
    _sphere(o[0], (0, 0, -4), 1.0); _sphere(o[1], (1.5, 0, -4), 0.7); o[1]["type"] = 7
    S.generate_aabbs(o); o[1]["bounds"] = o[0]["bounds"]
    _light(l[0], 9, (0, 3, -2), (0, -1, 0)); _light(l[1], S.POINT, (0, 3, -2), (0, -1, 0), shadow=5)
    out["unknown_types"] = Workload("unknown_types", o, l, W, H, 2)

    # 8. subsurface scattering (unreachable from the UI; a18) + diffuse bounce with frame_count > 0
    o = S.new_objects(3); l = S.new_lights(1)
    _sphere(o[0], (-1.1, 0, -4), 1.0, subsurfaceScatter=0.7, subsurfaceColor=(1.0, 0.4, 0.3), scatterDistance=0.6)
    _sphere(o[1], (1.1, 0, -4), 1.0, diffuseStrength=0.8, roughness=0.6)
    _plane(o[2], (0, -1, -4), (0, 1, 0), (12, 12), diffuseStrength=0.5, roughness=1.0)
    _light(l[0], S.POINT, (0, 4, -2), (0, -1, 0), pcf=3)
    out["sss_and_diffuse"] = Workload("sss_and_diffuse", S.generate_aabbs(o), l, W, H, 3, frame_count=5, use_skybox=True,
                                      skybox=synthetic.skybox_s1(32), noise=synthetic.noise_n1(64, 9))

    # 9. ragged: 1-pixel-wide and odd sizes are covered by the size parameter of the tests
    return out


def shipped_scene(name, size=(48, 48), depth=3):
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scenes.npz"))
    objs = g[name + "_objects"].view(S.OBJECT_DTYPE).copy()
    lights = g[name + "_lights"].view(S.LIGHT_DTYPE).copy()
    return Workload(name, objs, lights, size[0], size[1], depth)
