#!/usr/bin/env python3
"""Regenerates tests/golden/*.npz. Run in the build container (needs /root/reference):

  scenes.npz  the three shipped scene files decoded by this repo's restatement of SceneIO
              (the reference tree is absent on the GPU box)
  post.npz    bloom / TAA / equirect->cube outputs of oracle/_ref/libref_post.so (the reference's
              fragment shaders compiled as C++) on seeded inputs
  tiles.npz   colour / position / normal of small renders produced by oracle/_ref — the
              reference's OWN shader source compiled as C++ — for the five BASELINE configs,
              the shipped scenes and the edge cases. These pin the port (tests/test_oracle.py).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import oracle  # noqa: E402
from opengl_raytracing_b200 import scene as S, synthetic  # noqa: E402

REF_SCENES = "/root/reference/res/Scene"


def golden_workloads():
    from edge_cases import edge_workloads, shipped_scene
    w = {}
    for n, (W, H) in [(1, (64, 36)), (2, (64, 36)), (3, (48, 27)), (5, (48, 27))]:
        w[f"C{n}"] = synthetic.config(n, W, H)
    c4 = synthetic.config(4, 32, 18)
    c4.num_samples = 4
    w["C4"] = c4
    for name in ("SIMPLE", "performance_test", "default"):
        w["scene_" + name] = shipped_scene(name, (40, 40), 3)
    for k, v in edge_workloads().items():
        w["edge_" + k] = v
    return w


def main():
    scenes = {}
    for name in ("default", "SIMPLE", "performance_test"):
        sc = S.load_scene(os.path.join(REF_SCENES, name + ".scene"))
        scenes[name + "_objects"] = np.frombuffer(sc.objects.tobytes(), np.uint8)
        scenes[name + "_lights"] = np.frombuffer(sc.lights.tobytes(), np.uint8)
    np.savez_compressed(os.path.join(HERE, "scenes.npz"), **scenes)
    oracle.build_ref(force=True)
    tiles = {}
    for name, w in golden_workloads().items():
        r = oracle.render(w, "reference")
        for k in ("color", "position", "normal"):
            tiles[f"{name}.{k}"] = r[k]
    np.savez_compressed(os.path.join(HERE, "tiles.npz"), **tiles)
    print("wrote", len(scenes), "scene arrays and", len(tiles), "tile arrays")
    post_golden()


def post_golden():
    """post.npz: outputs of oracle/_ref/libref_post.so (the reference's fragment shaders compiled
    as C++) on the seeded inputs of tests/test_post_oracle.py::post_inputs."""
    from test_post_oracle import post_inputs
    from opengl_raytracing_b200 import post as P
    oracle.build_post_ref(force=True)
    R = oracle.Post("reference")
    color, normal, history = post_inputs()
    h, w = color.shape[:2]
    jx, jy = P.taa_jitter(3, w, h)
    g = {"jitter": np.array([jx, jy], np.float64)}
    g["extract"] = R.bloom_extract(color, 1.0)
    g["blur_h"] = R.gaussian_blur(g["extract"], True)
    g["blur_v"] = R.gaussian_blur(g["extract"], False)
    g["display"], g["history"] = R.post_chain(color, normal, history, blend=0.1, jx=jx, jy=jy)
    rng = np.random.default_rng(21)
    g["equirect"] = (rng.random((32, 64, 3), dtype=np.float32) ** 4 * 60).astype(np.float32)
    g["cube"] = R.equirect_to_cubemap(g["equirect"], 16)
    np.savez_compressed(os.path.join(HERE, "post.npz"), **g)
    print("wrote", len(g), "post arrays")


if __name__ == "__main__":
    main()
