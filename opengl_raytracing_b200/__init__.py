"""opengl_raytracing_b200 — the sm_100a replacement for the reference's per-pixel
ray-tracing compute pass (shader/raytracingCs.glsl) behind the C ABI in include/rt_abi.h.

Layout (only what the hot path needs):
  csrc/         CUDA kernels + the C-ABI shared library (librt_b200.so)
  abi.py        ctypes binding of the C ABI
  pipeline.py   host-side mirror of the reference's call site (SSBO / LightSSBO /
                RaytracingShader / ForwardShadingPipeline)
  scene.py      data contracts, SceneIO parser, GenerateAABBForObject, Camera
  synthetic.py  seeded BASELINE.json workloads (C1..C5), skybox S1, noise N1
  sharding.py   row-tile / sample-range sharding across GPUs (torch.distributed)
  build.py      nvcc build of csrc/ for sm_100a
"""
from . import abi, scene, synthetic  # noqa: F401

__all__ = ["abi", "scene", "synthetic"]
__version__ = "0.1.0"
