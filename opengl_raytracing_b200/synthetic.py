"""Seeded synthetic inputs for the BASELINE.json configs (SURVEY.md 8d "Concrete inputs").

Everything is generated from splitmix64 with the stated seeds so that the oracle, the
CUDA path and the bench all see identical bytes on any machine:
  C1  default scene, 1280x720, depth 1                       (CPU-runnable reference case)
  C2  default scene, 1920x1080, depth 3, PCSS area light, skybox S1, noise N1   (bench workload)
  C3  256 spheres + 4 planes, 4 mixed lights, depth 4 (Russian roulette live), 3840x2160
  C4  C3 with diffuse bounces, 64 samples (sample-range sharded)
  C5  1024 spheres, 7680x4320 (row-tile sharded)
The default scene is written out here as decoded by the reference's parser
(/root/reference/res/Scene/default.scene through src/SceneIO.h:145-186; the columns of the
shipped file are drifted relative to the parser) because the reference tree is
not present on the GPU box; tests/test_scene_io.py checks it against the file when it is.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import scene as S

MASK64 = (1 << 64) - 1


class SplitMix64:
    def __init__(self, seed: int):
        self.state = seed & MASK64

    def next(self) -> int:
        self.state = (self.state + 0x9E3779B97F4A7C15) & MASK64
        z = self.state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
        return z ^ (z >> 31)

    def u(self) -> float:
        """Uniform in [0,1) with 24 bits: (x >> 40) * 2^-24."""
        return (self.next() >> 40) * (1.0 / 16777216.0)

    def uniform(self, lo: float, hi: float) -> float:
        return lo + (hi - lo) * self.u()


def splitmix64_bytes(seed: int, n: int) -> np.ndarray:
    """Low 8 bits of the first n splitmix64 outputs (vectorised)."""
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z & np.uint64(0xFF)).astype(np.uint8)


# --------------------------------------------------------------------------- textures
def noise_n1(size: int = 1024, seed: int = 0xB1E0) -> np.ndarray:
    """N1: size x size R8 white noise (the reference's blue_noise.png is absent from the
    snapshot, .MISSING_LARGE_BLOBS; 1/1024 noiseScale suggests 1024^2)."""
    return splitmix64_bytes(seed, size * size).reshape(size, size)


def _face_dirs(size: int) -> np.ndarray:
    """Texel-centre directions of the 6 faces, GL convention (OpenGL 4.3 table 8.19):
    returns (6, size, size, 3) float64, row index = t, column index = s."""
    c = (np.arange(size, dtype=np.float64) + 0.5) / size * 2.0 - 1.0
    sc, tc = np.meshgrid(c, c)          # sc varies along columns, tc along rows
    one = np.ones_like(sc)
    faces = [
        (one, -tc, -sc),     # +X: sc = -rz, tc = -ry
        (-one, -tc, sc),     # -X: sc = +rz, tc = -ry
        (sc, one, tc),       # +Y: sc = +rx, tc = +rz
        (sc, -one, -tc),     # -Y: sc = +rx, tc = -rz
        (sc, -tc, one),      # +Z: sc = +rx, tc = -ry
        (-sc, -tc, -one),    # -Z: sc = -rx, tc = -ry
    ]
    d = np.stack([np.stack(f, axis=-1) for f in faces])
    return d / np.linalg.norm(d, axis=-1, keepdims=True)


def skybox_s1(size: int = 512) -> np.ndarray:
    """S1: (6, size, size, 4) float16. radiance = mix((0.9,0.8,0.7),(0.2,0.4,1.0),
    0.5(d.y+1)) + 40 * max(d . normalize(0.5,1,0.5), 0)^256; alpha 1."""
    d = _face_dirs(size)
    t = 0.5 * (d[..., 1] + 1.0)
    horizon = np.array([0.9, 0.8, 0.7])
    zenith = np.array([0.2, 0.4, 1.0])
    rgb = horizon * (1.0 - t[..., None]) + zenith * t[..., None]
    sun = np.array([0.5, 1.0, 0.5]) / np.linalg.norm([0.5, 1.0, 0.5])
    rgb = rgb + 40.0 * np.maximum(d @ sun, 0.0)[..., None] ** 256
    out = np.ones((6, size, size, 4), dtype=np.float16)
    out[..., :3] = rgb.astype(np.float16)
    return out


# --------------------------------------------------------------------------- scenes
def default_scene() -> S.Scene:
    """res/Scene/default.scene as the reference's parser decodes it."""
    o = S.new_objects(4)
    rows = [  # type, position, radius, normal, size, mtype, albedo, metallic, roughness, ior, transp, specular
        (S.SPHERE, (-2.5, 0.5, -5), 1.0, (0, 0, 0), (0, 0), 0, (0, 0.95, 0.9), 0.924, 0, 0, 0, 0),
        (S.SPHERE, (0, 0.5, -5), 0.8, (0, 0, 0), (0, 0), 0, (1, 0.1, 0.1), 0.1, 1, 0.05, 0, 0.95),
        (S.SPHERE, (2.5, 0.5, -5), 1.0, (0, 0, 0), (0, 0), 2, (0.2, 0.5, 0.8), 0, 0.5, 1, 0, 0.6),
        (S.PLANE, (0, -1, -5), 0.0, (0, 1, 0), (10, 10), 0, (0.8, 0.8, 0.8), 0, 0, 1, 0, 0),
    ]
    for i, (t, p, r, n, sz, mt, alb, met, rough, ior, tr, spec) in enumerate(rows):
        o[i]["type"], o[i]["position"], o[i]["radius"], o[i]["normal"], o[i]["size"] = t, p, r, n, sz
        m = o[i]["material"]
        m["type"], m["albedo"], m["metallic"], m["roughness"] = mt, alb, met, rough
        m["ior"], m["transparency"], m["specular"] = ior, tr, spec
    S.generate_aabbs(o)
    l = S.new_lights(3)
    lrows = [  # type, position, direction, color, intensity, radius, samples
        (S.DIRECTIONAL, (0, 5, 0), (0.5, -1, -0.5), (1, 1, 1), 3, 0, 1),
        (S.AREA, (0, 3.5, 0), (0, -1, 0), (1, 1, 0.9), 5, 0.5, 16),
        (S.POINT, (0, 2.5, -3), (0, 0, 0), (1, 0.8, 0.7), 8, 1.5, 0),
    ]
    for i, (t, p, d, c, inten, rad, smp) in enumerate(lrows):
        l[i]["type"], l[i]["position"], l[i]["direction"], l[i]["color"] = t, p, d, c
        l[i]["intensity"], l[i]["radius"], l[i]["samples"] = inten, rad, smp
    sc = S.Scene(objects=o, lights=l)
    sc.object_names = ["MetallicSphere", "GlassSphere", "PlasticSphere", "Ground"]
    sc.light_names = ["MainLight", "AreaLight", "BloomLight"]
    return sc


def _plane(o, pos, normal, size, albedo=(0.8, 0.8, 0.8)):
    o["type"], o["position"], o["radius"], o["normal"], o["size"] = S.PLANE, pos, 0.0, normal, size
    m = o["material"]
    m["type"], m["albedo"], m["metallic"], m["roughness"], m["ior"] = 0, albedo, 0.0, 0.5, 1.0
    m["transparency"], m["specular"], m["diffuseStrength"] = 0.0, 0.0, 0.0


def sphere_grid_scene(grid: int, x_half: float, z_near: float, z_far: float, r_lo: float, r_hi: float,
                      seed: int, planes: bool, diffuse_strength: float = 0.0) -> np.ndarray:
    """grid x grid jittered spheres resting on y = -1 (+ 4 room planes when ``planes``).
    Material class by index mod 3: 0 mirror (ior U[1.3,2], transparency 0), 1 glass
    (transparency 0.9, ior 1.5), 2 rough dielectric (roughness U[0.2,0.9], ior 1,
    diffuseStrength = ``diffuse_strength``). Draw order per sphere: jitter x, jitter z,
    radius, albedo r g b, metallic, class parameter."""
    rng = SplitMix64(seed)
    n = grid * grid
    o = S.new_objects(n + (4 if planes else 0))
    cell_x = 2.0 * x_half / grid
    cell_z = (z_near - z_far) / grid
    for idx in range(n):
        gx, gz = idx % grid, idx // grid
        x = -x_half + (gx + 0.25 + 0.5 * rng.u()) * cell_x
        z = z_near - (gz + 0.25 + 0.5 * rng.u()) * cell_z
        r = rng.uniform(r_lo, r_hi)
        ob = o[idx]
        ob["type"], ob["radius"], ob["normal"], ob["size"] = S.SPHERE, r, (0, 0, 0), (0, 0)
        ob["position"] = (x, np.float32(r) - np.float32(1.0), z)
        m = ob["material"]
        m["albedo"] = (rng.uniform(0.1, 1.0), rng.uniform(0.1, 1.0), rng.uniform(0.1, 1.0))
        m["metallic"] = rng.u()
        cls = idx % 3
        m["specular"], m["diffuseStrength"], m["transparency"] = 0.5, 0.0, 0.0
        if cls == 0:
            m["type"], m["ior"], m["roughness"] = 0, rng.uniform(1.3, 2.0), 0.1
        elif cls == 1:
            rng.u()
            m["type"], m["ior"], m["roughness"], m["transparency"] = 1, 1.5, 0.05, 0.9
        else:
            m["type"], m["ior"], m["roughness"] = 2, 1.0, rng.uniform(0.2, 0.9)
            m["diffuseStrength"] = diffuse_strength
    if planes:
        _plane(o[n + 0], (0, -1, -20), (0, 1, 0), (80, 80))
        _plane(o[n + 1], (0, 9, -40), (0, 0, 1), (80, 40), (0.7, 0.7, 0.75))
        _plane(o[n + 2], (-20, 9, -20), (1, 0, 0), (80, 40), (0.75, 0.7, 0.7))
        _plane(o[n + 3], (20, 9, -20), (-1, 0, 0), (80, 40), (0.7, 0.75, 0.7))
    return S.generate_aabbs(o)


def _light(l, typ, pos, direction, color, intensity, shadow_type, pcf, light_size=1.0):
    l["type"], l["position"], l["direction"], l["color"], l["intensity"] = typ, pos, direction, color, intensity
    l["shadowType"], l["pcfSamples"], l["lightSize"], l["radius"], l["samples"] = shadow_type, pcf, light_size, 0.5, 4


# --------------------------------------------------------------------------- configs
@dataclass
class Workload:
    """One BASELINE.json config: everything rt_upload_scene / rt_render need."""
    name: str
    objects: np.ndarray
    lights: np.ndarray
    width: int
    height: int
    max_ray_depth: int
    frame_count: int = 0
    num_samples: int = 1
    use_skybox: bool = False
    skybox: np.ndarray | None = None      # (6,S,S,4) float16
    noise: np.ndarray | None = None       # (h,w) uint8
    camera: S.Camera = field(default_factory=S.Camera)
    description: str = ""


def config(n: int, width: int | None = None, height: int | None = None) -> Workload:
    """BASELINE.json configs[n-1]; width/height override the resolution (parity tests run
    the same scene at sizes the CPU oracle finishes in seconds)."""
    if n == 1:
        sc = default_scene()
        w = Workload("C1", sc.objects, sc.lights, 1280, 720, 1,
                     description="default.scene, point/dir/area lights PCFx4, depth 1, 1 spp, no skybox, noise unbound")
    elif n == 2:
        sc = default_scene()
        # Documented override: the area light becomes PCSS and faces the scene
        # (direction (0,1,0): the shader's cosine uses the P->light vector, q12).
        a = sc.lights[1]
        a["shadowType"], a["lightSize"], a["direction"] = S.SHADOW_PCSS, 1.0, (0.0, 1.0, 0.0)
        w = Workload("C2", sc.objects, sc.lights, 1920, 1080, 3, use_skybox=True,
                     skybox=skybox_s1(), noise=noise_n1(),
                     description="default.scene, PBR + PCSS area light + PCF, skybox S1, noise N1, depth 3, 1 spp")
    elif n in (3, 4):
        o = sphere_grid_scene(16, 16.0, -4.0, -36.0, 0.3, 0.9, 0xC3, planes=True,
                              diffuse_strength=0.8 if n == 4 else 0.0)
        l = S.new_lights(4)
        _light(l[0], S.DIRECTIONAL, (0, 10, 0), (0.5, -1, -0.5), (1, 1, 1), 3, S.SHADOW_PCF, 4)
        _light(l[1], S.POINT, (0, 6, -10), (0, -1, 0), (1, 0.9, 0.8), 8, S.SHADOW_PCF, 4)
        _light(l[2], S.POINT, (-8, 4, -20), (0, -1, 0), (0.8, 0.9, 1), 6, S.SHADOW_PCF, 8)
        _light(l[3], S.AREA, (6, 8, -16), (0, 1, 0), (1, 1, 0.9), 40, S.SHADOW_PCSS, 4)
        if n == 3:
            w = Workload("C3", o, l, 3840, 2160, 4, use_skybox=True, skybox=skybox_s1(), noise=noise_n1(),
                         description="256 spheres + 4 planes, dir/point/point/area(PCSS) lights, refraction, depth 4 + RR")
        else:
            w = Workload("C4", o, l, 3840, 2160, 3, num_samples=64, use_skybox=True, skybox=skybox_s1(),
                         noise=noise_n1(),
                         description="C3 scene with diffuseStrength 0.8 on rough dielectrics, depth 3, 64 spp Hammersley")
    elif n == 5:
        o = sphere_grid_scene(32, 32.0, -4.0, -68.0, 0.2, 0.8, 0xC5, planes=False)
        l = S.new_lights(2)
        _light(l[0], S.DIRECTIONAL, (0, 10, 0), (0.5, -1, -0.5), (1, 1, 1), 3, S.SHADOW_PCF, 4)
        _light(l[1], S.POINT, (0, 10, -30), (0, -1, 0), (1, 0.95, 0.9), 60, S.SHADOW_PCF, 4)
        w = Workload("C5", o, l, 7680, 4320, 3, use_skybox=True, skybox=skybox_s1(), noise=noise_n1(),
                     description="1024 spheres (AABB-culled linear loop), dir + point lights PCFx4, depth 3")
    else:
        raise ValueError("config must be 1..5")
    if width is not None:
        w.width = width
    if height is not None:
        w.height = height
    return w
