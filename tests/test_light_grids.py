"""Host logic of the light grids (rt_abi.cu build_light_grids), without a GPU:
the grids a large scene gets are read back through rt_debug_light_grids and checked for soundness —
every sphere that a PCF shadow ray of a job really hits (double-precision ray/sphere test over
sampled jobs and sampled rays of the shader's own construction, glsl:342-397) must be among the
entries the job would read (its cell + the shared list). The GPU tests check the other half: the
image with the grids equals the image of the cluster scan bit for bit."""
import ctypes as C

import numpy as np
import pytest

from opengl_raytracing_b200 import abi, scene as S, synthetic


def _grids(w):
    lib = abi.load_library()
    objs = np.ascontiguousarray(w.objects)
    lights = np.ascontiguousarray(w.lights)
    off, ent = C.c_int(0), C.c_int(0)
    slots = np.full(len(objs), -1, dtype=np.int32)
    n = lib.rt_debug_light_grids(objs.ctypes.data, len(objs), lights.ctypes.data, len(lights), None, 0, C.byref(off), C.byref(ent), None)
    if n == 0:
        return None
    blob = np.zeros(n * 4, dtype=np.float32)
    assert lib.rt_debug_light_grids(objs.ctypes.data, len(objs), lights.ctypes.data, len(lights), blob.ctypes.data, n,
                                    C.byref(off), C.byref(ent), slots.ctypes.data) == n
    hdr = blob[:16 * len(lights)].reshape(len(lights), 4, 4)
    offsets = blob[4 * off.value:4 * ent.value].view(np.uint32)
    entries = blob[4 * ent.value:].view(np.uint16)
    return hdr, offsets, entries, slots


def _cell(h, L, o, dist):
    """numpy restatement of rt_packet.cuh grid_cell -> (cell, shared_cell) or None."""
    mode = int(h[0, 0].view(np.int32))
    if mode == 1:
        if not dist <= h[0, 2]:
            return None
        G = int(h[1, 3])
        x, y, z = (-L).astype(np.float32)
        ax, ay, az = abs(x), abs(y), abs(z)
        if ax >= ay and ax >= az:
            face, a, b, m = (0 if x >= 0 else 1), y, z, ax
        elif ay >= az:
            face, a, b, m = (2 if y >= 0 else 3), x, z, ay
        else:
            face, a, b, m = (4 if z >= 0 else 5), x, y, az
        ix = int(np.clip(np.floor(a / m * 0.5 * G + 0.5 * G), 0, G - 1))
        iy = int(np.clip(np.floor(b / m * 0.5 * G + 0.5 * G), 0, G - 1))
        return (face * G + iy) * G + ix, 6 * G * G
    v = o - h[3, :3]
    if not float(v @ v) <= h[0, 2]:
        return None
    G = 32
    ix = int(np.clip(np.floor((float(o @ h[1, :3]) - h[1, 3]) * h[3, 3]), 0, G - 1))
    iy = int(np.clip(np.floor((float(o @ h[2, :3]) - h[2, 3]) * h[3, 3]), 0, G - 1))
    return iy * G + ix, G * G


@pytest.mark.parametrize("cfg", [3, 5])
def test_light_grids_hold_every_sphere_a_shadow_ray_hits(cfg):
    w = synthetic.config(cfg)
    got = _grids(w)
    assert got is not None, "a large scene with PCF lights gets grids"
    hdr, offsets, entries, slots = got
    objs, lights = w.objects, w.lights
    sph = np.flatnonzero(objs["type"] == S.SPHERE)
    slot_of = {int(slots[k]): k for k in range(len(sph))}            # object index -> sphere slot
    centres = objs["position"][sph].astype(np.float64)
    radii = np.abs(objs["radius"][sph].astype(np.float64))
    slot_ids = np.array([slot_of[int(i)] for i in sph])
    rng = np.random.default_rng(1234 + cfg)
    gridded = 0
    for j in range(len(lights)):
        h = hdr[j]
        mode = int(h[0, 0].view(np.int32))
        lt = lights[j]
        if int(lt["shadowType"]) != S.SHADOW_PCF:
            assert mode == 0                                       # PCSS / unshadowed lights: no grid
            continue
        assert mode in (1, 2)
        gridded += 1
        first = int(h[0, 3].view(np.int32))
        fs = float(lt["shadowSoftness"]) * 0.005
        checked = 0
        for _ in range(400):
            # a shading point: on a random sphere (outward normal) or, for scenes with planes, on the ground
            k = int(rng.integers(len(sph)))
            n = rng.standard_normal(3); n /= np.linalg.norm(n)
            o = centres[k] + (radii[k] + 0.001) * n
            if cfg == 3 and rng.random() < 0.3:
                o = np.array([rng.uniform(-20, 20), -1.0 + 0.001, rng.uniform(-40, 0)])
            if int(lt["type"]) == S.DIRECTIONAL:
                d = -lt["direction"].astype(np.float64); L = d / np.linalg.norm(d); dist, lim = 1e6, 114514.0
            else:
                d = lt["position"].astype(np.float64) - o; dist = float(np.linalg.norm(d)); L = d / dist; lim = dist
            cell = _cell(h, L.astype(np.float32), o.astype(np.float32), np.float32(dist))
            if cell is None:
                continue                                             # the device takes the cluster scan for this job
            c, shared = cell
            cand = set(entries[offsets[first + c]:offsets[first + c + 1]].tolist()) | \
                set(entries[offsets[first + shared]:offsets[first + shared + 1]].tolist())
            # the shader's PCF rays (glsl:349-378): tangent frame from cross(L, Y), offsets in [0,1)^2 * filterSize
            T = np.cross(L, [0.0, 1.0, 0.0])
            if np.linalg.norm(T) < 1e-6:
                continue
            T /= np.linalg.norm(T); B = np.cross(L, T)
            for _r in range(8):
                v = L + rng.random() * T * fs + rng.random() * B * fs
                if int(lt["type"]) != S.DIRECTIONAL:
                    v = v / np.linalg.norm(v)
                oc = o - centres
                a = float(v @ v); b = 2.0 * (oc @ v); cc = np.einsum("ij,ij->i", oc, oc) - radii ** 2
                disc = b * b - 4 * a * cc
                with np.errstate(invalid="ignore"):
                    t = (-b - np.sqrt(disc)) / (2 * a)
                hit = (disc >= 0) & (t > 0) & (t < lim)
                for s in slot_ids[hit]:
                    assert int(s) in cand, (cfg, j, "slot", int(s), "not in cell", c)
                checked += int(hit.sum())
        assert checked > 50, "the sampled jobs exercise real occluders"
    assert gridded >= 2


def test_small_scenes_and_wide_cones_get_no_grid():
    assert _grids(synthetic.config(2)) is None                       # 4 objects: the small-scene kernel scans them
    w = synthetic.config(5)
    w.lights["shadowSoftness"][:] = 6.0                              # sin bound 0.0606 > 0.05: too wide for a grid
    assert _grids(w) is None
