"""Host-side scene layer: the producers of the hot path's inputs.

Mirrors, in the reference's own vocabulary:
  * the std430 data contracts  Object / Material / AABB / Light
        (/root/reference/src/Object.h:8-21, src/Material.h:11-23, src/Light.h:7-20)
  * ``SceneIO::Load`` / ``ParseObject`` / ``ParseLight`` (src/SceneIO.h:108-122,145-186),
    including the C++ ``istream >>`` extraction semantics the shipped scene files rely on
    (a malformed field stores 0 and sets failbit; a missing field, or any field after a
    failure, keeps its default)
  * ``GenerateAABBForObject`` (src/SceneIO.h:75-104), evaluated in FP32 exactly as glm does
  * ``Camera`` defaults and ``UpdateVectors`` (src/Camera.h:9-34)

Nothing here touches the GPU; the arrays produced are handed to ``rt_upload_scene``
byte-for-byte.
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass, field

import numpy as np

F32 = np.float32

# --------------------------------------------------------------------------- dtypes
MATERIAL_DTYPE = np.dtype({
    "names": ["type", "albedo", "metallic", "roughness", "diffuseStrength", "ior", "transparency",
              "specular", "subsurfaceScatter", "subsurfaceColor", "scatterDistance"],
    "formats": ["<i4", ("<f4", 3), "<f4", "<f4", "<f4", "<f4", "<f4", "<f4", "<f4", ("<f4", 3), "<f4"],
    "offsets": [0, 16, 28, 32, 36, 40, 44, 48, 52, 64, 76],
    "itemsize": 80,
})
AABB_DTYPE = np.dtype({
    "names": ["min", "max"], "formats": [("<f4", 3), ("<f4", 3)], "offsets": [0, 16], "itemsize": 32,
})
OBJECT_DTYPE = np.dtype({
    "names": ["type", "position", "radius", "normal", "size", "material", "bounds"],
    "formats": ["<i4", ("<f4", 3), "<f4", ("<f4", 3), ("<f4", 2), MATERIAL_DTYPE, AABB_DTYPE],
    "offsets": [0, 16, 28, 32, 48, 64, 144],
    "itemsize": 176,
})
LIGHT_DTYPE = np.dtype({
    "names": ["type", "position", "direction", "color", "intensity", "radius", "samples",
              "shadowSoftness", "shadowType", "pcfSamples", "lightSize", "angularRadius"],
    "formats": ["<i4", ("<f4", 3), ("<f4", 3), ("<f4", 3), "<f4", "<f4", "<i4", "<f4", "<i4", "<i4", "<f4", "<f4"],
    "offsets": [0, 16, 32, 48, 60, 64, 68, 72, 76, 80, 84, 88],
    "itemsize": 96,
})
assert OBJECT_DTYPE.itemsize == 176 and LIGHT_DTYPE.itemsize == 96 and MATERIAL_DTYPE.itemsize == 80

SPHERE, PLANE = 0, 1                      # enum class ObjectType (src/Object.h:6)
POINT, DIRECTIONAL, AREA = 0, 1, 2        # enum class LightType  (src/Light.h:5)
MATERIAL_METALLIC, MATERIAL_DIELECTRIC, MATERIAL_PLASTIC = 0, 1, 2   # src/Material.h:5-9
SHADOW_NONE, SHADOW_PCF, SHADOW_PCSS = 0, 1, 2                        # src/Light.h:16


def new_objects(n: int) -> np.ndarray:
    """n Objects carrying the C++ default member initialisers (src/Object.h:13-21,
    src/Material.h:11-23). ``diffuseStrength`` has no initialiser in the reference
    (stack garbage); it is pinned to 0 here."""
    o = np.zeros(n, dtype=OBJECT_DTYPE)
    o["radius"] = 1.0
    o["normal"] = (0.0, 1.0, 0.0)
    o["size"] = (1.0, 1.0)
    m = o["material"]
    m["type"] = MATERIAL_PLASTIC
    m["albedo"] = 1.0
    m["roughness"] = 0.5
    m["ior"] = 1.0
    m["specular"] = 0.5
    m["subsurfaceColor"] = 1.0
    m["scatterDistance"] = 0.1
    return o


def new_lights(n: int) -> np.ndarray:
    """n Lights carrying the defaults of src/Light.h:7-20; ``angularRadius`` (no
    initialiser, dead in the shader) is pinned to 0."""
    l = np.zeros(n, dtype=LIGHT_DTYPE)
    l["direction"] = (0.0, -1.0, 0.0)
    l["color"] = 1.0
    l["intensity"] = 1.0
    l["radius"] = 0.5
    l["samples"] = 4
    l["shadowSoftness"] = 1.0
    l["shadowType"] = SHADOW_PCF
    l["pcfSamples"] = 4
    l["lightSize"] = 1.0
    return l


# --------------------------------------------------------------------------- glm in FP32
def _dot3(a, b):
    t = (a * b).astype(F32)
    return F32(F32(t[0] + t[1]) + t[2])          # glm: tmp.x + tmp.y + tmp.z


def _cross(x, y):
    return np.array([F32(F32(x[1] * y[2]) - F32(y[1] * x[2])),
                     F32(F32(x[2] * y[0]) - F32(y[2] * x[0])),
                     F32(F32(x[0] * y[1]) - F32(y[0] * x[1]))], dtype=F32)


def _normalize(v):
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = F32(F32(1.0) / np.sqrt(_dot3(v, v), dtype=F32))   # glm::inversesqrt
        return (v * inv).astype(F32)


def generate_aabb_for_object(obj: np.void) -> None:
    """``GenerateAABBForObject`` (src/SceneIO.h:75-104), in place, FP32 like glm.

    Sphere: centre -/+ radius. Plane: the in-plane rectangle, then BOTH corners shifted by
    +0.01*normal — a zero-thickness box 1 cm off the plane, with min > max on some axes
    for some orientations. The shader's slab test (glsl:91-103) tolerates both."""
    pos = obj["position"].astype(F32)
    if obj["type"] == SPHERE:
        r = F32(obj["radius"])
        obj["bounds"]["min"] = pos - r
        obj["bounds"]["max"] = pos + r
    elif obj["type"] == PLANE:
        n = obj["normal"].astype(F32)
        if abs(float(n[1])) > 0.9:                       # double compare, as in C++
            right = np.array([1, 0, 0], dtype=F32)
            forward = np.array([0, 0, 1], dtype=F32)
        else:
            right = _normalize(_cross(n, np.array([0, 1, 0], dtype=F32)))
            forward = _normalize(_cross(right, n))
        size = obj["size"].astype(F32)
        hx = (right * F32(size[0] / F32(2.0))).astype(F32)
        hy = (forward * F32(size[1] / F32(2.0))).astype(F32)
        bmin = ((pos - hx).astype(F32) - hy).astype(F32)
        bmax = ((pos + hx).astype(F32) + hy).astype(F32)
        shift = (n * F32(0.01)).astype(F32)
        obj["bounds"]["min"] = (bmin + shift).astype(F32)
        obj["bounds"]["max"] = (bmax + shift).astype(F32)


def generate_aabbs(objects: np.ndarray) -> np.ndarray:
    for i in range(len(objects)):
        generate_aabb_for_object(objects[i])
    return objects


# --------------------------------------------------------------------------- SceneIO
_INT_RE = re.compile(r"[+-]?\d+")
_FLT_RE = re.compile(r"[+-]?(?:\d+\.?\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?)")


class _IStream:
    """Just enough of std::istringstream: ``>>`` for string / int / float with failbit."""

    def __init__(self, text: str):
        self.s, self.i, self.fail = text, 0, False

    def _skip(self):
        while self.i < len(self.s) and self.s[self.i].isspace():
            self.i += 1

    def word(self) -> str:
        if self.fail:
            return ""
        self._skip()
        j = self.i
        while j < len(self.s) and not self.s[j].isspace():
            j += 1
        if j == self.i:
            self.fail = True
            return ""
        w, self.i = self.s[self.i:j], j
        return w

    def _num(self, rx, conv, default):
        """Returns the extracted value. A field that is present but not a number stores 0
        (C++11 num_get) and sets failbit; when the line has simply run out (the sentry hits
        end-of-stream) or the stream had already failed, nothing is stored: ``default`` stays."""
        if self.fail:
            return default
        self._skip()
        if self.i >= len(self.s):
            self.fail = True
            return default
        m = rx.match(self.s, self.i)
        if not m:
            self.fail = True
            return conv(0)
        self.i = m.end()
        return conv(m.group(0))

    def int_(self, default=0):
        return self._num(_INT_RE, int, default)

    def float_(self, default=0.0):
        return self._num(_FLT_RE, float, default)


@dataclass
class Scene:
    """What SceneIO::Load fills: the SSBO vectors plus the UI names."""
    objects: np.ndarray = field(default_factory=lambda: new_objects(0))
    lights: np.ndarray = field(default_factory=lambda: new_lights(0))
    object_names: list = field(default_factory=list)
    light_names: list = field(default_factory=list)


_OBJ_TYPES = {"SPHERE": SPHERE, "PLANE": PLANE}            # StringToObjectType, default SPHERE
_LIGHT_TYPES = {"DIRECTIONAL": DIRECTIONAL, "AREA": AREA}   # StringToLightType, default POINT
_LIGHT_TYPES_V2 = {"POINT": POINT, "DIRECTIONAL": DIRECTIONAL, "AREA": AREA}


def _type_v2(word: str, names: dict, default: int) -> int:
    """Type token of a v2 record: a known name, or ``TYPE<n>`` for a tag the reference's enums do
    not name (the shader treats such an object as never hit and such a light as black, and both
    must survive a save/load cycle)."""
    if word in names:
        return names[word]
    if word.startswith("TYPE"):
        try:
            return int(word[4:])
        except ValueError:
            pass
    return default


def parse_scene(text: str) -> Scene:
    """``SceneIO::Load`` (src/SceneIO.h:108-122) on the text of a .scene file."""
    objs, lights, onames, lnames = [], [], [], []
    for line in text.splitlines():
        iss = _IStream(line)
        kind = iss.word()
        if kind == "OBJECT":                                   # ParseObject :145-170
            o = new_objects(1)
            type_str, name = iss.word(), iss.word()
            o["position"][0] = [iss.float_(), iss.float_(), iss.float_()]
            o["type"] = _OBJ_TYPES.get(type_str, SPHERE)
            o["radius"] = iss.float_(float(o["radius"][0]))
            o["normal"][0] = [iss.float_(float(o["normal"][0][k])) for k in range(3)]
            o["size"][0] = [iss.float_(float(o["size"][0][k])) for k in range(2)]
            m = o["material"]
            m["type"] = iss.int_(0)   # `int matType;` is uninitialised in the reference; pinned 0
            m["albedo"][0] = [iss.float_(float(m["albedo"][0][k])) for k in range(3)]
            for f in ("metallic", "roughness", "ior", "transparency", "specular"):
                m[f] = iss.float_(float(m[f][0]))
            generate_aabb_for_object(o[0])
            objs.append(o)
            onames.append(name)
        elif kind == "LIGHT":                                  # ParseLight :172-186
            l = new_lights(1)
            type_str, name = iss.word(), iss.word()
            l["type"] = _LIGHT_TYPES.get(type_str, POINT)
            for f in ("position", "direction", "color"):
                l[f][0] = [iss.float_(float(l[f][0][k])) for k in range(3)]
            l["intensity"] = iss.float_(float(l["intensity"][0]))
            l["radius"] = iss.float_(float(l["radius"][0]))
            l["samples"] = iss.int_(int(l["samples"][0]))
            lights.append(l)
            lnames.append(name)
        elif kind == "OBJECT2":                                # the lossless record (save_scene_v2)
            o = new_objects(1)
            type_str, name = iss.word(), iss.word()
            o["type"] = _type_v2(type_str, _OBJ_TYPES, SPHERE)
            o["position"][0] = [iss.float_(float(o["position"][0][k])) for k in range(3)]
            o["radius"] = iss.float_(float(o["radius"][0]))
            o["normal"][0] = [iss.float_(float(o["normal"][0][k])) for k in range(3)]
            o["size"][0] = [iss.float_(float(o["size"][0][k])) for k in range(2)]
            m = o["material"]
            m["type"] = iss.int_(int(m["type"][0]))
            m["albedo"][0] = [iss.float_(float(m["albedo"][0][k])) for k in range(3)]
            for f in ("metallic", "roughness", "diffuseStrength", "ior", "transparency", "specular", "subsurfaceScatter"):
                m[f] = iss.float_(float(m[f][0]))
            m["subsurfaceColor"][0] = [iss.float_(float(m["subsurfaceColor"][0][k])) for k in range(3)]
            m["scatterDistance"] = iss.float_(float(m["scatterDistance"][0]))
            generate_aabb_for_object(o[0])
            objs.append(o)
            onames.append(name)
        elif kind == "LIGHT2":
            l = new_lights(1)
            type_str, name = iss.word(), iss.word()
            l["type"] = _type_v2(type_str, _LIGHT_TYPES_V2, POINT)
            for f in ("position", "direction", "color"):
                l[f][0] = [iss.float_(float(l[f][0][k])) for k in range(3)]
            l["intensity"] = iss.float_(float(l["intensity"][0]))
            l["radius"] = iss.float_(float(l["radius"][0]))
            l["samples"] = iss.int_(int(l["samples"][0]))
            l["shadowSoftness"] = iss.float_(float(l["shadowSoftness"][0]))
            l["shadowType"] = iss.int_(int(l["shadowType"][0]))
            l["pcfSamples"] = iss.int_(int(l["pcfSamples"][0]))
            l["lightSize"] = iss.float_(float(l["lightSize"][0]))
            l["angularRadius"] = iss.float_(float(l["angularRadius"][0]))
            lights.append(l)
            lnames.append(name)
    sc = Scene()
    sc.objects = new_objects(len(objs))      # (np.concatenate would drop the std430 padding)
    for i, o in enumerate(objs):
        sc.objects[i] = o[0]
    sc.lights = new_lights(len(lights))
    for i, l in enumerate(lights):
        sc.lights[i] = l[0]
    sc.object_names, sc.light_names = onames, lnames
    return sc


def load_scene(path: str) -> Scene:
    with open(path, encoding="utf-8", errors="replace") as f:
        return parse_scene(f.read())


def save_scene(scene: Scene) -> str:
    """``SceneIO::Save`` (src/SceneIO.h:124-142): the lossy text format (no diffuseStrength,
    SSS or shadow fields). Numbers use the default ostream format (%g, 6 significant)."""
    onames = {SPHERE: "SPHERE", PLANE: "PLANE"}
    lnames = {DIRECTIONAL: "DIRECTIONAL", AREA: "AREA"}
    g = lambda v: "%g" % float(v)
    out = []
    for i, o in enumerate(scene.objects):
        m = o["material"]
        name = scene.object_names[i] if i < len(scene.object_names) else "Object%d" % i
        vals = [*o["position"], o["radius"], *o["normal"], *o["size"]]
        tail = [*m["albedo"], m["metallic"], m["roughness"], m["ior"], m["transparency"], m["specular"]]
        out.append(" ".join(["OBJECT", onames.get(int(o["type"]), "UNKNOWN"), name] + [g(v) for v in vals]
                            + [str(int(m["type"]))] + [g(v) for v in tail]))
    for i, l in enumerate(scene.lights):
        name = scene.light_names[i] if i < len(scene.light_names) else "Light%d" % i
        vals = [*l["position"], *l["direction"], *l["color"], l["intensity"], l["radius"]]
        out.append(" ".join(["LIGHT", lnames.get(int(l["type"]), "POINT"), name] + [g(v) for v in vals]
                            + [str(int(l["samples"]))]))
    return "\n".join(out) + "\n"


def save_scene_v2(scene: Scene) -> str:
    """The lossless text format: one ``OBJECT2`` / ``LIGHT2`` line per record holding
    EVERY field the shader reads — the reference's ``OBJECT`` / ``LIGHT`` columns in their order,
    plus ``diffuseStrength``, the subsurface fields and the five shadow fields the reference's
    format drops (src/SceneIO.h:124-142). Floats are written with 9 significant
    digits, which round-trips every finite FP32 value; ``inf`` / ``nan`` are not representable
    in an istream extraction and are rejected. Bounds are not stored: ``GenerateAABBForObject``
    recomputes them on load, as the reference does for every object it touches
    (src/ImGUIManager.cpp:197). ``parse_scene`` and the C++ ``SceneIO::Parse`` read both record
    kinds, in any mix; the reference's own loader skips lines whose first token it does not know
    (src/SceneIO.h:114-119), so a v2 file is simply empty to it rather than mis-decoded."""
    onames = {SPHERE: "SPHERE", PLANE: "PLANE"}
    lnames = {POINT: "POINT", DIRECTIONAL: "DIRECTIONAL", AREA: "AREA"}

    def g(v) -> str:
        v = float(v)
        if not math.isfinite(v):
            raise ValueError("non-finite value in a scene record")
        return "%.9g" % v

    out = ["# rt-scene 2"]
    for i, o in enumerate(scene.objects):
        m = o["material"]
        name = scene.object_names[i] if i < len(scene.object_names) else "Object%d" % i
        vals = [*o["position"], o["radius"], *o["normal"], *o["size"]]
        tail = [*m["albedo"], m["metallic"], m["roughness"], m["diffuseStrength"], m["ior"], m["transparency"],
                m["specular"], m["subsurfaceScatter"], *m["subsurfaceColor"], m["scatterDistance"]]
        t = int(o["type"])
        out.append(" ".join(["OBJECT2", onames.get(t, "TYPE%d" % t), name] + [g(v) for v in vals]
                            + [str(int(m["type"]))] + [g(v) for v in tail]))
    for i, l in enumerate(scene.lights):
        name = scene.light_names[i] if i < len(scene.light_names) else "Light%d" % i
        t = int(l["type"])
        vals = [*l["position"], *l["direction"], *l["color"], l["intensity"], l["radius"]]
        out.append(" ".join(["LIGHT2", lnames.get(t, "TYPE%d" % t), name] + [g(v) for v in vals]
                            + [str(int(l["samples"])), g(l["shadowSoftness"]), str(int(l["shadowType"])),
                               str(int(l["pcfSamples"])), g(l["lightSize"]), g(l["angularRadius"])]))
    return "\n".join(out) + "\n"


# --------------------------------------------------------------------------- Camera
@dataclass
class Camera:
    """src/Camera.h:7-42 (the fields the dispatch site uploads, :158-162)."""
    position: tuple = (0.0, 0.0, 0.0)
    front: tuple = (0.0, 0.0, -1.0)
    up: tuple = (0.0, 1.0, 0.0)
    right: tuple = (1.0, 0.0, 0.0)
    fov: float = 45.0
    yaw: float = -90.0
    pitch: float = 0.0
    focal_length: float = 1.0

    def update_vectors(self) -> None:
        """``Camera::UpdateVectors`` (src/Camera.h:26-34) in FP32."""
        ry, rp = F32(math.radians(self.yaw)), F32(math.radians(self.pitch))
        front = np.array([F32(np.cos(ry)) * F32(np.cos(rp)), F32(np.sin(rp)),
                          F32(np.sin(ry)) * F32(np.cos(rp))], dtype=F32)
        front = _normalize(front)
        right = _normalize(_cross(front, np.array([0, 1, 0], dtype=F32)))
        up = _normalize(_cross(right, front))
        self.front, self.right, self.up = tuple(map(float, front)), tuple(map(float, right)), tuple(map(float, up))
