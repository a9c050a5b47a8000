#!/usr/bin/env python3
"""make_ref.py — TEST INFRASTRUCTURE: build oracle/_ref/libref_shader.so.

Compiles the reference's OWN implementation of the hot path — the GLSL source
/root/reference/shader/raytracingCs.glsl, read where it lies — as C++ against
oracle/glsl_shim.h, after a purely mechanical, rule-based rewrite.
The rewritten text goes to a temp directory and is deleted; only the .so lands in
oracle/_ref/ (git-ignored, but it travels to the GPU box). No reference source is
copied into the repository.

Rewrite rules (each a regex; nothing is re-derived or re-ordered):
  1. drop `#version` and `layout(local_size...) in;`
  2. `layout(std430, binding=N) buffer X { T xs[]; };`  ->  `T* xs;`
  3. strip `layout(...)` and `uniform` qualifiers  ->  plain globals (defaults kept)
  4. `out T x` / `inout T x` parameters  ->  `T& x`
  5. append `f` to every floating literal (all arithmetic stays FP32, as in GLSL)
  6. swizzles `.xy` `.rg` `.rgb`  ->  method calls
  7. `imageSize(outputImage)` -> `imageSize_(outputImage)` (a local shadows the builtin)
  8. `void main()` -> `void shader_main()`
  9. `#define MAX_RAY_DEPTH n` -> `int MAX_RAY_DEPTH = n;` (run-time)

The fragment shaders of the stages either side of the path (bloom, TAA, equirect -> cube; built
into oracle/_ref/libref_post.so against post_shim.h + post_harness.cpp) take the same rules plus:
 10. global-scope `in T x;` / `out T x;`  ->  `thread_local T x;` (stage inputs / outputs)
 11. `T name[n] = T[] (a, b, ...);`  ->  `T name[n] = {a, b, ...};`
 12. `ivec2(gl_FragCoord)`  ->  `to_ivec2(gl_FragCoord)` (a vec4 -> ivec2 constructor)
"""
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SHADER = os.environ.get("RT_REFERENCE_SHADER", "/root/reference/shader/raytracingCs.glsl")
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "libref_shader.so")

FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")


def rewrite(src: str) -> str:
    # strip comments first so no rule fires inside one (the comments are not code)
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    # rule 1
    src = re.sub(r"^\s*#version[^\n]*\n", "\n", src, flags=re.M)
    src = re.sub(r"layout\s*\(\s*local_size[^)]*\)\s*in\s*;", "", src)
    # rule 9
    src = re.sub(r"^\s*#define\s+MAX_RAY_DEPTH\s+(\d+)\s*$", r"int MAX_RAY_DEPTH = \1;", src, flags=re.M)
    # rule 2
    src = re.sub(r"layout\s*\(\s*std430[^)]*\)\s*buffer\s+\w+\s*\{\s*(\w+)\s+(\w+)\s*\[\s*\]\s*;\s*\}\s*;",
                 r"\1* \2;", src)
    # rule 3
    src = re.sub(r"layout\s*\([^)]*\)\s*", "", src)
    src = re.sub(r"^\s*uniform\s+", "", src, flags=re.M)
    # rule 4
    src = re.sub(r"\b(?:inout|out)\s+(\w+)\s+(\w+)", r"\1& \2", src)
    # rule 7 (before rule 6 so the identifier is intact)
    src = src.replace("imageSize(outputImage)", "imageSize_(outputImage)")
    # rule 6
    src = re.sub(r"\.(xy|rg|rgb)\b(?!\s*\()", r".\1()", src)
    # rule 8
    src = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", src)
    # rule 5
    src = FLOAT_LIT.sub(lambda m: m.group(1) + "f", src)
    return src


def rewrite_fragment(src: str) -> str:
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"layout\s*\([^)]*\)\s*", "", src)                                       # rule 3 (first half)
    src = re.sub(r"^\s*(?:in|out)\s+(\w+)\s+(\w+)\s*;", r"thread_local \1 \2;", src, flags=re.M)   # rule 10
    src = re.sub(r"(\w+)\s+(\w+)\s*\[\s*(\d+)\s*\]\s*=\s*\1\s*\[\s*\]\s*\(([^)]*)\)\s*;",
                 r"\1 \2[\3] = {\4};", src)                                              # rule 11
    src = src.replace("ivec2(gl_FragCoord)", "to_ivec2(gl_FragCoord)")                    # rule 12
    return rewrite(src)


POST_SHADERS = {                      # macro -> file under /root/reference/shader
    "EXTRACT_INC": "brightness_extractFS.glsl",
    "BLUR_INC": "gaussian_blurFs.glsl",
    "COMBINE_INC": "bloom_combineFs.glsl",
    "TAA_INC": "taaFs.glsl",
    "SKYBOX_INC": "skyboxFs.glsl",
}
OUT_POST_SO = os.path.join(OUT_DIR, "libref_post.so")


def post_sources() -> list:
    d = os.path.dirname(REF_SHADER)
    return [os.path.join(d, f) for f in POST_SHADERS.values()]


def build_post(verbose: bool = False) -> str:
    for f in post_sources():
        if not os.path.isfile(f):
            raise FileNotFoundError(f"reference shader not found: {f}")
    os.makedirs(OUT_DIR, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="rtref_") as tmp:
        defs = []
        for (macro, name), path in zip(POST_SHADERS.items(), post_sources()):
            with open(path, encoding="utf-8", errors="replace") as f:
                text = f.read()
            inc = os.path.join(tmp, name + ".inc")
            with open(inc, "w") as f:
                f.write(rewrite_fragment(text))
            defs.append(f'-D{macro}="{inc}"')
        cmd = ["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared",
               "-fno-fast-math", "-Wno-unused-variable"] + defs + ["-I", HERE,
               os.path.join(HERE, "post_harness.cpp"), "-o", OUT_POST_SO]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)
    return OUT_POST_SO


def build(verbose: bool = False) -> str:
    if not os.path.isfile(REF_SHADER):
        raise FileNotFoundError(f"reference shader not found: {REF_SHADER}")
    with open(REF_SHADER, encoding="utf-8", errors="replace") as f:
        text = f.read()
    os.makedirs(OUT_DIR, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="rtref_") as tmp:
        inc = os.path.join(tmp, "shader.inc")
        with open(inc, "w") as f:
            f.write(rewrite(text))
        cmd = ["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared",
               "-fno-fast-math", "-Wno-unused-variable", "-Wno-maybe-uninitialized",
               f'-DREF_SHADER_INC="{inc}"', "-I", HERE,
               os.path.join(HERE, "ref_harness.cpp"), "-o", OUT_SO]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)
    return OUT_SO


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
    print(build_post(verbose="-v" in sys.argv))
