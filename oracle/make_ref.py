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
