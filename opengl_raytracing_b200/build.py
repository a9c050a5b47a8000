"""nvcc build of csrc/ -> csrc/librt_b200.so for sm_100a (in-tree, so the .so travels to
the GPU box). Two builds of the kernel TU (RT_FLAG_PARITY_MATH selects at run time), BOTH with
-fmad=false: the geometric chain (ray -> t -> P -> N -> next ray) must round exactly like the
shader's unfused arithmetic or mirror/glass paths diverge after a bounce. The fast build fuses
explicitly (fmaf) and uses SFU approximations only in colour-only shading code."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "librt_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-Xptxas", "-v"]
UNITS = [  # (source, extra flags)
    ("rt_kernel_fast.cu", ["-fmad=false"]),
    ("rt_kernel_parity.cu", ["-fmad=false"]),
    ("rt_regroup_fast_small.cu", ["-fmad=false"]),      # the regrouping megakernel: own units, own CTA sizes (rt_regroup_tu.inc)
    ("rt_regroup_fast_big.cu", ["-fmad=false"]),
    ("rt_regroup_parity_small.cu", ["-fmad=false"]),
    ("rt_regroup_parity_big.cu", ["-fmad=false"]),
    ("rt_abi.cu", []),
    ("rt_post.cu", ["-fmad=false"]),
    ("rt_frame.cu", ["-fmad=false"]),
    ("rt_peak.cu", []),
    ("rt_multi.cu", []),
]
HEADERS = ["rt_kernel.cuh", "rt_packet.cuh", "rt_wavefront.cuh", "rt_regroup.cuh", "rt_regroup_tu.inc", "rt_math.cuh", "rt_device.h", os.path.join("..", "..", "include", "rt_abi.h"),
           os.path.join("..", "..", "include", "rt_post.h"), os.path.join("..", "..", "include", "rt_frame.h"),
           os.path.join("..", "..", "include", "rt_multi.h")]


def _stale(target: str, deps: list) -> bool:
    if not os.path.isfile(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, variant: str = "", defines: tuple = ()) -> str:
    """variant/defines: dev-only A/B builds (csrc/librt_b200<variant>.so with extra -D flags,
    loaded through the RT_B200_LIB environment variable); the product build takes neither."""
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    objs, rebuilt = [], False
    lib = LIB if not variant else LIB[:-3] + variant + ".so"
    jobs = []
    for src, extra in UNITS:
        s = os.path.join(CSRC, src)
        o = s[:-3] + variant + ".o"
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append((src, o, [NVCC] + ARCH + COMMON + extra + [f"-D{d}" for d in defines] + ["-c", s, "-o", o]))

    def compile_one(job):
        src, o, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        with open(o + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        return src, r

    if jobs:                                  # the translation units compile side by side
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as ex:
            for src, r in ex.map(compile_one, jobs):
                if verbose or r.returncode != 0:
                    sys.stderr.write(r.stdout + r.stderr)
                if r.returncode != 0:
                    raise RuntimeError(f"nvcc failed on {src}")
        rebuilt = True
    if force or rebuilt or _stale(lib, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", lib] + objs
        subprocess.run(cmd, check=True)
    return lib


HOST_DIR = os.path.join(HERE, "..", "host")
HEADLESS = os.path.join(HOST_DIR, "rt_headless")


def build_host(force: bool = False) -> str:
    """host/rt_headless: the C++ host layer (host/rt_host.hpp) + the windowless frame loop, linked
    against csrc/librt_b200.so through an $ORIGIN-relative rpath (so it runs on the GPU box)."""
    src = [os.path.join(HOST_DIR, "rt_headless.cpp"), os.path.join(HOST_DIR, "rt_host.hpp"),
           os.path.join(HERE, "..", "include", "rt_abi.h"), LIB]
    if force or _stale(HEADLESS, src):
        cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-Wall", "-ffp-contract=off",
               "-I" + os.path.join(HERE, "..", "include"), "-I" + HOST_DIR, src[0], "-o", HEADLESS,
               "-L" + CSRC, "-lrt_b200", "-Wl,-rpath,$ORIGIN/../opengl_raytracing_b200/csrc"]
        subprocess.run(cmd, check=True)
    return HEADLESS


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("-")]
    print(build(force="-f" in sys.argv, verbose="-v" in sys.argv, variant=args[0] if args else "",
                defines=tuple(args[1:])))
    if not args:
        print(build_host(force="-f" in sys.argv))
