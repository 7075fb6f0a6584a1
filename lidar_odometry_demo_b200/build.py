"""Builds liblo_b200.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc. No torch involved."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "liblo_b200.so")
REPLAY = os.path.join(HERE, "lo_replay")   # headless driver (apps/lo_replay.cpp): host C++ over the C ABI
SOURCES = ["lo_api.cu", "lo_grid.cu", "lo_robin.cu", "lo_match.cu", "lo_points.cu", "lo_planefit.cu", "lo_odometry.cu", "lo_pipeline.cu", "lo_sharded.cu"]
# -fmad=false: no FMA contraction, the reference is unfused x86-64 arithmetic (DESIGN.md "numerics")
COMPILE_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
                 "-Xcompiler", "-fPIC,-O3,-ffp-contract=off"]
LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "--shared", "-cudart", "shared", "-ldl"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "lo_b200.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


def build_apps(force=False):
    """g++ only: the driver sees nothing but include/lo_b200.h and the header-only host facade."""
    deps = [os.path.join(HERE, "apps", "lo_replay.cpp"), os.path.join(HERE, "host", "lo_io.hpp"), os.path.join(HERE, "host", "lo_b200.hpp"), OUT]
    if not force and os.path.exists(REPLAY) and all(os.path.getmtime(d) <= os.path.getmtime(REPLAY) for d in deps):
        return REPLAY
    cmd = ["g++", "-std=c++20", "-O2", "-Wall", "-o", REPLAY, deps[0], f"-L{HERE}", "-l:liblo_b200.so", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("g++ failed on lo_replay")
    return REPLAY


def _stale(obj, src, headers):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src, __file__] + headers)


def build(force=False, verbose=False):
    """One nvcc per translation unit, all at once (objects cached under csrc/../build/), then one link."""
    if not force and not needs_build():
        build_apps()
        return OUT
    from concurrent.futures import ThreadPoolExecutor
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    extra = os.environ.get("LO_B200_NVCC_EXTRA", "").split()   # developer builds, e.g. -DLO_TILE_DEBUG
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + [os.path.join(HERE, "..", "include", "lo_b200.h")]
    tag = os.path.join(objdir, "flags.txt")
    flags_now = " ".join(COMPILE_FLAGS + extra)
    if not os.path.exists(tag) or open(tag).read() != flags_now:
        force = True

    def compile_one(name):
        src, obj = os.path.join(CSRC, name), os.path.join(objdir, name[:-3] + ".o")
        if not force and not _stale(obj, src, headers):
            return ""
        cmd = ["nvcc"] + COMPILE_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {name}\n{r.stdout}{r.stderr}")
        return r.stderr
    names = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    with ThreadPoolExecutor(max_workers=min(len(names), os.cpu_count() or 1)) as pool:
        logs = list(pool.map(compile_one, names))
    if verbose:
        sys.stderr.write("".join(logs))
    objs = [os.path.join(objdir, n[:-3] + ".o") for n in names]
    r = subprocess.run(["nvcc"] + LINK_FLAGS + ["-o", OUT] + objs, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    with open(tag, "w") as f:
        f.write(flags_now)
    build_apps(force=True)
    return OUT


if __name__ == "__main__":
    build(force=True, verbose="-v" in sys.argv)
    print(OUT)
