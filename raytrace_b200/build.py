"""Builds raytrace_b200/libvxrt.so (hand-written CUDA for sm_100a) with nvcc, in-tree.

    python raytrace_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  -lineinfo keeps ncu's source page mapped to the .cu files; the
host side is compiled with -ffp-contract=off because the per-frame constants computed on the host
(sun direction/colour, diffuse LUTs) must round exactly like the oracle's.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libvxrt.so")
HOST_OUT = os.path.join(HERE, "vxrt_main")
SOURCES = ["api.cu", "trace.cu", "trace_fast.cu", "trace_pool.cu", "trace_wave.cu", "post.cu", "temporal.cu", "world.cu", "comm.cu", "peer.cu", "chunk_store.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-shared"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def needs_build():
    if not os.path.exists(OUT) or not os.path.exists(HOST_OUT):
        return True
    t = min(os.path.getmtime(OUT), os.path.getmtime(HOST_OUT))
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    deps += [os.path.join(CSRC, "host", f) for f in os.listdir(os.path.join(CSRC, "host"))]
    deps = deps + [os.path.join(HERE, "..", "include", "vxrt.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_variant(tag, extra):
    """libvxrt_<tag>.so with extra nvcc flags (kernel-variant A/B runs; picked up through VXRT_LIB)."""
    out = os.path.join(HERE, f"libvxrt_{tag}.so")
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.run([_nvcc()] + NVCC_FLAGS + list(extra) + ["-o", out] + srcs + ["-ldl"], cwd=CSRC, check=True)
    return out


def build(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    if not force and not needs_build():
        return OUT
    extra = os.environ.get("VXRT_NVCC_EXTRA", "").split()
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + srcs + ["-ldl"]
    r = subprocess.run(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libvxrt.so")
    build_host(verbose)
    return OUT


def build_host(verbose=False):
    """C++ host driver above the C ABI (csrc/host/: the reference's render/game/world interfaces + a headless
    src/bin/main.rs), linked against the in-tree libvxrt.so."""
    cmd = ["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I", os.path.join(HERE, "..", "include"), "-o", HOST_OUT,
           os.path.join(CSRC, "host", "vxrt_main.cpp"), "-L", HERE, "-lvxrt", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("g++ failed building vxrt_main")
    return HOST_OUT


if __name__ == "__main__" and "--variant" in sys.argv:
    i = sys.argv.index("--variant")
    print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
    sys.exit(0)
if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(OUT)
