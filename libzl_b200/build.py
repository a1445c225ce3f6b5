"""Builds libzl_b200/libzl_b200.so (hand-written CUDA for sm_100a) in-tree with nvcc."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["zl_api.cu", "zl_render.cu", "zl_thumb.cu", "zl_wav.cu", "zl_sched.cpp"]
HEADERS = ["zl_device.cuh", "zl_voice.cuh", os.path.join("..", "..", "include", "zl_b200.h")]
OUT = os.path.join(HERE, "libzl_b200.so")
# host side above the C ABI: the reference's C++ classes (include/zl/*.h) and its libzl.h-compatible C symbols, plain g++
HOST_SOURCES = [os.path.join(HERE, "host", s) for s in ("libzl_facade.cpp", "libzl_capi.cpp")]
HOST_OUT = os.path.join(HERE, "libzl_b200_host.so")
# -fmad=false: no a*b+c contraction anywhere (the scalar audio math uses _rn intrinsics already; the packed f32x2
# intrinsics __fmul2_rn/__fadd2_rn would otherwise be fused into FFMA2 and break bit-exactness); explicit __fma_rn stays an FMA
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "shared"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)] + HOST_SOURCES \
        + [os.path.join(HERE, "..", "include", "zl", "libzl_facade.h")]
    if not os.path.exists(HOST_OUT):
        return True
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=None):
    """defines/out: kernel-variant experiments (e.g. defines=["ZL_KT=128"], out="gpurun_out/v1.so")."""
    out = out or OUT
    if out == OUT and not force and not needs_build():
        return OUT
    cmd = ([_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-D" + d for d in defines]
           + ["-o", out] + [os.path.join(CSRC, s) for s in SOURCES])
    subprocess.check_call(cmd)
    if out == OUT:
        build_host()
    return out


def build_host():
    """libzl_b200_host.so: the facade classes over the C ABI (links libzl_b200.so from its own directory)."""
    cxx = shutil.which("g++") or "g++"
    subprocess.check_call([cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wextra", "-fvisibility=hidden",
                           "-fvisibility-inlines-hidden", "-DZL_FACADE_BUILD", "-o", HOST_OUT] + HOST_SOURCES
                          + ["-L" + HERE, "-l:libzl_b200.so", "-Wl,-rpath,$ORIGIN"])
    return HOST_OUT


if __name__ == "__main__":
    build(force=True, verbose=True)
    print(OUT)
