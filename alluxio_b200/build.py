"""Builds libaps.so (the C-ABI CUDA library) and libaps_host.so in-tree for sm_100a."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIBAPS = os.path.join(HERE, "libaps.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-shared",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_libaps(force=False, verbose=False):
    csrc = os.path.join(HERE, "csrc")
    sources = [os.path.join(csrc, f) for f in os.listdir(csrc)] + [os.path.join(ROOT, "include", "aps.h")]
    if force or _newer(LIBAPS, sources):
        cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
            "-I", os.path.join(ROOT, "include"), "-o", LIBAPS] + sorted(
                os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cu"))
        subprocess.check_call(cmd)
    return LIBAPS


LIBHOST = os.path.join(HERE, "libaps_host.so")


def build_host(force=False):
    """g++: the host-side cache layer (alluxio_b200/host) on top of libaps.so."""
    build_libaps()
    host = os.path.join(HERE, "host")
    sources = [os.path.join(host, f) for f in os.listdir(host)] + [os.path.join(ROOT, "include", "aps_cache.h"), LIBAPS]
    if force or _newer(LIBHOST, sources):
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wall", "-Wextra", "-pthread",
               "-I", os.path.join(ROOT, "include"), "-o", LIBHOST, os.path.join(host, "cache.cc"),
               "-L", HERE, "-laps", "-Wl,-rpath,$ORIGIN"]
        subprocess.check_call(cmd)
    return LIBHOST


if __name__ == "__main__":
    print(build_libaps(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
