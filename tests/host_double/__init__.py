"""TEST INFRASTRUCTURE ONLY: builds and loads tests/_build/libaps_host_double.so = the product's host layer
(alluxio_b200/host/cache.cc, unchanged) linked against fake_aps.cc, a host-memory stand-in for the libaps.so entry
points that layer calls.  It lets `pytest -m "not gpu"` run the host layer's logic (manager, evictors, stream, time-bound
store, sharded routing, replays) against the oracle on a machine without a GPU.  Nothing under alluxio_b200/ knows about
it; the product libraries have no CPU path."""
import contextlib
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT_DIR = os.path.join(os.path.dirname(HERE), "_build")
LIB = os.path.join(OUT_DIR, "libaps_host_double.so")
SOURCES = [os.path.join(ROOT, "alluxio_b200", "host", "cache.cc"), os.path.join(ROOT, "alluxio_b200", "host", "cache.h"),
           os.path.join(HERE, "fake_aps.cc"), os.path.join(ROOT, "include", "aps.h"), os.path.join(ROOT, "include", "aps_cache.h")]


def build():
    if os.path.exists(LIB) and all(os.path.getmtime(s) <= os.path.getmtime(LIB) for s in SOURCES):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    # -Bsymbolic + -z defs: every aps_* the host layer calls is bound to the double inside this one library (a missing
    # entry point fails the link here), whatever else the process has loaded.
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wall", "-Wextra",
                           "-pthread", "-I", os.path.join(ROOT, "include"), "-o", LIB, SOURCES[0], SOURCES[2],
                           "-Wl,-Bsymbolic", "-Wl,-z,defs"])
    return LIB


_lib = None


def lib():
    """The double library with the argument types of alluxio_b200/cache.py applied."""
    global _lib
    if _lib is None:
        from alluxio_b200 import cache
        L = C.CDLL(build())
        for name, (res, args) in cache.SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        L.aps_shard_of.restype, L.aps_shard_of.argtypes = C.c_uint32, [C.c_void_p, C.c_uint32]
        _lib = L
    return _lib


class _CapiOverDouble:
    """What alluxio_b200.capi.lib() returns inside the block below: the double's entry points under the argument types
    capi declares for the real ones; anything the double does not have (every kernel-backed call) raises."""

    def __init__(self, real, double):
        self._real, self._double = real, double

    def __getattr__(self, name):
        try:
            fn = getattr(self._double, name)
        except AttributeError:
            raise AttributeError("%s is not part of the host-layer double: a GPU-only entry point" % name) from None
        real = getattr(self._real, name)
        fn.restype, fn.argtypes = real.restype, real.argtypes
        return fn


@contextlib.contextmanager
def host_layer_over_double():
    """Inside the block, the Python mirrors (alluxio_b200/cache.py, page_store.py) talk to the double library."""
    from alluxio_b200 import cache, capi
    saved, saved_capi = cache._lib, capi.lib
    proxy = _CapiOverDouble(capi.lib(), lib())
    cache._lib = lib()
    capi.lib = lambda: proxy
    try:
        yield cache
    finally:
        cache._lib, capi.lib = saved, saved_capi
