"""CPU: the host layer under ThreadSanitizer and under AddressSanitizer + UndefinedBehaviorSanitizer.

alluxio_b200/host/cache.cc (unchanged) is compiled with the sanitizer together with the store double and a multi-threaded
driver of its C API (tests/host_double/stress.cc): threads putting, getting and deleting pages of one cache under
pressure -- with synchronous puts, with the asynchronous write pool, behind TimeBoundPageStore -- streams of several
threads over one small cache (per page and batched), and the factory asked by many threads at once.  The driver checks
the invariants that hold for any interleaving (a get returns that page's bytes; meta store and page store agree once the
threads are done; the byte budget holds); the sanitizer reports data races, leaks, overflows and use after free.  This is
the race-detection leg of the test strategy for the layer that carries the reference's lock hierarchy
(CACHE/LocalCacheManager.java:172-203); the kernels' concurrency invariants are in tests/test_gpu_concurrency.py."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(HERE, "_build")
SOURCES = [os.path.join(HERE, "host_double", "stress.cc"), os.path.join(ROOT, "alluxio_b200", "host", "cache.cc"),
           os.path.join(HERE, "host_double", "fake_aps.cc")]
HEADERS = [os.path.join(ROOT, "alluxio_b200", "host", "cache.h"), os.path.join(ROOT, "include", "aps.h"),
           os.path.join(ROOT, "include", "aps_cache.h")]


def _build(sanitize):
    exe = os.path.join(OUT, "stress_" + sanitize.replace(",", "_"))
    if os.path.exists(exe) and all(os.path.getmtime(s) <= os.path.getmtime(exe) for s in SOURCES + HEADERS):
        return exe
    os.makedirs(OUT, exist_ok=True)
    cmd = ["g++", "-O1", "-g", "-std=c++17", "-fsanitize=" + sanitize, "-fno-omit-frame-pointer", "-pthread",
           "-I", os.path.join(ROOT, "include"), "-o", exe] + SOURCES
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and ("cannot find -l" in r.stderr or "unrecognized" in r.stderr):
        pytest.skip("this toolchain has no runtime for -fsanitize=%s" % sanitize)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


@pytest.mark.parametrize("sanitize,env", [
    ("thread", {"TSAN_OPTIONS": "halt_on_error=1 second_deadlock_stack=1"}),
    ("address,undefined", {"ASAN_OPTIONS": "detect_leaks=1", "UBSAN_OPTIONS": "halt_on_error=1 print_stacktrace=1"}),
])
def test_host_layer_is_clean_under(sanitize, env):
    exe = _build(sanitize)
    r = subprocess.run([exe, "8", "6000"], capture_output=True, text=True, timeout=600, env={**os.environ, **env})
    if "LeakSanitizer has encountered a fatal error" in r.stderr:  # no ptrace in this sandbox: everything but the leak check
        r = subprocess.run([exe, "8", "6000"], capture_output=True, text=True, timeout=600,
                           env={**os.environ, **env, "ASAN_OPTIONS": "detect_leaks=0"})
    if "unexpected memory mapping" in r.stderr or "ReExec" in r.stderr:  # a sanitizer runtime that cannot start on this kernel
        pytest.skip("the %s runtime cannot map its shadow memory here: %s" % (sanitize, r.stderr.strip().splitlines()[0][:120]))
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), (r.stdout[-2000:], r.stderr[-6000:])
    assert "Sanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-6000:]
