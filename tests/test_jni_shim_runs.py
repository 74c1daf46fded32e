"""CPU: the JNI shim (integration/jni/aps_jni.c) is not only type-checked against the ABI (tests/test_c_abi.py) but RUN:
every native of GpuPageStore.java is called the way the Java class calls it, over a mock JNIEnv whose array accessors
always copy (the strictest behaviour the JNI specification allows: results written into an array that is then released
with JNI_ABORT are lost, an unreleased array is counted) and over the store double of tests/host_double -- no JVM and no
GPU needed.  Built with AddressSanitizer + UBSan.  The driver is tests/jni_stub/run_shim.c."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(HERE, "_build")
SAN = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-omit-frame-pointer"]
INC = ["-I", os.path.join(HERE, "jni_stub", "nocuda"), "-I", os.path.join(HERE, "jni_stub"), "-I", os.path.join(ROOT, "include"),
       "-I", os.path.join(ROOT, "integration", "jni")]


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and "cannot find -l" in r.stderr:
        pytest.skip("this toolchain has no AddressSanitizer runtime")
    assert r.returncode == 0, r.stderr[-3000:]


def test_every_native_of_the_shim_runs_over_a_mock_jvm():
    os.makedirs(OUT, exist_ok=True)
    objs = []
    for src, extra in ((os.path.join(HERE, "jni_stub", "run_shim.c"), ["-std=c99", "-Wall", "-Wextra", "-Wno-unused-parameter", "-Werror"]),
                       (os.path.join(HERE, "jni_stub", "fake_env.c"), ["-std=c99", "-Wall", "-Wextra", "-Werror"]),
                       (os.path.join(ROOT, "oracle", "checksums.c"), ["-msse4.2"])):
        obj = os.path.join(OUT, "shim_" + os.path.basename(src) + ".o")
        _run(["gcc"] + SAN + extra + INC + ["-c", src, "-o", obj])
        objs.append(obj)
    exe = os.path.join(OUT, "run_shim")
    _run(["g++", "-std=c++17"] + SAN + ["-I", os.path.join(ROOT, "include"), "-o", exe] + objs +
         [os.path.join(HERE, "host_double", "fake_aps.cc")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120,
                       env={**os.environ, "ASAN_OPTIONS": "detect_leaks=1", "UBSAN_OPTIONS": "halt_on_error=1 print_stacktrace=1"})
    assert r.returncode == 0 and r.stdout.strip() == "ok", (r.stdout[-1000:], r.stderr[-5000:])
    assert "Sanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-5000:]
