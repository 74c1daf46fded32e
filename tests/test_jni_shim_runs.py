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


STRICT = ["-std=c99", "-Wall", "-Wextra", "-Werror"]
MD5 = [(os.path.join(HERE, "host_double", "md5_via_oracle.c"), STRICT), (os.path.join(ROOT, "oracle", "checksums.c"), ["-msse4.2"])]


def _build_over_the_double(name, c_sources):
    """gcc for the C sources, g++ for the link with the store double; everything under ASan + UBSan."""
    os.makedirs(OUT, exist_ok=True)
    objs = []
    for src, extra in c_sources:
        obj = os.path.join(OUT, name + "_" + os.path.basename(src) + ".o")
        _run(["gcc"] + SAN + extra + INC + ["-c", src, "-o", obj])
        objs.append(obj)
    exe = os.path.join(OUT, name)
    _run(["g++", "-std=c++17"] + SAN + ["-I", os.path.join(ROOT, "include"), "-o", exe] + objs +
         [os.path.join(HERE, "host_double", "fake_aps.cc")])
    return exe


def _run_clean(exe):
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120,
                       env={**os.environ, "ASAN_OPTIONS": "detect_leaks=1", "UBSAN_OPTIONS": "halt_on_error=1 print_stacktrace=1"})
    if "LeakSanitizer has encountered a fatal error" in r.stderr:  # no ptrace in this sandbox: everything but the leak check
        r = subprocess.run([exe], capture_output=True, text=True, timeout=120,
                           env={**os.environ, "ASAN_OPTIONS": "detect_leaks=0", "UBSAN_OPTIONS": "halt_on_error=1 print_stacktrace=1"})
    assert r.returncode == 0, (r.stdout[-1000:], r.stderr[-5000:])
    assert "Sanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-5000:]
    return r.stdout


def test_every_native_of_the_shim_runs_over_a_mock_jvm():
    exe = _build_over_the_double("run_shim", [
        (os.path.join(HERE, "jni_stub", "run_shim.c"), STRICT + ["-Wno-unused-parameter"]),
        (os.path.join(HERE, "jni_stub", "fake_env.c"), STRICT)] + MD5)
    assert _run_clean(exe).strip() == "ok"


@pytest.mark.parametrize("example", ["example", "batch_example"])
def test_plain_c_callers_run_over_the_double(example):
    """integration/c/example.c (single-page calls) and batch_example.c (the batched JNI marshalling with a malloc'd
    "direct buffer") are run against libaps.so on a GPU by tests/test_c_abi.py; here the same programs run over the store
    double, under ASan + UBSan, so that their own logic is exercised in the CPU suite too."""
    exe = _build_over_the_double("c_" + example, [(os.path.join(ROOT, "integration", "c", example + ".c"), ["-std=c99", "-Wall"])] + MD5)
    out = _run_clean(exe)
    assert "ok" in out.lower() or "pass" in out.lower(), out[-500:]
