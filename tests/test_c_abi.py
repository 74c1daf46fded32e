"""The boundary is a C ABI ("plain C, callable from JNI and from a JVM-free harness"): both headers must
compile as strict C99, and a C program with no Python or C++ in it must link against libaps.so and, on a GPU, run."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "alluxio_b200")


def _gcc():
    cc = shutil.which("gcc")
    if not cc:
        pytest.skip("no gcc")
    return cc


def _build_example(tmp_path, name="example"):
    from alluxio_b200 import build
    build.build_libaps()
    exe = str(tmp_path / name)
    subprocess.run([_gcc(), "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                    "-I", os.path.join(ROOT, "integration", "jni"),
                    os.path.join(ROOT, "integration", "c", name + ".c"), "-L", LIBDIR, "-laps", "-Wl,-rpath," + LIBDIR,
                    "-o", exe], check=True, capture_output=True)
    return exe


def test_headers_are_strict_c99(tmp_path):
    src = tmp_path / "hdr.c"
    src.write_text('#include "aps.h"\n#include "aps_cache.h"\nint main(void) { return sizeof(aps_get_req) == 40 ? 0 : 1; }\n')
    subprocess.run([_gcc(), "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), "-c",
                    str(src), "-o", str(tmp_path / "hdr.o")], check=True, capture_output=True)


def test_c_example_links_against_libaps(tmp_path):
    assert os.path.exists(_build_example(tmp_path))


@pytest.mark.gpu
def test_c_example_runs(tmp_path):
    out = subprocess.run([_build_example(tmp_path)], capture_output=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == b"ok", out.stderr


def test_batched_shim_marshalling_links_against_libaps(tmp_path):
    assert os.path.exists(_build_example(tmp_path, "batch_example"))


@pytest.mark.gpu
def test_batched_shim_marshalling_runs_with_a_malloced_direct_buffer(tmp_path):
    """The batched JNI calls minus the JVM: integration/jni/aps_marshal.h (what aps_jni.c calls) driven by plain C with a
    malloc'd "direct buffer" -- put, get (offset read, absent id, short pages), MD5, delete, double close."""
    out = subprocess.run([_build_example(tmp_path, "batch_example")], capture_output=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == b"ok", out.stderr


def test_jni_shim_type_checks_against_the_abi(tmp_path):
    """The JNI shim of INTEGRATION.md cannot be built for real here (no JDK), but its calls into include/aps.h can be
    type-checked: compile it against a minimal stand-in jni.h."""
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime_api.h")):
        pytest.skip("no CUDA headers")
    subprocess.run([_gcc(), "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "tests", "jni_stub"), "-I",
                    os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "integration", "jni"), "-I", cuda_inc, "-c",
                    os.path.join(ROOT, "integration", "jni", "aps_jni.c"),
                    "-o", str(tmp_path / "aps_jni.o")], check=True, capture_output=True)
