"""The patches of integration/patches/ must apply to the reference tree they were cut against.  Runs only where that tree
is present (the build container); `patch --dry-run` writes nothing.  Also checks that every native method the Java class
declares has its JNI function in the shim, with the name JNI derives from the class."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = "/root/reference"
PATCHES = sorted(f for f in os.listdir(os.path.join(ROOT, "integration", "patches")) if f.endswith(".patch"))


@pytest.mark.parametrize("name", PATCHES)
def test_patch_applies_to_the_reference_tree(name):
    if not os.path.isdir(os.path.join(REFERENCE, "core")) or not shutil.which("patch"):
        pytest.skip("reference tree or patch(1) not available here")
    out = subprocess.run(["patch", "--dry-run", "-p1", "-d", REFERENCE, "-i", os.path.join(ROOT, "integration", "patches", name)],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "FAILED" not in out.stdout and "fuzz" not in out.stdout, out.stdout


def test_every_native_method_has_its_jni_function():
    java = open(os.path.join(ROOT, "integration", "java", "alluxio", "client", "file", "cache", "store", "GpuPageStore.java")).read()
    shim = open(os.path.join(ROOT, "integration", "jni", "aps_jni.c")).read()
    natives = re.findall(r"private static native [\w\[\]<>]+ (native\w+)\(", java)
    assert len(natives) >= 15
    for n in natives:
        assert re.search(r"JFN\(%s\)" % n, shim), "no JNI function for %s" % n
    # and nothing in the shim that the class does not declare
    for n in re.findall(r"JFN\((native\w+)\)", shim):
        assert n in natives, "%s has no declaration in GpuPageStore.java" % n
