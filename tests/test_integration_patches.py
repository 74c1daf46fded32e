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


def _strip_java(text):
    """Java source without comments, string and char literals (enough for bracket counting)."""
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r'"(\\.|[^"\\])*"', '""', text)
    return re.sub(r"'(\\.|[^'\\])'", "''", text)


@pytest.mark.parametrize("name", ["GpuPageStore.java", "GpuPageStoreOptions.java"])
def test_java_sources_are_well_formed_as_far_as_can_be_told_without_a_compiler(name):
    """No JDK here: what can be checked is structure -- brackets balance and nest, every statement-looking line ends, every
    native the class CALLS is declared with that many parameters, and the JNI function takes the same number (+ env, cls)."""
    java = _strip_java(open(os.path.join(ROOT, "integration", "java", "alluxio", "client", "file", "cache", "store", name)).read())
    stack, pairs = [], {")": "(", "]": "[", "}": "{"}
    for ch in java:
        if ch in "([{":
            stack.append(ch)
        elif ch in pairs:
            assert stack and stack.pop() == pairs[ch], "unbalanced %s in %s" % (ch, name)
    assert not stack, "unclosed %s in %s" % (stack, name)
    if name != "GpuPageStore.java":
        return
    shim = open(os.path.join(ROOT, "integration", "jni", "aps_jni.c")).read()

    def arity(args):
        depth, n, seen = 0, 0, False
        for ch in args:
            if ch in "([{":
                depth += 1
            elif ch in ")]}":
                depth -= 1
            elif ch == "," and depth == 0:
                n += 1
            seen = seen or not ch.isspace()
        return n + 1 if seen else 0

    def call_args(text, start):
        depth, i = 1, start
        while depth:
            depth += {"(": 1, ")": -1}.get(text[i], 0)
            i += 1
        return text[start:i - 1]

    declared = {}
    for m in re.finditer(r"private static native [\w\[\]<>]+ (native\w+)\(", java):
        declared[m.group(1)] = arity(call_args(java, m.end()))
    for m in re.finditer(r"(?<!native )(?<![\w\]>] )\b(native\w+)\(", java):
        name_, n = m.group(1), arity(call_args(java, m.end()))
        assert name_ in declared, "%s is called but not declared" % name_
        assert n == declared[name_], "%s is called with %d arguments, declared with %d" % (name_, n, declared[name_])
    for m in re.finditer(r"JFN\((native\w+)\)\(", shim):
        assert arity(call_args(shim, m.end())) == declared[m.group(1)] + 2, "%s: JNI parameters do not match the declaration" % m.group(1)


def test_patched_reference_files_are_well_formed(tmp_path):
    """Applies both patches for real to copies of the files they touch and checks the results as far as can be done without
    a compiler: brackets balance, and the calls the patched LocalCacheManager makes into GpuPageStore have the arity of the
    methods GpuPageStore.java declares."""
    if not os.path.isdir(os.path.join(REFERENCE, "core")) or not shutil.which("patch"):
        pytest.skip("reference tree or patch(1) not available here")
    touched = set()
    for name in PATCHES:
        for line in open(os.path.join(ROOT, "integration", "patches", name)):
            if line.startswith("+++ b/"):
                touched.add(line[6:].split("\t")[0].strip())
    assert len(touched) == 9
    for rel in touched:
        os.makedirs(os.path.dirname(tmp_path / rel), exist_ok=True)
        shutil.copy(os.path.join(REFERENCE, rel), tmp_path / rel)
    for name in PATCHES:
        out = subprocess.run(["patch", "-p1", "-d", str(tmp_path), "-i", os.path.join(ROOT, "integration", "patches", name)],
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stdout + out.stderr
    pairs = {")": "(", "]": "[", "}": "{"}
    for rel in touched:
        java, stack = _strip_java(open(tmp_path / rel).read()), []
        for ch in java:
            if ch in "([{":
                stack.append(ch)
            elif ch in pairs:
                assert stack and stack.pop() == pairs[ch], "unbalanced %s in patched %s" % (ch, rel)
        assert not stack, "unclosed %s in patched %s" % (stack, rel)
    store = _strip_java(open(os.path.join(ROOT, "integration", "java", "alluxio", "client", "file", "cache", "store", "GpuPageStore.java")).read())
    manager = _strip_java(open(tmp_path / "core/client/fs/src/main/java/alluxio/client/file/cache/LocalCacheManager.java").read())

    def params(text, name):
        m = re.search(r"public [\w\[\]<>]+ %s\(" % name, text)
        depth, i = 1, m.end()
        while depth:
            depth += {"(": 1, ")": -1}.get(text[i], 0)
            i += 1
        return text[m.end():i - 1].count(",") + 1

    for method in ("getBatch", "allocPinned"):
        call = re.search(r"store\.%s\(([^;]*)\);" % method, manager)
        assert call, "the patched LocalCacheManager no longer calls GpuPageStore.%s" % method
        depth = n = 0
        for ch in call.group(1):
            depth += {"(": 1, ")": -1}.get(ch, 0)
            n += ch == "," and depth == 0
        assert n + 1 == params(store, method), "GpuPageStore.%s is called with %d arguments" % (method, n + 1)
    # the registration patch and the options class agree on the setters it calls
    options = open(os.path.join(ROOT, "integration", "java", "alluxio", "client", "file", "cache", "store", "GpuPageStoreOptions.java")).read()
    created = open(tmp_path / "core/client/fs/src/main/java/alluxio/client/file/cache/store/PageStoreOptions.java").read()
    gpu_case = created[created.index("case GPU:"):]
    gpu_case = gpu_case[:gpu_case.index("break;")]
    setters = re.findall(r"\.(set\w+)\(", gpu_case)
    assert len(setters) >= 4
    for setter in setters:
        assert re.search(r"public GpuPageStoreOptions %s\(" % setter, options), setter
    # every key the patch reads is defined by the patch
    keys = open(tmp_path / "core/common/src/main/java/alluxio/conf/PropertyKey.java").read()
    for key in set(re.findall(r"PropertyKey\.(USER_CLIENT_CACHE_GPU_\w+)", created)):
        assert re.search(r"public static final PropertyKey %s =" % key, keys) and re.search(r"public static final String %s =" % key, keys), key
