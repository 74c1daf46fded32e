"""The C-ABI library loads and exports every symbol include/aps.h declares (no GPU needed), and the
data path fails loudly -- never falls back -- when no CUDA device is present."""
import ctypes as C
import os
import re

import pytest

from alluxio_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "aps.h")).read()
    return sorted(set(re.findall(r"APS_API\s+[\w\s\*]+?\b(aps_\w+)\s*\(", text)))


def test_header_declares_the_surface():
    syms = declared_symbols()
    for must in ("aps_open", "aps_close", "aps_put", "aps_get", "aps_delete", "aps_iterate", "aps_get_batch",
                 "aps_put_batch", "aps_md5_batch", "aps_stats", "aps_strerror"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    L = C.CDLL(capi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), name


def test_binding_covers_every_declared_symbol():
    assert sorted(capi.SIGNATURES) == declared_symbols()


def test_host_only_entry_points():
    L = capi.lib()
    assert b"sm_100a" in L.aps_version()
    assert L.aps_strerror(capi.APS_ENOTFOUND) == b"page not found"
    assert L.aps_total_kernel_launches() >= 0


def test_open_rejects_bad_config_and_fails_loudly_without_gpu():
    import torch
    L = capi.lib()
    h = C.c_void_p()
    for bad_size in (0, (1 << 30) + 1):  # (any size in between is accepted: the reference's page size is a free byte count)
        bad = capi.aps_config(device=0, page_size=bad_size, cache_bytes=1 << 20)
        assert L.aps_open(C.byref(bad), C.byref(h)) == capi.APS_EINVAL
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the no-device branch cannot be shown here")
    cfg = capi.aps_config(device=0, page_size=4096, cache_bytes=1 << 20)
    rc = L.aps_open(C.byref(cfg), C.byref(h))
    assert rc == capi.APS_ECUDA and not h.value
    assert L.aps_last_cuda_error() != b""
    from alluxio_b200.page_store import GpuPageStore, IOExceptionJ
    with pytest.raises(IOExceptionJ):
        GpuPageStore(page_size=4096, cache_size=1 << 20)


def declared_cache_symbols():
    text = open(os.path.join(ROOT, "include", "aps_cache.h")).read()
    return sorted(set(re.findall(r"APS_API\s+[\w\s\*]+?\b(apc_\w+)\s*\(", text)))


def test_host_library_exports_and_binds_every_symbol_of_aps_cache_h():
    from alluxio_b200 import cache
    syms = declared_cache_symbols()
    assert "apc_manager_create" in syms and "apc_replay_batched" in syms and "apc_evictor_evict" in syms
    capi.lib()  # libaps.so first: the host library links against it
    L = C.CDLL(cache.LIB_PATH)
    for name in syms:
        assert hasattr(L, name), name
    assert sorted(cache.SIGNATURES) == syms


def test_product_has_no_path_through_the_oracle_or_the_test_double():
    """The checker and the double are test infrastructure: no source of the product names them, the host library's one
    dependency for the store entry points is libaps.so itself, and the CUDA library exports none of the double's state."""
    import subprocess
    offenders = []
    for top in ("alluxio_b200", "include", "integration"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if not f.endswith((".py", ".cc", ".cu", ".cuh", ".h", ".c", ".java")):
                    continue
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for line in text.splitlines():
                    code = line.split("//")[0].split("#")[0] if not line.lstrip().startswith(("*", "/*")) else ""
                    if re.search(r"\b(import\s+oracle|from\s+oracle|host_double|fake_aps|liboracle|orc_\w+\s*\()", code):
                        offenders.append((os.path.join(dirpath, f), line.strip()))
    assert not offenders, offenders
    from alluxio_b200 import cache
    needed = subprocess.run(["readelf", "-d", cache.LIB_PATH], capture_output=True, text=True).stdout
    assert "libaps.so" in needed and "oracle" not in needed and "double" not in needed
    undefined = subprocess.run(["nm", "-D", "--undefined-only", cache.LIB_PATH], capture_output=True, text=True).stdout
    assert "aps_open" in undefined and "aps_get_batch" in undefined   # resolved by libaps.so, not defined in the host library
