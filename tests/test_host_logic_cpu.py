"""CPU: the host layer's logic (alluxio_b200/host/cache.cc -- LocalCacheManager, evictors, meta store,
LocalCacheFileInStream, TimeBoundPageStore, factory, sharded routing, replays) against the oracle, on a machine
without a GPU.

How: tests/host_double links the UNCHANGED host layer against a host-memory stand-in for the handful of libaps.so entry
points it calls (a test double: tests/host_double/fake_aps.cc, built into tests/_build/ and loaded by nothing else).
The bodies run here are the `-m gpu` tests themselves (tests/test_gpu_cache_manager.py, test_gpu_stream_factory.py,
test_gpu_sharded_manager.py, test_gpu_z_reference_manager.py,
test_gpu_z_host_layer_vs_oracle.py): the same assertions, the same oracle comparisons, over the double here and over the CUDA
store on the B200.  A case is left to the GPU run only when it drives the CUDA store directly (named below)."""
import importlib

import numpy as np
import pytest

import host_double
from oracle import binding as orc

# Cases that drive the CUDA store directly (device buffers, kernels) instead of going through the host layer.
GPU_ONLY = {
    "test_gpu_cache_manager": set(),
    "test_gpu_stream_factory": set(),
    "test_gpu_sharded_manager": set(),
    "test_gpu_z_reference_manager": set(),
    "test_gpu_z_host_layer_vs_oracle": set(),
    # the PageStore contract cases (PageStoreTest ports, random ops against the oracle's LocalPageStore): over the double
    # they show that the stand-in itself follows the contract the GPU store is tested against
    "test_gpu_page_store": {"test_crc_detects_a_flipped_byte"},   # flips a byte in the HBM slab
}


def _param_sets(fn):
    """The argument sets of a function's own @pytest.mark.parametrize marks."""
    sets = [{}]
    for mark in getattr(fn, "pytestmark", []):
        if mark.name != "parametrize":
            continue
        names = [n.strip() for n in mark.args[0].split(",")]
        sets = [{**base, **dict(zip(names, v if len(names) > 1 else (v,)))} for base in sets for v in mark.args[1]]
    return sets


def _cases():
    out = []
    for mod_name, skip in GPU_ONLY.items():
        mod = importlib.import_module(mod_name)
        for name in sorted(vars(mod)):
            fn = getattr(mod, name)
            if not (name.startswith("test_") and callable(fn)) or name in skip or getattr(fn, "__module__", mod_name) != mod_name:
                continue
            for params in _param_sets(fn):
                tag = "-".join("".join(map(str, v)) if isinstance(v, list) else str(v) for v in params.values())
                out.append(pytest.param(mod_name, name, params, id="%s::%s%s" % (mod_name, name, "[%s]" % tag if tag else "")))
    return out


def _restore_cases():
    """Cases that walk the store's listing (restore, getPages): run them again under the double's other iteration order.
    The GPU store walks its hash table, so no case may depend on the order pages are listed in."""
    return [p for p in _cases() if any(w in p.values[1] for w in ("restore", "load_cache", "get_pages", "populated"))]


@pytest.mark.parametrize("mod_name,name,params", _restore_cases())
def test_listing_order_does_not_matter(mod_name, name, params, request, monkeypatch):
    monkeypatch.setenv("APS_DOUBLE_ITERATE", "reverse")
    test_gpu_case_over_the_double(mod_name, name, params, request)


@pytest.mark.parametrize("mod_name,name,params", _cases())
def test_gpu_case_over_the_double(mod_name, name, params, request):
    fn = getattr(importlib.import_module(mod_name), name)
    kwargs = dict(params)
    with host_double.host_layer_over_double():
        for arg in fn.__code__.co_varnames[:fn.__code__.co_argcount]:
            if arg in kwargs:
                continue
            if arg == "store":                                   # the fixture of test_gpu_page_store.py, made over the double
                from alluxio_b200.page_store import GpuPageStore
                kwargs[arg] = GpuPageStore(page_size=1024, cache_size=65536, slab_slots=256)
            else:
                kwargs[arg] = request.getfixturevalue(arg)
        fn(**kwargs)
        if "store" in kwargs:
            kwargs["store"].close()


def test_the_double_routes_like_the_library():
    """The double restates the library's shard hash; libaps.so's own aps_shard_of is host code and runs here."""
    import ctypes as C
    from alluxio_b200 import capi
    real, fake = capi.lib(), host_double.lib()
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 4, 8):
        for _ in range(500):
            hi, lo, idx = (int(x) for x in rng.integers(0, 2**63, 3))
            k = capi.aps_key(hi, lo, idx)
            assert real.aps_shard_of(C.byref(k), n) == fake.aps_shard_of(C.byref(k), n)
