"""ctypes view of libaps_host.so: the native mirror of the reference's classes above the PageStore
boundary (include/aps_cache.h).  Names follow the Java classes (CACHE/LocalCacheManager.java,
CACHE/DefaultMetaStore.java, CACHE/CacheEvictor.java) so tests read like the reference's."""
import ctypes as C
import os

from . import capi
from .page_store import PageId  # noqa: F401  (re-export)

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaps_host.so")

_vp, _cp, _i32, _i64, _u64, _dbl = C.c_void_p, C.c_char_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
EXTERNAL_READ = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32)

class apc_options(C.Structure):
    _fields_ = [("store", capi.aps_config), ("evictor", C.c_char_p), ("lfu_logbase", C.c_double), ("n_shards", C.c_uint32),
                ("shard_devices", C.POINTER(C.c_int32)), ("over", C.c_void_p), ("async_write_threads", C.c_int32),
                ("async_restore", C.c_int32), ("timeout_ms", C.c_int64), ("timeout_threads", C.c_int32)]


SIGNATURES = {
    "apc_manager_create_ex": (C.c_int, [C.POINTER(apc_options), C.POINTER(_vp)]),
    "apc_wait_async": (None, [_vp]),
    "apc_store_timeout_counters": (C.c_int, [_vp, _vp]),
    "apc_set_store_delays": (C.c_int, [_vp, _i32, _i32, _i32]),
    "apc_put_async_rejections": (_i64, [_vp]),
    "apc_put_not_ready_errors": (_i64, [_vp]),
    "apc_factory_get": (C.c_int, [C.POINTER(apc_options), C.POINTER(_vp)]),
    "apc_factory_clear": (None, []),
    "apc_factory_create_errors": (_i64, []),
    "apc_file_open": (_vp, [_vp, _cp, _i64, EXTERNAL_READ, _vp]),
    "apc_stream_read": (_i32, [_vp, _vp, _i32, _i32, _i32]),
    "apc_stream_positioned_read": (_i32, [_vp, _i64, _vp, _i32, _i32, _i32]),
    "apc_stream_skip": (_i64, [_vp, _i64]),
    "apc_stream_seek": (C.c_int, [_vp, _i64]),
    "apc_stream_remaining": (_i64, [_vp]),
    "apc_stream_get_pos": (_i64, [_vp]),
    "apc_stream_bypass": (C.c_int, [_vp]),
    "apc_stream_counters": (None, [_vp, _vp]),
    "apc_stream_set_batched": (None, [_vp, C.c_int]),
    "apc_stream_close": (None, [_vp]),
    "apc_stream_destroy": (None, [_vp]),
    "apc_manager_create": (C.c_int, [C.POINTER(capi.aps_config), _cp, _dbl, C.POINTER(_vp)]),
    "apc_manager_create_sharded": (C.c_int, [C.POINTER(capi.aps_config), C.c_uint32, C.POINTER(_i32), _cp, _dbl, C.POINTER(_vp)]),
    "apc_shard_count": (C.c_uint32, [_vp]),
    "apc_shard_store": (_vp, [_vp, C.c_uint32]),
    "apc_shard_of": (C.c_uint32, [_vp, _cp, _i64]),
    "apc_manager_create_over": (C.c_int, [_vp, C.c_uint32, _cp, _dbl, C.POINTER(_vp)]),
    "apc_manager_close": (C.c_int, [_vp]),
    "apc_get_noexcept": (_i32, [_vp, _cp, _i64, _i32, _i32, _vp, _i32, _i32]),
    "apc_metrics_read": (None, [_vp, _vp]),
    "apc_metrics_read_more": (None, [_vp, _vp]),
    "apc_positioned_read_batched": (_i32, [_vp, _cp, _i64, _i64, _vp, _i32, EXTERNAL_READ, _vp, C.POINTER(_i64),
                                           C.POINTER(_i64)]),
    "apc_put": (C.c_int, [_vp, _cp, _i64, _vp, _i32]),
    "apc_get": (_i32, [_vp, _cp, _i64, _i32, _i32, _vp, _i32, _i32]),
    "apc_delete": (C.c_int, [_vp, _cp, _i64]),
    "apc_state": (C.c_int, [_vp]),
    "apc_manager_snapshot_save": (C.c_int, [_vp, _cp, C.POINTER(_u64)]),
    "apc_manager_create_over_snapshot": (C.c_int, [_vp, C.c_uint32, _cp, _dbl, _cp, C.POINTER(_vp)]),
    "apc_bytes": (_i64, [_vp]),
    "apc_pages": (_i64, [_vp]),
    "apc_page_lock_id": (_i32, [_vp, _cp, _i64]),
    "apc_set_store_faults": (C.c_int, [_vp, C.c_uint32]),
    "apc_store": (_vp, [_vp]),
    "apc_positioned_read": (_i32, [_vp, _cp, _i64, _i64, _vp, _i32, EXTERNAL_READ, _vp, C.POINTER(_i64), C.POINTER(_i64)]),
    "apc_replay": (_i64, [_vp, _i64, _vp, _vp, _vp, _u64, C.c_int, _vp, _vp, _vp, _vp, _i64, C.POINTER(_i64),
                          C.POINTER(_i64), C.POINTER(_i64)]),
    "apc_replay_batched": (_i64, [_vp, _i64, _vp, _vp, _vp, _u64, C.c_uint32, _vp, _vp, _vp, _vp, _i64, C.POINTER(_i64),
                                  C.POINTER(_i64), _vp, C.POINTER(_dbl), C.POINTER(_dbl)]),
    "apc_evictor_create": (_vp, [_cp, _dbl]),
    "apc_evictor_destroy": (None, [_vp]),
    "apc_evictor_update_on_get": (None, [_vp, _cp, _i64]),
    "apc_evictor_update_on_put": (None, [_vp, _cp, _i64]),
    "apc_evictor_update_on_delete": (None, [_vp, _cp, _i64]),
    "apc_evictor_reset": (None, [_vp]),
    "apc_evictor_evict": (C.c_int, [_vp, _vp, C.c_size_t, C.POINTER(_i64)]),
    "apc_meta_create": (_vp, [_cp, _dbl]),
    "apc_meta_destroy": (None, [_vp]),
    "apc_meta_has_page": (C.c_int, [_vp, _cp, _i64]),
    "apc_meta_add_page": (None, [_vp, _cp, _i64, _i64]),
    "apc_meta_get_page_info": (C.c_int, [_vp, _cp, _i64, C.POINTER(_i64)]),
    "apc_meta_remove_page": (C.c_int, [_vp, _cp, _i64]),
    "apc_meta_bytes": (_i64, [_vp]),
    "apc_meta_pages": (_i64, [_vp]),
    "apc_meta_reset": (None, [_vp]),
    "apc_meta_evict": (C.c_int, [_vp, _vp, C.c_size_t, C.POINTER(_i64), C.POINTER(_i64)]),
    "apc_file_key": (None, [_vp, _cp, C.POINTER(_u64), C.POINTER(_u64)]),
    "apc_java_string_hashcode": (_i32, [_cp]),
    "apc_hadoop_file_identifier": (None, [_cp, _vp]),
    "apc_file_key_of": (None, [_cp, C.POINTER(_u64), C.POINTER(_u64)]),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libaps_host.so is not built: run `python -m alluxio_b200.build`")
        capi.lib()  # libaps.so first (the host library links against it)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


class PageNotFoundException(Exception):
    pass


def _f(p):
    return p.file_id.encode("utf-8")


class CacheEvictor:
    """CACHE/CacheEvictor.java:26-69; kind in {"LRU", "LFU", "FIFO"}."""

    def __init__(self, kind, lfu_logbase=2.0):
        self._h = lib().apc_evictor_create(kind.encode(), lfu_logbase)
        if not self._h:
            raise ValueError(kind)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().apc_evictor_destroy(self._h)
            self._h = None

    def update_on_get(self, p):
        lib().apc_evictor_update_on_get(self._h, _f(p), p.page_index)

    def update_on_put(self, p):
        lib().apc_evictor_update_on_put(self._h, _f(p), p.page_index)

    def update_on_delete(self, p):
        lib().apc_evictor_update_on_delete(self._h, _f(p), p.page_index)

    def reset(self):
        lib().apc_evictor_reset(self._h)

    def evict(self):
        name = C.create_string_buffer(512)
        idx = C.c_int64()
        if not lib().apc_evictor_evict(self._h, name, 512, C.byref(idx)):
            return None
        return PageId(name.value.decode(), idx.value)


class MetaStore:
    """CACHE/DefaultMetaStore.java:33-132."""

    def __init__(self, evictor="LRU", lfu_logbase=2.0):
        self._h = lib().apc_meta_create(evictor.encode(), lfu_logbase)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().apc_meta_destroy(self._h)
            self._h = None

    def has_page(self, p):
        return bool(lib().apc_meta_has_page(self._h, _f(p), p.page_index))

    def add_page(self, p, size):
        lib().apc_meta_add_page(self._h, _f(p), p.page_index, size)

    def get_page_info(self, p):
        size = C.c_int64()
        if not lib().apc_meta_get_page_info(self._h, _f(p), p.page_index, C.byref(size)):
            raise PageNotFoundException(repr(p))
        return (p, size.value)

    def remove_page(self, p):
        if not lib().apc_meta_remove_page(self._h, _f(p), p.page_index):
            raise PageNotFoundException(repr(p))

    def bytes(self):
        return lib().apc_meta_bytes(self._h)

    def pages(self):
        return lib().apc_meta_pages(self._h)

    def reset(self):
        lib().apc_meta_reset(self._h)

    def evict(self):
        name = C.create_string_buffer(512)
        idx, size = C.c_int64(), C.c_int64()
        if not lib().apc_meta_evict(self._h, name, 512, C.byref(idx), C.byref(size)):
            return None
        return (PageId(name.value.decode(), idx.value), size.value)


class LocalCacheManager:
    """CACHE/LocalCacheManager.java:75-627 over a GpuPageStore (sync write, sync restore)."""
    NOT_IN_USE, READ_ONLY, READ_WRITE = 0, 1, 2

    def __init__(self, page_size, cache_size, evictor="LRU", lfu_logbase=2.0, device=0, slab_slots=0, crc=True,
                 max_batch=0, shard_devices=None):
        """shard_devices: CUDA ordinals, one per shard, for ONE cache spread over several GPUs: the
        manager, its meta store and its byte budget stay single and global, pages live on GPU hash(PageId) mod n."""
        cfg = capi.aps_config(device=device, page_size=page_size, cache_bytes=cache_size, slab_slots=slab_slots,
                              table_capacity=0, max_batch=max_batch, flags=capi.APS_FLAG_CRC if crc else 0)
        h = C.c_void_p()
        if shard_devices:
            devs = (C.c_int32 * len(shard_devices))(*shard_devices)
            capi.check(lib().apc_manager_create_sharded(C.byref(cfg), len(shard_devices), devs, evictor.encode(), lfu_logbase,
                                                        C.byref(h)), "apc_manager_create_sharded")
        else:
            capi.check(lib().apc_manager_create(C.byref(cfg), evictor.encode(), lfu_logbase, C.byref(h)), "apc_manager_create")
        self._h = h
        self.page_size = page_size

    @staticmethod
    def _options(page_size, cache_size, evictor="LRU", lfu_logbase=2.0, device=0, slab_slots=0, crc=True, max_batch=0,
                 shard_devices=None, over=None, async_write_threads=0, async_restore=False, timeout_ms=-1,
                 timeout_threads=0):
        o = apc_options()
        o.store = capi.aps_config(device=device, page_size=page_size, cache_bytes=cache_size, slab_slots=slab_slots,
                                  table_capacity=0, max_batch=max_batch, flags=capi.APS_FLAG_CRC if crc else 0)
        o.evictor, o.lfu_logbase = evictor.encode(), lfu_logbase
        keep = None
        if shard_devices:
            keep = (C.c_int32 * len(shard_devices))(*shard_devices)
            o.n_shards, o.shard_devices = len(shard_devices), keep
        if over is not None:
            o.over = over.handle
        o.async_write_threads, o.async_restore = async_write_threads, int(async_restore)
        o.timeout_ms, o.timeout_threads = timeout_ms, timeout_threads
        return o, keep

    @classmethod
    def create(cls, page_size, cache_size, **kw):
        """LocalCacheManager.create(conf): every configuration knob of the reference's factory method, including
        async_write_threads (alluxio.user.client.cache.async.write.*) and async_restore (...async.restore.enabled)."""
        self = cls.__new__(cls)
        o, keep = cls._options(page_size, cache_size, **kw)
        h = C.c_void_p()
        capi.check(lib().apc_manager_create_ex(C.byref(o), C.byref(h)), "apc_manager_create_ex")
        self._h, self.page_size, self._keep = h, page_size, kw.get("over")
        return self

    def store_timeouts(self):
        """Client.CacheStore{Put,Get,Delete}Timeout and ...ThreadsRejected of the TimeBoundPageStore wrapper."""
        import numpy as np
        v = np.zeros(4, np.int64)
        capi.check(lib().apc_store_timeout_counters(self._h, v.ctypes.data), "no timeout configured")
        return dict(zip(("put", "get", "delete", "rejected"), v.tolist()))

    def set_store_delays(self, put_ms=0, get_ms=0, delete_ms=0):
        capi.check(lib().apc_set_store_delays(self._h, put_ms, get_ms, delete_ms), "no timeout configured")

    def wait_async(self):
        lib().apc_wait_async(self._h)

    def put_async_rejections(self):
        return lib().apc_put_async_rejections(self._h)

    def put_not_ready_errors(self):
        return lib().apc_put_not_ready_errors(self._h)

    def open_file(self, file_id, file_len, external):
        """LocalCacheFileSystem.openFile(status): a LocalCacheFileInStream, or the external stream when the cache
        is NOT_IN_USE."""
        return LocalCacheFileInStream(self, file_id, file_len, external)

    def shard_count(self):
        return lib().apc_shard_count(self._h)

    def shard_of(self, p):
        return lib().apc_shard_of(self._h, _f(p), p.page_index)

    def shard_stats(self, i):
        st = capi.aps_stats_t()
        capi.check(capi.lib().aps_stats(C.c_void_p(lib().apc_shard_store(self._h, i)), C.byref(st)), "aps_stats")
        return {n: getattr(st, n) for n, _ in capi.aps_stats_t._fields_}

    @classmethod
    def over_store(cls, store, page_size, evictor="LRU", lfu_logbase=2.0):
        """LocalCacheManager.create(conf, metaStore, pageStore) over a store that already holds pages
        (a GpuPageStore from alluxio_b200.page_store): runs the reference's restore."""
        self = cls.__new__(cls)
        h = C.c_void_p()
        capi.check(lib().apc_manager_create_over(store.handle, page_size, evictor.encode(), lfu_logbase, C.byref(h)))
        self._h, self.page_size, self._keep = h, page_size, store
        return self

    @classmethod
    def over_snapshot(cls, store, page_size, path, evictor="LRU", lfu_logbase=2.0):
        """Warm restart: loads the snapshot at `path` into the freshly opened `store`, takes over the table of interned
        file ids it carries and runs the reference's restore over the loaded pages."""
        import os
        self = cls.__new__(cls)
        h = C.c_void_p()
        capi.check(lib().apc_manager_create_over_snapshot(store.handle, page_size, evictor.encode(), lfu_logbase,
                                                          os.fsencode(path), C.byref(h)), "restore from %s" % path)
        self._h, self.page_size, self._keep = h, page_size, store
        return self

    def save_snapshot(self, path):
        import os
        n = _u64()
        capi.check(lib().apc_manager_snapshot_save(self._h, os.fsencode(path), C.byref(n)), "snapshot save %s" % path)
        return n.value

    METRIC_NAMES = ("BytesReadCache", "BytesRequestedExternal", "BytesReadExternal", "BytesWrittenCache", "BytesEvicted",
                    "PagesEvicted", "GetErrors", "PutErrors", "DeleteErrors", "PutEvictionErrors", "PutStoreWriteErrors",
                    "PutStoreDeleteErrors", "GetStoreReadErrors", "DeleteNonExistingPageErrors")

    MORE_METRIC_NAMES = ("GetNotReadyErrors", "DeleteNotReadyErrors", "DeleteStoreDeleteErrors", "PutBenignRacingErrors",
                         "CleanupGetErrors", "CleanupPutErrors", "PutNotReadyErrors", "PutAsyncRejectionErrors")

    def metrics(self):
        """The Client.Cache* counters of LocalCacheManager.Metrics (CACHE/LocalCacheManager.java:557-627)."""
        import numpy as np
        v, w = np.zeros(14, np.int64), np.zeros(8, np.int64)
        lib().apc_metrics_read(self._h, v.ctypes.data)
        lib().apc_metrics_read_more(self._h, w.ctypes.data)
        return dict(zip(self.METRIC_NAMES + self.MORE_METRIC_NAMES, v.tolist() + w.tolist()))

    def get_noexcept(self, p, bytes_to_read, buffer, offset_in_buffer=0, page_offset=0):
        cbuf = (C.c_uint8 * max(1, len(buffer))).from_buffer(buffer) if len(buffer) else (C.c_uint8 * 1)()
        r = lib().apc_get_noexcept(self._h, _f(p), p.page_index, page_offset, bytes_to_read, C.addressof(cbuf), len(buffer),
                                   offset_in_buffer)
        del cbuf
        return r

    def close(self):
        if self._h is not None:
            if not getattr(self, "_factory_owned", False):  # the factory's manager goes away with Factory.clear()
                lib().apc_manager_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def put(self, p, page):
        data = bytes(page)
        buf = (C.c_uint8 * max(1, len(data))).from_buffer_copy(data or b"\0")
        return bool(lib().apc_put(self._h, _f(p), p.page_index, buf, len(data)))

    def get(self, p, bytes_to_read, buffer, offset_in_buffer=0, page_offset=0):
        cbuf = (C.c_uint8 * max(1, len(buffer))).from_buffer(buffer) if len(buffer) else (C.c_uint8 * 1)()
        r = lib().apc_get(self._h, _f(p), p.page_index, page_offset, bytes_to_read, C.addressof(cbuf), len(buffer),
                          offset_in_buffer)
        del cbuf
        if r == -2:
            raise ValueError("IllegalArgumentException")
        return r

    def delete(self, p):
        return bool(lib().apc_delete(self._h, _f(p), p.page_index))

    def state(self):
        return lib().apc_state(self._h)

    def bytes(self):
        return lib().apc_bytes(self._h)

    def pages(self):
        return lib().apc_pages(self._h)

    def page_lock_id(self, p):
        return lib().apc_page_lock_id(self._h, _f(p), p.page_index)

    def set_store_faulty(self, put=False, delete=False):
        mask = (capi.APS_FAULT_PUT if put else 0) | (capi.APS_FAULT_DELETE if delete else 0)
        capi.check(lib().apc_set_store_faults(self._h, mask))

    def store_stats(self):
        st = capi.aps_stats_t()
        capi.check(capi.lib().aps_stats(lib().apc_store(self._h), C.byref(st)))
        return {n: getattr(st, n) for n, _ in capi.aps_stats_t._fields_}

    def positioned_read(self, file_id, file_len, pos, length, external, batched=False):
        """LocalCacheFileInStream.positionedRead: `external(pos, n) -> bytes` stands for the
        external FileInStream.  Returns (data or None at EOF, bytes_from_cache, bytes_from_external)."""
        import numpy as np
        out = np.zeros(max(1, length), np.uint8)
        fc, fe = C.c_int64(0), C.c_int64(0)

        def ext(user, file_pos, dst, n):
            data = external(file_pos, n)
            C.memmove(dst, data, len(data))
            return len(data)

        cb = EXTERNAL_READ(ext)
        fn = lib().apc_positioned_read_batched if batched else lib().apc_positioned_read
        r = fn(self._h, str(file_id).encode(), file_len, pos, out.ctypes.data, length, cb, None, C.byref(fc), C.byref(fe))
        if r == -2:
            raise ValueError("IllegalArgumentException")
        return (None if r < 0 else out[:r].tobytes()), fc.value, fe.value

    def replay(self, files, pages, page_lens, seed, verify=True, victims_cap=None):
        """Reference read flow (get; on a miss put the synthetic page) over a trace."""
        import numpy as np
        files = np.ascontiguousarray(files, dtype=np.int64)
        pages = np.ascontiguousarray(pages, dtype=np.int64)
        page_lens = np.ascontiguousarray(page_lens, dtype=np.int32)
        n = files.size
        cap = n if victims_cap is None else victims_cap
        hit, put_ok = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        vf, vp = np.zeros(cap, np.int64), np.zeros(cap, np.int64)
        fp, fb, bad = C.c_int64(), C.c_int64(), C.c_int64()
        nv = lib().apc_replay(self._h, n, files.ctypes.data, pages.ctypes.data, page_lens.ctypes.data, seed, int(verify),
                              hit.ctypes.data, put_ok.ctypes.data, vf.ctypes.data, vp.ctypes.data, cap, C.byref(fp),
                              C.byref(fb), C.byref(bad))
        k = min(nv, cap)
        return dict(hit=hit, put_ok=put_ok, victims=np.stack([vf[:k], vp[:k]], axis=1), n_victims=nv,
                    final_pages=fp.value, final_bytes=fb.value, verify_failures=bad.value)


def _replay_batched(self, files, pages, page_lens, seed, window=4096, victims_cap=None):
    """apc_replay_batched: same decisions as replay(), store operations issued in batches."""
    import numpy as np
    files = np.ascontiguousarray(files, dtype=np.int64)
    pages = np.ascontiguousarray(pages, dtype=np.int64)
    page_lens = np.ascontiguousarray(page_lens, dtype=np.int32)
    n = files.size
    cap = n if victims_cap is None else victims_cap
    hit, put_ok = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
    vf, vp = np.zeros(max(cap, 1), np.int64), np.zeros(max(cap, 1), np.int64)
    fp, fb = C.c_int64(), C.c_int64()
    ops = np.zeros(5, np.int64)
    th, ts = C.c_double(), C.c_double()
    nv = lib().apc_replay_batched(self._h, n, files.ctypes.data, pages.ctypes.data, page_lens.ctypes.data, seed, window,
                                  hit.ctypes.data, put_ok.ctypes.data, vf.ctypes.data, vp.ctypes.data, cap, C.byref(fp),
                                  C.byref(fb), ops.ctypes.data, C.byref(th), C.byref(ts))
    if nv < 0:
        raise RuntimeError("apc_replay_batched failed")
    k = min(nv, cap)
    return dict(hit=hit, put_ok=put_ok, victims=np.stack([vf[:k], vp[:k]], axis=1), n_victims=nv,
                final_pages=fp.value, final_bytes=fb.value, deletes=int(ops[0]), puts=int(ops[1]), gets=int(ops[2]),
                flushes=int(ops[3]), bad_gets=int(ops[4]), seconds_host=th.value, seconds_store=ts.value)


LocalCacheManager.replay_batched = _replay_batched


class CacheManagerFactory:
    """CacheManager.Factory (CACHE/CacheManager.java:69-127): the process-wide manager."""

    @staticmethod
    def get(page_size, cache_size, **kw):
        o, keep = LocalCacheManager._options(page_size, cache_size, **kw)
        h = C.c_void_p()
        capi.check(lib().apc_factory_get(C.byref(o), C.byref(h)), "Failed to create CacheManager")
        if not h.value:
            return None  # another thread is still creating it
        m = LocalCacheManager.__new__(LocalCacheManager)
        m._h, m.page_size, m._factory_owned = h, page_size, True
        return m

    @staticmethod
    def clear():
        lib().apc_factory_clear()

    @staticmethod
    def create_errors():
        return lib().apc_factory_create_errors()


class LocalCacheFileInStream:
    """CACHE/LocalCacheFileInStream.java:43-330 (+ the bypass of LocalCacheFileSystem.openFile, :51-70).
    `external(pos, n) -> bytes` stands for the delegated file system's stream."""

    def __init__(self, manager, file_id, file_len, external):
        def ext(user, file_pos, dst, n):
            data = external(file_pos, n)
            C.memmove(dst, data, len(data))
            return len(data)

        self._cb = EXTERNAL_READ(ext)
        self._h = C.c_void_p(lib().apc_file_open(manager._h if manager is not None else None, str(file_id).encode(),
                                                 file_len, self._cb, None))
        if not self._h.value:
            raise ValueError("apc_file_open")

    @staticmethod
    def _result(r, buf, off):
        if r == -2:
            raise ValueError("IllegalArgumentException")
        if r == -3:
            raise IOError("Failed to read complete page from external storage")
        return r

    def read(self, b, off=0, length=None):
        """int read(byte[] b, int off, int len): fills the bytearray in place; -1 at end of file."""
        length = len(b) - off if length is None else length
        cbuf = (C.c_uint8 * max(1, len(b))).from_buffer(b) if len(b) else (C.c_uint8 * 1)()
        r = lib().apc_stream_read(self._h, C.addressof(cbuf), len(b), off, length)
        del cbuf
        return self._result(r, b, off)

    def positioned_read(self, pos, b, off=0, length=None):
        length = len(b) - off if length is None else length
        cbuf = (C.c_uint8 * max(1, len(b))).from_buffer(b) if len(b) else (C.c_uint8 * 1)()
        r = lib().apc_stream_positioned_read(self._h, pos, C.addressof(cbuf), len(b), off, length)
        del cbuf
        return self._result(r, b, off)

    def set_batched(self, on=True):
        """one batched store call per read instead of one get per page."""
        lib().apc_stream_set_batched(self._h, int(on))

    def skip(self, n):
        r = lib().apc_stream_skip(self._h, n)
        if r == -4:
            raise RuntimeError("Cannot operate on a closed stream")
        return r

    def seek(self, pos):
        r = lib().apc_stream_seek(self._h, pos)
        if r == -2:
            raise ValueError("IllegalArgumentException")
        if r == -4:
            raise RuntimeError("Cannot operate on a closed stream")

    def remaining(self):
        return lib().apc_stream_remaining(self._h)

    def get_pos(self):
        return lib().apc_stream_get_pos(self._h)

    def bypass(self):
        return bool(lib().apc_stream_bypass(self._h))

    def counters(self):
        import numpy as np
        v = np.zeros(3, np.int64)
        lib().apc_stream_counters(self._h, v.ctypes.data)
        return dict(zip(("BytesReadCache", "BytesRequestedExternal", "BytesReadExternal"), v.tolist()))

    def close(self):
        lib().apc_stream_close(self._h)

    def __del__(self):
        try:
            if self._h and self._h.value:
                lib().apc_stream_destroy(self._h)
                self._h = None
        except Exception:
            pass


def hadoop_file_identifier(path):
    """HadoopUtils.toAlluxioUriStatus: md5 hex of the path string."""
    out = C.create_string_buffer(33)
    lib().apc_hadoop_file_identifier(path.encode("utf-8"), out)
    return out.value.decode()


def file_key_of(file_id):
    hi, lo = C.c_uint64(), C.c_uint64()
    lib().apc_file_key_of(str(file_id).encode(), C.byref(hi), C.byref(lo))
    return hi.value, lo.value
