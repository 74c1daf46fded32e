"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the CPU restatement of the reference).  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this module; nothing under
alluxio_b200/ does.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


_SOURCES = ("checksums.c", "page_cache.cc", "page_cache.h", "capi.cc", "Makefile")


def _stale():
    if not os.path.exists(_LIB_PATH):
        return True
    built = os.path.getmtime(_LIB_PATH)
    return any(os.path.getmtime(os.path.join(_HERE, f)) > built for f in _SOURCES)


def build(force=False):
    """Compile the restatement with the committed Makefile when the library is missing or older than a source.  Several
    processes may get here at once (bench.py under torchrun: every rank checks its pages against the oracle), so the
    build runs under a file lock: one builds, the others wait and then find it up to date."""
    if not force and not _stale():
        return _LIB_PATH
    import fcntl
    with open(os.path.join(_HERE, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if force or _stale():
            subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
            os.utime(_LIB_PATH)  # make may have had nothing to do (objects newer than sources): mark the library current
    return _LIB_PATH


_lib = None
EXTERNAL_READ = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, cp, i32, i64, u64, u8p, dbl = C.c_void_p, C.c_char_p, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p, C.c_double
        sig = {
            "orc_crc32c_bitwise": (C.c_uint32, [vp, C.c_size_t]),
            "orc_crc32c": (C.c_uint32, [vp, C.c_size_t]),
            "orc_crc32c_hw": (C.c_uint32, [vp, C.c_size_t, C.POINTER(C.c_int)]),
            "orc_crc32c_combine": (C.c_uint32, [C.c_uint32, C.c_uint32, u64]),
            "orc_md5": (None, [vp, C.c_size_t, vp]),
            "orc_md5_hex": (None, [vp, C.c_size_t, vp]),
            "orc_java_string_hashcode": (i32, [cp]),
            "orc_floor_mod": (i32, [i32, i32]),
            "orc_evictor_create": (vp, [cp, dbl]),
            "orc_evictor_destroy": (None, [vp]),
            "orc_evictor_update_on_get": (None, [vp, cp, i64]),
            "orc_evictor_update_on_put": (None, [vp, cp, i64]),
            "orc_evictor_update_on_delete": (None, [vp, cp, i64]),
            "orc_evictor_reset": (None, [vp]),
            "orc_evictor_evict": (C.c_int, [vp, vp, C.c_size_t, C.POINTER(i64)]),
            "orc_meta_create": (vp, [cp, dbl]),
            "orc_meta_destroy": (None, [vp]),
            "orc_meta_has_page": (C.c_int, [vp, cp, i64]),
            "orc_meta_add_page": (None, [vp, cp, i64, i64]),
            "orc_meta_get_page_info": (C.c_int, [vp, cp, i64, C.POINTER(i64)]),
            "orc_meta_remove_page": (C.c_int, [vp, cp, i64]),
            "orc_meta_bytes": (i64, [vp]),
            "orc_meta_pages": (i64, [vp]),
            "orc_meta_reset": (None, [vp]),
            "orc_meta_evict": (C.c_int, [vp, vp, C.c_size_t, C.POINTER(i64), C.POINTER(i64)]),
            "orc_store_open_local": (vp, [cp, i64, i64, i32, C.c_int]),
            "orc_store_open_mem": (vp, [i64, i64, C.c_int]),
            "orc_store_open_external": (vp, [vp, vp, vp, vp, vp, i64]),
            "orc_store_destroy": (None, [vp]),
            "orc_store_put": (C.c_int, [vp, cp, i64, u8p, i32]),
            "orc_store_get": (C.c_int, [vp, cp, i64, i32, i32, u8p, i32, i32, C.POINTER(i32)]),
            "orc_store_delete": (C.c_int, [vp, cp, i64]),
            "orc_store_set_faulty": (None, [vp, C.c_int, C.c_int]),
            "orc_store_get_pages": (C.c_int, [vp, vp, vp]),
            "orc_manager_create": (vp, [vp, i64, cp, dbl]),
            "orc_manager_destroy": (None, [vp]),
            "orc_manager_put": (C.c_int, [vp, cp, i64, u8p, i32]),
            "orc_manager_get": (i32, [vp, cp, i64, i32, i32, u8p, i32, i32]),
            "orc_manager_delete": (C.c_int, [vp, cp, i64]),
            "orc_manager_state": (C.c_int, [vp]),
            "orc_manager_page_lock_id": (i32, [vp, cp, i64]),
            "orc_manager_metrics": (None, [vp, vp]),
            "orc_manager_bytes": (i64, [vp]),
            "orc_manager_pages": (i64, [vp]),
            "orc_stream_open": (vp, [vp, cp, i64, i64, EXTERNAL_READ, vp]),
            "orc_stream_destroy": (None, [vp]),
            "orc_stream_read": (i32, [vp, vp, i32, i32, i32]),
            "orc_stream_positioned_read": (i32, [vp, i64, vp, i32, i32, i32]),
            "orc_stream_skip": (i64, [vp, i64]),
            "orc_stream_seek": (C.c_int, [vp, i64]),
            "orc_stream_remaining": (i64, [vp]),
            "orc_stream_get_pos": (i64, [vp]),
            "orc_stream_close": (None, [vp]),
            "orc_stream_counters": (None, [vp, vp]),
            "orc_replay": (i64, [cp, dbl, i64, i64, i64, vp, vp, vp, vp, vp, vp, vp, i64, C.POINTER(i64), C.POINTER(i64)]),
            "orc_fill_page": (None, [u64, u64, u64, vp, i64]),
            "orc_bench_local_get": (dbl, [cp, i64, i64, i64, C.c_int, i64, u64, u64, i32, C.c_int, C.POINTER(i64)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


# ----------------------------------------------------------------------------- small helpers

def _buf(b):
    """bytes / bytearray / numpy array -> (ctypes pointer value, length, keepalive)."""
    import numpy as np
    if isinstance(b, np.ndarray):
        a = np.ascontiguousarray(b).view(np.uint8).reshape(-1)
        return a.ctypes.data, a.size, a
    if isinstance(b, (bytes, bytearray)):
        a = (C.c_uint8 * len(b)).from_buffer_copy(bytes(b)) if len(b) else (C.c_uint8 * 1)()
        return C.addressof(a), len(b), a
    raise TypeError(type(b))


def crc32c(data, impl="sb8"):
    p, n, keep = _buf(data)
    if impl == "bitwise":
        return lib().orc_crc32c_bitwise(p, n)
    if impl == "hw":
        ok = C.c_int(0)
        v = lib().orc_crc32c_hw(p, n, C.byref(ok))
        return v if ok.value else None
    return lib().orc_crc32c(p, n)


def crc32c_combine(crc1, crc2, len2):
    return lib().orc_crc32c_combine(crc1, crc2, len2)


def md5(data):
    p, n, keep = _buf(data)
    out = (C.c_uint8 * 16)()
    lib().orc_md5(p, n, out)
    return bytes(out)


def md5_hex(data):
    p, n, keep = _buf(data)
    out = C.create_string_buffer(33)
    lib().orc_md5_hex(p, n, out)
    return out.value.decode()


def fill_page(seed, file, index, length):
    import numpy as np
    out = np.empty(length, dtype=np.uint8)
    lib().orc_fill_page(seed, file, index, out.ctypes.data, length)
    return out


def java_string_hashcode(s):
    return lib().orc_java_string_hashcode(s.encode("utf-8"))


def floor_mod(x, y):
    return lib().orc_floor_mod(x, y)


# ----------------------------------------------------------------------------- mirrors of the Java classes

class PageId:
    """CACHE/PageId.java:23-74."""
    __slots__ = ("file_id", "page_index")

    def __init__(self, file_id, page_index):
        self.file_id, self.page_index = str(file_id), int(page_index)

    def __eq__(self, o):
        return isinstance(o, PageId) and (self.file_id, self.page_index) == (o.file_id, o.page_index)

    def __hash__(self):
        return hash((self.file_id, self.page_index))

    def __repr__(self):
        return "PageId{FileId=%s, PageIndex=%d}" % (self.file_id, self.page_index)

    @property
    def _f(self):
        return self.file_id.encode("utf-8")


class PageNotFoundException(Exception):
    pass


class IOExceptionJ(Exception):
    pass


class Evictor:
    """CACHE/CacheEvictor.java:26-69 (kind: "LRU", "LFU", "FIFO")."""

    def __init__(self, kind, lfu_logbase=2.0):
        self._h = lib().orc_evictor_create(kind.encode(), lfu_logbase)
        assert self._h

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_evictor_destroy(self._h)
            self._h = None

    def update_on_get(self, p):
        lib().orc_evictor_update_on_get(self._h, p._f, p.page_index)

    def update_on_put(self, p):
        lib().orc_evictor_update_on_put(self._h, p._f, p.page_index)

    def update_on_delete(self, p):
        lib().orc_evictor_update_on_delete(self._h, p._f, p.page_index)

    def reset(self):
        lib().orc_evictor_reset(self._h)

    def evict(self):
        name = C.create_string_buffer(512)
        idx = C.c_int64()
        if not lib().orc_evictor_evict(self._h, name, 512, C.byref(idx)):
            return None
        return PageId(name.value.decode(), idx.value)


class MetaStore:
    """CACHE/DefaultMetaStore.java:33-132."""

    def __init__(self, evictor_kind="LRU", lfu_logbase=2.0):
        self._h = lib().orc_meta_create(evictor_kind.encode(), lfu_logbase)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_meta_destroy(self._h)
            self._h = None

    def has_page(self, p):
        return bool(lib().orc_meta_has_page(self._h, p._f, p.page_index))

    def add_page(self, p, size):
        lib().orc_meta_add_page(self._h, p._f, p.page_index, size)

    def get_page_info(self, p):
        size = C.c_int64()
        if not lib().orc_meta_get_page_info(self._h, p._f, p.page_index, C.byref(size)):
            raise PageNotFoundException(repr(p))
        return (p, size.value)

    def remove_page(self, p):
        if not lib().orc_meta_remove_page(self._h, p._f, p.page_index):
            raise PageNotFoundException(repr(p))

    def bytes(self):
        return lib().orc_meta_bytes(self._h)

    def pages(self):
        return lib().orc_meta_pages(self._h)

    def reset(self):
        lib().orc_meta_reset(self._h)

    def evict(self):
        name = C.create_string_buffer(512)
        idx, size = C.c_int64(), C.c_int64()
        if not lib().orc_meta_evict(self._h, name, 512, C.byref(idx), C.byref(size)):
            return None
        return (PageId(name.value.decode(), idx.value), size.value)


_PAGE_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_char_p, C.c_int64, C.c_int64)


class PageStore:
    """CACHE/PageStore.java:37-181 over LocalPageStore (a directory) or the in-RAM variant."""

    def __init__(self, handle, page_size, cache_size):
        if not handle:
            raise IOExceptionJ("cannot open page store")
        self._h, self.page_size, self.cache_size = handle, page_size, cache_size

    @classmethod
    def create_local(cls, root, page_size=1 << 20, cache_size=512 << 20, file_buckets=1000):
        """PageStore.create(options): wipe the directory, then open (PageStore.java:49-52)."""
        return cls(lib().orc_store_open_local(str(root).encode(), page_size, cache_size, file_buckets, 1), page_size, cache_size)

    @classmethod
    def open_local(cls, root, page_size=1 << 20, cache_size=512 << 20, file_buckets=1000):
        return cls(lib().orc_store_open_local(str(root).encode(), page_size, cache_size, file_buckets, 0), page_size, cache_size)

    @classmethod
    def open_mem(cls, page_size, cache_size, keep_bytes=True):
        return cls(lib().orc_store_open_mem(page_size, cache_size, int(keep_bytes)), page_size, cache_size)

    @classmethod
    def over(cls, impl, page_size, cache_size, make_id=None):
        """A PageStore implemented by `impl` -- any object with the interface of CACHE/PageStore.java:119-171 as the
        Python mirrors spell it: put(id, bytes), get(id, bytearray, page_offset, bytes_to_read, buffer_offset) -> n,
        delete(id), get_pages() -> [(id, size)].  The oracle's LocalCacheManager then drives `impl` exactly as the
        reference's LocalCacheManager drives a PageStore in the JVM.  Exceptions map as in the reference:
        *PageNotFoundException -> 1, ValueError (IllegalArgumentException) -> 2, anything else (IOException) -> 3."""
        make_id = make_id or PageId

        def status(e):
            if "PageNotFound" in type(e).__name__:
                return 1
            return 2 if isinstance(e, ValueError) else 3

        def put(user, f, idx, page, n):
            try:
                impl.put(make_id(f.decode(), idx), C.string_at(page, n) if n else b"")
                return 0
            except Exception as e:  # noqa: BLE001
                return status(e)

        def get(user, f, idx, page_offset, bytes_to_read, buffer, buf_len, buffer_offset, bytes_read):
            try:
                tmp = bytearray(max(buf_len, 0))
                n = impl.get(make_id(f.decode(), idx), tmp, page_offset, bytes_to_read, buffer_offset)
                if n > 0 and buffer:
                    C.memmove(buffer + buffer_offset, bytes(tmp[buffer_offset:buffer_offset + n]), n)
                bytes_read[0] = n
                return 0
            except Exception as e:  # noqa: BLE001
                return status(e)

        def delete(user, f, idx):
            try:
                impl.delete(make_id(f.decode(), idx))
                return 0
            except Exception as e:  # noqa: BLE001
                return status(e)

        def get_pages(user, emit, emit_user):
            try:
                for pid, size in impl.get_pages():
                    emit(emit_user, 1, str(pid.file_id).encode(), pid.page_index, size)
                return 0
            except Exception as e:  # noqa: BLE001
                return status(e)

        emit_t = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_char_p, C.c_int64, C.c_int64)
        cbs = (C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_int64, C.c_void_p, C.c_int32)(put),
               C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                           C.POINTER(C.c_int32))(get),
               C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_int64)(delete),
               C.CFUNCTYPE(C.c_int, C.c_void_p, emit_t, C.c_void_p)(get_pages))
        self = cls(lib().orc_store_open_external(None, *[C.cast(c, C.c_void_p) for c in cbs], cache_size), page_size, cache_size)
        self._keep = (impl, cbs)
        return self

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_store_destroy(self._h)
            self._h = None

    @staticmethod
    def _raise(st, what):
        if st == 1:
            raise PageNotFoundException(what)
        if st == 2:
            raise ValueError(what)  # IllegalArgumentException
        if st != 0:
            raise IOExceptionJ(what)

    def put(self, p, page):
        ptr, n, keep = _buf(page)
        self._raise(lib().orc_store_put(self._h, p._f, p.page_index, ptr, n), "put %r" % p)

    def get(self, p, buffer, page_offset=0, bytes_to_read=None, buffer_offset=0):
        """buffer: bytearray, filled in place.  Returns bytes read."""
        if bytes_to_read is None:
            bytes_to_read = len(buffer)
        cbuf = (C.c_uint8 * max(1, len(buffer))).from_buffer(buffer) if len(buffer) else (C.c_uint8 * 1)()
        got = C.c_int32()
        st = lib().orc_store_get(self._h, p._f, p.page_index, page_offset, bytes_to_read, C.addressof(cbuf),
                                 len(buffer), buffer_offset, C.byref(got))
        del cbuf
        self._raise(st, "get %r" % p)
        return got.value

    def delete(self, p):
        self._raise(lib().orc_store_delete(self._h, p._f, p.page_index), "delete %r" % p)

    def set_faulty(self, put=False, delete=False):
        lib().orc_store_set_faulty(self._h, int(put), int(delete))

    def get_pages(self):
        """List of (PageId, size), or None in place of an unrecognised file (Java null)."""
        out = []

        def cb(user, valid, f, idx, size):
            out.append((PageId(f.decode(), idx), size) if valid else None)

        c = _PAGE_CB(cb)
        self._raise(lib().orc_store_get_pages(self._h, C.cast(c, C.c_void_p), None), "getPages")
        return out

    def get_cache_size(self):
        return self.cache_size


class LocalCacheManager:
    """CACHE/LocalCacheManager.java:75-627 with sync write and sync restore."""
    NOT_IN_USE, READ_ONLY, READ_WRITE = 0, 1, 2

    def __init__(self, store, page_size, evictor="LRU", lfu_logbase=2.0):
        self._store = store
        self._h = lib().orc_manager_create(store._h, page_size, evictor.encode(), lfu_logbase)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_manager_destroy(self._h)
            self._h = None

    def put(self, p, page):
        ptr, n, keep = _buf(page)
        return bool(lib().orc_manager_put(self._h, p._f, p.page_index, ptr, n))

    def get(self, p, bytes_to_read, buffer, offset_in_buffer=0, page_offset=0):
        cbuf = (C.c_uint8 * max(1, len(buffer))).from_buffer(buffer) if len(buffer) else (C.c_uint8 * 1)()
        r = lib().orc_manager_get(self._h, p._f, p.page_index, page_offset, bytes_to_read, C.addressof(cbuf),
                                  len(buffer), offset_in_buffer)
        del cbuf
        if r == -2:
            raise ValueError("IllegalArgumentException")
        return r

    def delete(self, p):
        return bool(lib().orc_manager_delete(self._h, p._f, p.page_index))

    def state(self):
        return lib().orc_manager_state(self._h)

    def page_lock_id(self, p):
        return lib().orc_manager_page_lock_id(self._h, p._f, p.page_index)

    def bytes(self):
        return lib().orc_manager_bytes(self._h)

    METRICS = ("BytesWrittenCache", "BytesEvicted", "PagesEvicted", "GetErrors", "GetNotReadyErrors", "GetStoreReadErrors",
               "CleanupGetErrors", "PutErrors", "PutNotReadyErrors", "PutEvictionErrors", "PutStoreWriteErrors",
               "PutStoreDeleteErrors", "PutBenignRacingErrors", "CleanupPutErrors", "DeleteErrors", "DeleteNotReadyErrors",
               "DeleteNonExistingPageErrors", "DeleteStoreDeleteErrors")

    def metrics(self):
        """The Client.Cache* counters of LocalCacheManager.Metrics (LocalCacheManager.java:557-627)."""
        v = (C.c_int64 * len(self.METRICS))()
        lib().orc_manager_metrics(self._h, v)
        return dict(zip(self.METRICS, list(v)))

    def pages(self):
        return lib().orc_manager_pages(self._h)


class LocalCacheFileInStream:
    """CACHE/LocalCacheFileInStream.java:37-336 over an oracle LocalCacheManager.  `external(pos, n) -> bytes` stands
    for the external FileInStream (it may return fewer than n bytes; b"" ends the page read).  Exceptions as in the
    Python mirror of the product: ValueError = IllegalArgumentException, IOError = IOException, RuntimeError =
    IllegalStateException (closed stream)."""

    def __init__(self, manager, file_id, file_len, external, page_size):
        def ext(user, file_pos, dst, n):
            data = external(file_pos, n)
            C.memmove(dst, data, len(data))
            return len(data)

        self._manager = manager
        self._cb = EXTERNAL_READ(ext)
        self._h = lib().orc_stream_open(manager._h, str(file_id).encode(), file_len, page_size, self._cb, None)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_stream_destroy(self._h)
            self._h = None

    @staticmethod
    def _result(r):
        if r == -2:
            raise ValueError("IllegalArgumentException")
        if r == -3:
            raise IOError("IOException")
        return r

    def read(self, b, off=0, length=None):
        length = len(b) - off if length is None else length
        cbuf = (C.c_uint8 * max(1, len(b))).from_buffer(b) if len(b) else (C.c_uint8 * 1)()
        r = lib().orc_stream_read(self._h, C.addressof(cbuf), len(b), off, length)
        del cbuf
        return self._result(r)

    def read_byte(self):
        """int read(): CACHE/LocalCacheFileInStream.java:96-103."""
        one = bytearray(1)
        return -1 if self.read(one) == -1 else one[0]

    def positioned_read(self, pos, b, off=0, length=None):
        length = len(b) - off if length is None else length
        cbuf = (C.c_uint8 * max(1, len(b))).from_buffer(b) if len(b) else (C.c_uint8 * 1)()
        r = lib().orc_stream_positioned_read(self._h, pos, C.addressof(cbuf), len(b), off, length)
        del cbuf
        return self._result(r)

    def skip(self, n):
        r = lib().orc_stream_skip(self._h, n)
        if r == -4:
            raise RuntimeError("Cannot operate on a closed stream")
        return r

    def seek(self, pos):
        r = lib().orc_stream_seek(self._h, pos)
        if r == -2:
            raise ValueError("IllegalArgumentException")
        if r == -4:
            raise RuntimeError("Cannot operate on a closed stream")

    def remaining(self):
        return lib().orc_stream_remaining(self._h)

    def get_pos(self):
        return lib().orc_stream_get_pos(self._h)

    def close(self):
        lib().orc_stream_close(self._h)

    def counters(self):
        v = (C.c_int64 * 4)()
        lib().orc_stream_counters(self._h, v)
        return dict(zip(("BytesReadCache", "BytesRequestedExternal", "BytesReadExternal", "ExternalReadCalls"), list(v)))


def replay(evictor, page_size, cache_size, files, pages, page_lens, lfu_logbase=2.0, victims_cap=None):
    """Reference read flow (get; on miss put) over an access trace.  numpy int64/int32 inputs."""
    import numpy as np
    files = np.ascontiguousarray(files, dtype=np.int64)
    pages = np.ascontiguousarray(pages, dtype=np.int64)
    page_lens = np.ascontiguousarray(page_lens, dtype=np.int32)
    n = files.size
    cap = n if victims_cap is None else victims_cap
    hit = np.zeros(n, dtype=np.uint8)
    put_ok = np.zeros(n, dtype=np.uint8)
    vf = np.zeros(cap, dtype=np.int64)
    vp = np.zeros(cap, dtype=np.int64)
    fp, fb = C.c_int64(), C.c_int64()
    nv = lib().orc_replay(evictor.encode(), lfu_logbase, page_size, cache_size, n, files.ctypes.data, pages.ctypes.data,
                          page_lens.ctypes.data, hit.ctypes.data, put_ok.ctypes.data, vf.ctypes.data, vp.ctypes.data, cap,
                          C.byref(fp), C.byref(fb))
    k = min(nv, cap)
    return dict(hit=hit, put_ok=put_ok, victims=np.stack([vf[:k], vp[:k]], axis=1), n_victims=nv,
                final_pages=fp.value, final_bytes=fb.value)


def bench_local_get(root, page_size, n_files, pages_per_file, populate, n_gets, data_seed, access_seed, n_threads,
                    check=True):
    b = C.c_int64()
    secs = lib().orc_bench_local_get(str(root).encode(), page_size, n_files, pages_per_file, int(populate), n_gets,
                                     data_seed, access_seed, n_threads, int(check), C.byref(b))
    if secs < 0:
        raise RuntimeError("orc_bench_local_get failed: %r" % secs)
    return secs, b.value
