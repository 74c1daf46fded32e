"""Python mirror of the reference's PageStore interface over the libaps C ABI.

`GpuPageStore` has the same methods, argument meaning and error behaviour as
alluxio.client.file.cache.PageStore / LocalPageStore (CACHE/PageStore.java:37-181,
CACHE/store/LocalPageStore.java:44-212), so the parity tests read like the reference's
PageStoreTest.  It is a thin veneer: every call goes through include/aps.h via ctypes.
(CACHE/ = /root/reference/core/client/fs/src/main/java/alluxio/client/file/cache/.)
"""
import ctypes as C
import os
import re
import threading

from . import capi
from .capi import aps_config, aps_get_req, aps_key, aps_put_req, aps_result, aps_stats_t


class PageNotFoundException(Exception):
    """alluxio.exception.PageNotFoundException"""


class IOExceptionJ(Exception):
    """java.io.IOException"""


class PageId:
    """CACHE/PageId.java:23-74: (String fileId, long pageIndex)."""
    __slots__ = ("file_id", "page_index")

    def __init__(self, file_id, page_index):
        self.file_id, self.page_index = str(file_id), int(page_index)

    def __eq__(self, o):
        return isinstance(o, PageId) and (self.file_id, self.page_index) == (o.file_id, o.page_index)

    def __hash__(self):
        return hash((self.file_id, self.page_index))

    def __repr__(self):
        return "PageId{FileId=%s, PageIndex=%d}" % (self.file_id, self.page_index)


_DEC = re.compile(r"(0|-?[1-9][0-9]{0,18})\Z")   # \Z, not $: "123\n" is not a decimal long (Java's String.matches)
_HEX32 = re.compile(r"[0-9a-f]{32}\Z")
_MASK64 = (1 << 64) - 1
_INTERNED_HI = _MASK64


class FileKeyCodec:
    """String fileId -> 128 bits without collisions.

    * canonical decimal longs (Long.toString(fileId), wire/FileInfo.java:99-100)  -> (0, value)
    * 32 lower-case hex chars (MD5 of the path, hadoop/HadoopUtils.java:230) whose high word is not
      one of the two reserved values                                               -> (hi, lo)
    * everything else (UUIDs, raw paths, ...) is interned                          -> (2^64-1, n)
    The three ranges are disjoint, so equal keys imply equal strings.
    """

    def __init__(self):
        self._lock = threading.Lock()
        self._ids = {}
        self._names = []

    def encode(self, file_id):
        if _DEC.match(file_id):
            v = int(file_id)
            if -(1 << 63) <= v < (1 << 63):
                return 0, v & _MASK64
        if _HEX32.match(file_id):
            hi, lo = int(file_id[:16], 16), int(file_id[16:], 16)
            if hi not in (0, _INTERNED_HI):
                return hi, lo
        with self._lock:
            n = self._ids.get(file_id)
            if n is None:
                n = len(self._names)
                self._ids[file_id] = n
                self._names.append(file_id)
            return _INTERNED_HI, n

    def names(self):
        """The interned ids in index order (what a snapshot has to carry, see save_snapshot)."""
        with self._lock:
            return list(self._names)

    def adopt_names(self, names):
        """After a restart: take over the intern table a snapshot carries.  The table here must be empty or agree."""
        with self._lock:
            for i, nm in enumerate(names):
                if i < len(self._names):
                    if self._names[i] != nm:
                        raise ValueError("snapshot name table disagrees with ids already interned: %r vs %r" % (self._names[i], nm))
                else:
                    self._ids[nm] = i
                    self._names.append(nm)

    def decode(self, hi, lo):
        if hi == 0:
            return str(lo - (1 << 64) if lo >> 63 else lo)
        if hi == _INTERNED_HI:
            with self._lock:
                return self._names[lo]
        return "%016x%016x" % (hi, lo)


class GpuPageStoreOptions:
    """Counterpart of LocalPageStoreOptions (CACHE/store/LocalPageStoreOptions.java:19-67)."""

    def __init__(self, page_size=1 << 20, cache_size=512 << 20, device=0, slab_slots=0, table_capacity=0,
                 max_batch=0, crc=True):
        self.page_size, self.cache_size, self.device = page_size, cache_size, device
        self.slab_slots, self.table_capacity, self.max_batch, self.crc = slab_slots, table_capacity, max_batch, crc


class GpuPageStore:
    """PageStore backed by an HBM slab + hash table on one B200 (the new PageStoreType.GPU)."""

    def __init__(self, options=None, **kw):
        o = options or GpuPageStoreOptions(**kw)
        self.options = o
        self.codec = FileKeyCodec()
        cfg = aps_config(device=o.device, page_size=o.page_size, cache_bytes=o.cache_size, slab_slots=o.slab_slots,
                         table_capacity=o.table_capacity, max_batch=o.max_batch,
                         flags=capi.APS_FLAG_CRC if o.crc else 0)
        h = C.c_void_p()
        rc = capi.lib().aps_open(C.byref(cfg), C.byref(h))
        if rc != capi.APS_OK:
            raise IOExceptionJ(str(capi.ApsError(rc, "aps_open")))
        self._h = h
        self.last_crc = None

    # ---- helpers
    def key(self, page_id):
        hi, lo = self.codec.encode(page_id.file_id)
        return aps_key(hi, lo, page_id.page_index & _MASK64)

    @property
    def handle(self):
        if self._h is None:
            raise IOExceptionJ("store is closed")
        return self._h

    @staticmethod
    def _raise(rc, what):
        if rc == capi.APS_ENOTFOUND:
            raise PageNotFoundException(what)
        if rc == capi.APS_EINVAL:
            raise ValueError(what)  # IllegalArgumentException
        if rc != capi.APS_OK:
            raise IOExceptionJ(str(capi.ApsError(rc, what)))

    # ---- PageStore interface
    def put(self, page_id, page):
        """void put(PageId, byte[]) throws IOException  (PageStore.java:119)"""
        data = bytes(page)
        buf = (C.c_uint8 * max(1, len(data))).from_buffer_copy(data or b"\0")
        k = self.key(page_id)
        self._raise(capi.lib().aps_put(self.handle, C.byref(k), buf, len(data)), "put %r" % page_id)

    def get(self, page_id, buffer, page_offset=0, bytes_to_read=None, buffer_offset=0):
        """int get(PageId, pageOffset, bytesToRead, byte[] buffer, bufferOffset)  (PageStore.java:147).
        `buffer` is a bytearray filled in place; returns the number of bytes read."""
        if bytes_to_read is None:
            bytes_to_read = len(buffer)
        # Preconditions of LocalPageStore.get, LocalPageStore.java:91-93 (done shim-side, as the JNI shim does)
        if page_offset < 0:
            raise ValueError("page offset should be non-negative")
        if len(buffer) < buffer_offset:
            raise ValueError("buffer offset %d exceeds buffer length %d" % (buffer_offset, len(buffer)))
        room = len(buffer) - buffer_offset
        want = max(0, min(bytes_to_read, room))
        tmp = (C.c_uint8 * max(1, want))()
        out_len, out_crc = C.c_uint32(), C.c_uint32()
        k = self.key(page_id)
        rc = capi.lib().aps_get(self.handle, C.byref(k), page_offset, want, tmp, want, C.byref(out_len), C.byref(out_crc))
        self._raise(rc, "get %r" % page_id)
        n = out_len.value
        buffer[buffer_offset:buffer_offset + n] = bytes(tmp[:n])
        self.last_crc = out_crc.value
        return n

    def delete(self, page_id):
        """void delete(PageId) throws IOException, PageNotFoundException  (PageStore.java:157)"""
        k = self.key(page_id)
        self._raise(capi.lib().aps_delete(self.handle, C.byref(k)), "delete %r" % page_id)

    def get_pages(self):
        """Stream<PageInfo> getPages()  (PageStore.java:166): list of (PageId, length)."""
        out = []

        def cb(user, key, length):
            k = key.contents
            idx = k.page_index - (1 << 64) if k.page_index >> 63 else k.page_index
            out.append((PageId(self.codec.decode(k.file_hi, k.file_lo), idx), length))
            return 0

        c = capi.ITER_CB(cb)
        self._raise(capi.lib().aps_iterate(self.handle, c, None), "getPages")
        return out

    def get_cache_size(self):
        """long getCacheSize()  (PageStore.java:171)"""
        return capi.lib().aps_cache_size(self.handle)

    def close(self):
        if self._h is not None:
            capi.lib().aps_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- additive surface
    def set_faulty(self, put=False, delete=False):
        mask = (capi.APS_FAULT_PUT if put else 0) | (capi.APS_FAULT_DELETE if delete else 0)
        capi.check(capi.lib().aps_set_faults(self.handle, mask))

    def stats(self):
        st = aps_stats_t()
        capi.check(capi.lib().aps_stats(self.handle, C.byref(st)), "aps_stats")
        return {n: getattr(st, n) for n, _ in aps_stats_t._fields_}

    def save_snapshot(self, path):
        """Warm restart, SURVEY.md 8f-4: writes every resident page to `path`; returns the number written."""
        n = C.c_uint64()
        # ids that are neither decimal nor 32-hex live in this process's intern table: the snapshot carries it, so that
        # getPages() returns the same ids after a restart as LocalPageStore's directory walk would
        names = [nm.encode("utf-8") for nm in self.codec.names()]
        arr = (C.c_char_p * max(1, len(names)))(*names)
        self._raise(capi.lib().aps_snapshot_save_names(self.handle, os.fsencode(path), len(names), arr, C.byref(n)),
                    "snapshot save %s" % path)
        return n.value

    def load_snapshot(self, path):
        """Puts the pages of a snapshot into this store; returns (loaded, rejected).  getPages() lists them
        afterwards, which is what LocalCacheManager.restore walks (LocalCacheManager.java:466-509)."""
        n, r = C.c_uint64(), C.c_uint64()
        names = []
        cb_t = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_char_p)
        cb = cb_t(lambda user, index, name: names.append(name.decode("utf-8")) or 0)
        self._raise(capi.lib().aps_snapshot_names(os.fsencode(path), C.cast(cb, C.c_void_p), None), "snapshot names %s" % path)
        self.codec.adopt_names(names)
        self._raise(capi.lib().aps_snapshot_load(self.handle, os.fsencode(path), C.byref(n), C.byref(r)),
                    "snapshot load %s" % path)
        return n.value, r.value

    def debug_corrupt(self, page_id, byte_offset):
        k = self.key(page_id)
        self._raise(capi.lib().aps_debug_corrupt(self.handle, C.byref(k), byte_offset), "corrupt %r" % page_id)

    def get_batch(self, requests, stream=None):
        """requests: list of (PageId, page_offset, nbytes, dst_pointer).  Returns [(len, crc)];
        len < 0 is -(aps status)."""
        n = len(requests)
        reqs = (aps_get_req * max(1, n))()
        for i, (pid, off, nb, dst) in enumerate(requests):
            reqs[i] = aps_get_req(self.key(pid), off, nb, dst)
        res = (aps_result * max(1, n))()
        capi.check(capi.lib().aps_get_batch(self.handle, n, reqs, res, stream), "aps_get_batch")
        return [(res[i].len, res[i].crc) for i in range(n)]

    def put_batch(self, requests, stream=None):
        """requests: list of (PageId, length, src_pointer).  Returns [(len, crc)]."""
        n = len(requests)
        reqs = (aps_put_req * max(1, n))()
        for i, (pid, ln, src) in enumerate(requests):
            reqs[i] = aps_put_req(self.key(pid), ln, 0, src)
        res = (aps_result * max(1, n))()
        capi.check(capi.lib().aps_put_batch(self.handle, n, reqs, res, stream), "aps_put_batch")
        return [(res[i].len, res[i].crc) for i in range(n)]

    def delete_batch(self, page_ids, stream=None):
        n = len(page_ids)
        keys = (aps_key * max(1, n))()
        for i, pid in enumerate(page_ids):
            keys[i] = self.key(pid)
        res = (aps_result * max(1, n))()
        capi.check(capi.lib().aps_delete_batch(self.handle, n, keys, res, stream), "aps_delete_batch")
        return [res[i].len for i in range(n)]
