"""ctypes binding of libaps.so, the C ABI declared in include/aps.h.

This is the only way Python code in this repository reaches the data path; tests and bench.py call
through it so that what they exercise is exactly what a JNI shim would bind.  There is no fallback:
if libaps.so is missing or no CUDA device answers, calls fail loudly.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaps.so")

(APS_OK, APS_ENOTFOUND, APS_EINVAL, APS_ENOSPC, APS_ECUDA, APS_ECORRUPT, APS_ECLOSED, APS_ENOMEM, APS_EFAULT,
 APS_ELIMIT, APS_EIO, APS_ETIMEOUT) = range(12)
APS_NOALLOC_MAX_RANGES, APS_NOALLOC_MAX_LEN = 1024, 4 << 20
APS_FLAG_CRC, APS_FLAG_STABLE_REQS, APS_FLAG_DEVICE_CLOCK = 1, 4, 8
APS_FAULT_PUT, APS_FAULT_DELETE = 1, 2


class aps_key(C.Structure):
    _fields_ = [("file_hi", C.c_uint64), ("file_lo", C.c_uint64), ("page_index", C.c_uint64)]


class aps_config(C.Structure):
    _fields_ = [("device", C.c_int32), ("page_size", C.c_uint32), ("cache_bytes", C.c_uint64),
                ("slab_slots", C.c_uint64), ("table_capacity", C.c_uint64), ("max_batch", C.c_uint32),
                ("flags", C.c_uint32)]


class aps_get_req(C.Structure):
    _fields_ = [("key", aps_key), ("page_off", C.c_uint32), ("nbytes", C.c_uint32), ("dst", C.c_void_p)]


class aps_put_req(C.Structure):
    _fields_ = [("key", aps_key), ("len", C.c_uint32), ("reserved", C.c_uint32), ("src", C.c_void_p)]


class aps_result(C.Structure):
    _fields_ = [("len", C.c_int32), ("crc", C.c_uint32)]


class aps_shard_handle(C.Structure):
    _fields_ = [("opaque", C.c_ubyte * 256)]


class aps_push_handle(C.Structure):
    _fields_ = [("opaque", C.c_ubyte * 192)]


class aps_stats_t(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("pages", "bytes", "slots_free", "table_capacity", "tombstones", "lookups",
                                           "probe_buckets", "bytes_copied", "crc_mismatches", "kernel_launches", "slow_sms",
                                           "clock_evictions")]


ITER_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(aps_key), C.c_uint32)

# every symbol include/aps.h declares: name -> (restype, argtypes)
_vp, _u32, _u64, _i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
SIGNATURES = {
    "aps_open": (C.c_int, [C.POINTER(aps_config), C.POINTER(_vp)]),
    "aps_close": (C.c_int, [_vp]),
    "aps_put": (C.c_int, [_vp, C.POINTER(aps_key), _vp, _u32]),
    "aps_get": (C.c_int, [_vp, C.POINTER(aps_key), _u32, _u32, _vp, _u32, C.POINTER(_u32), C.POINTER(_u32)]),
    "aps_delete": (C.c_int, [_vp, C.POINTER(aps_key)]),
    "aps_cache_size": (_u64, [_vp]),
    "aps_iterate": (C.c_int, [_vp, ITER_CB, _vp]),
    "aps_get_batch": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_put_batch": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_delete_batch": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_get_batch_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_put_batch_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_delete_batch_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_lookup_batch_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_shard_of": (_u32, [C.POINTER(aps_key), _u32]),
    "aps_shard_export": (C.c_int, [_vp, C.POINTER(aps_shard_handle)]),
    "aps_shard_attach": (C.c_int, [_vp, _u32, _u32, _vp]),
    "aps_shard_attach_local": (C.c_int, [_vp, _u32, _u32, _vp]),
    "aps_get_batch_sharded_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_get_batch_sharded": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_push_setup": (C.c_int, [_vp, _u64, _u32]),
    "aps_push_landing": (_vp, [_vp, C.POINTER(_u64)]),
    "aps_push_export": (C.c_int, [_vp, C.POINTER(aps_push_handle)]),
    "aps_push_attach": (C.c_int, [_vp, _u32, _u32, C.POINTER(aps_push_handle)]),
    "aps_push_attach_local": (C.c_int, [_vp, _u32, _u32, _vp]),
    "aps_get_batch_push_begin_dev": (C.c_int, [_vp, _u32, _vp, _vp]),
    "aps_get_batch_push_end_dev": (C.c_int, [_vp, _vp, _vp]),
    "aps_get_batch_push_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_md5_batch_dev": (C.c_int, [_i32, _u32, _vp, _vp, _vp, _vp]),
    "aps_md5_batch": (C.c_int, [_i32, _u32, _vp, _vp, _vp]),
    "aps_md5_stream_dev": (C.c_int, [_i32, _u32, _vp, _vp, _vp, _vp, C.c_int, _vp, _vp]),
    "aps_crc32c_batch_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp]),
    "aps_crc32c_noalloc": (C.c_int, [_i32, _u32, _vp, _vp, _u32, _vp, _vp]),
    "aps_fill_pages_dev": (C.c_int, [_i32, _vp, _u64, _u32, _u32, _vp, _vp, _u64, _vp]),
    "aps_fill_pages": (C.c_int, [_vp, _vp, _u64, _u32, _u32, _vp, _vp, _u64, _vp]),
    "aps_device_alloc": (C.c_int, [_i32, _u64, C.POINTER(_vp)]),
    "aps_device_free": (C.c_int, [_i32, _vp]),
    "aps_host_alloc_pinned": (C.c_int, [_u64, C.POINTER(_vp)]),
    "aps_host_free_pinned": (C.c_int, [_vp]),
    "aps_compact": (C.c_int, [_vp]),
    "aps_snapshot_save": (C.c_int, [_vp, C.c_char_p, C.POINTER(_u64)]),
    "aps_snapshot_load": (C.c_int, [_vp, C.c_char_p, C.POINTER(_u64), C.POINTER(_u64)]),
    "aps_snapshot_save_names": (C.c_int, [_vp, C.c_char_p, C.c_uint32, C.POINTER(C.c_char_p), C.POINTER(_u64)]),
    "aps_snapshot_names": (C.c_int, [C.c_char_p, _vp, _vp]),
    "aps_clock_evict_dev": (C.c_int, [_vp, _u32, _vp, _vp, _vp]),
    "aps_stats": (C.c_int, [_vp, C.POINTER(aps_stats_t)]),
    "aps_set_faults": (C.c_int, [_vp, _u32]),
    "aps_profile_begin": (C.c_int, [_vp, _u32]),
    "aps_profile_end": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(_u32)]),
    "aps_debug_corrupt": (C.c_int, [_vp, C.POINTER(aps_key), _u32]),
    "aps_debug_timeline": (C.c_int, [_vp, C.c_int, _vp, _u32]),
    "aps_debug_crc_constants": (C.c_int, [_vp, _vp, _vp, _vp]),
    "aps_debug_crc_fold4": (C.c_int, [_vp]),
    "aps_strerror": (C.c_char_p, [C.c_int]),
    "aps_last_cuda_error": (C.c_char_p, []),
    "aps_version": (C.c_char_p, []),
    "aps_total_kernel_launches": (_u64, []),
}

_lib = None


def lib():
    """Loads libaps.so (built in-tree by alluxio_b200/build.py).  Raises if it is not there."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libaps.so is not built: run `python -m alluxio_b200.build` (needs nvcc); "
                               "there is no CPU fallback for the page store")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


class ApsError(RuntimeError):
    def __init__(self, code, where=""):
        self.code = code
        msg = lib().aps_strerror(code).decode()
        if code == APS_ECUDA:
            msg += ": " + lib().aps_last_cuda_error().decode()
        super().__init__("%s%s (aps status %d)" % (where + ": " if where else "", msg, code))


def check(code, where=""):
    if code != APS_OK:
        raise ApsError(code, where)
