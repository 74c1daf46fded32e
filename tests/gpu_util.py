"""Helpers for the -m gpu tests: device buffers via torch (plumbing only) and request arrays in the
exact layout of include/aps.h."""
import ctypes as C

import numpy as np
import torch

from alluxio_b200 import capi

GET_REQ = np.dtype([("file_hi", "<u8"), ("file_lo", "<u8"), ("page_index", "<u8"), ("page_off", "<u4"),
                    ("nbytes", "<u4"), ("dst", "<u8")])
PUT_REQ = np.dtype([("file_hi", "<u8"), ("file_lo", "<u8"), ("page_index", "<u8"), ("len", "<u4"),
                    ("reserved", "<u4"), ("src", "<u8")])
KEY = np.dtype([("file_hi", "<u8"), ("file_lo", "<u8"), ("page_index", "<u8")])
RESULT = np.dtype([("len", "<i4"), ("crc", "<u4")])
assert GET_REQ.itemsize == C.sizeof(capi.aps_get_req) == 40
assert PUT_REQ.itemsize == C.sizeof(capi.aps_put_req) == 40
assert KEY.itemsize == C.sizeof(capi.aps_key) == 24
assert RESULT.itemsize == C.sizeof(capi.aps_result) == 8


def to_dev(arr, device="cuda:0"):
    """numpy (structured) array -> uint8 device tensor with the same bytes."""
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy()).to(device)


def from_dev(t, dtype):
    return t.cpu().numpy().view(dtype)


def cur_stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def fill_pages(dst, stride, page_len, files, indices, seed, device=0):
    """aps_fill_pages_dev over a uint8 device tensor."""
    f = torch.as_tensor(np.asarray(files, dtype=np.uint64).view(np.int64), device=dst.device)
    i = torch.as_tensor(np.asarray(indices, dtype=np.uint64).view(np.int64), device=dst.device)
    capi.check(capi.lib().aps_fill_pages_dev(device, dst.data_ptr(), stride, page_len, len(files), f.data_ptr(),
                                             i.data_ptr(), seed, cur_stream()), "aps_fill_pages_dev")
    torch.cuda.synchronize()


class _RawDeviceMemory:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def raw_tensor(ptr, nbytes, device="cuda:0"):
    """A uint8 tensor over device memory somebody else allocated (e.g. the library's landing buffer); no copy."""
    with torch.cuda.device(device):
        return torch.as_tensor(_RawDeviceMemory(ptr, nbytes), device=device)
