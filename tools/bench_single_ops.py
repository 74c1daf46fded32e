#!/usr/bin/env python
"""The reference-facing calls one page at a time (PageStore.put / get / delete, as the JNI shim issues them;
modelled on PageStoreTest.perfTest, CTEST/store/PageStoreTest.java:187-220): latency per call for 1 MiB and 64 KiB
pages with pinned and with pageable caller buffers.  Prints one JSON line.   python tools/bench_single_ops.py"""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    from alluxio_b200 import capi
    L = capi.lib()
    rows = []
    for page in (1 << 20, 64 << 10):
        n = 512
        cfg = capi.aps_config(device=0, page_size=page, cache_bytes=n * page, slab_slots=n + 8, table_capacity=0,
                              max_batch=64, flags=capi.APS_FLAG_CRC)
        h = C.c_void_p()
        capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "open")
        pinned = torch.empty(page, dtype=torch.uint8).pin_memory()
        pinned.numpy()[:] = np.random.default_rng(1).integers(0, 256, page, dtype=np.uint8)
        pageable = np.array(pinned.numpy(), copy=True)
        out_len, out_crc = C.c_uint32(), C.c_uint32()
        for kind, buf in (("pinned", pinned.data_ptr()), ("pageable", pageable.ctypes.data)):
            keys = [capi.aps_key(0, 7, i) for i in range(n)]
            t0 = time.perf_counter()
            for k in keys:
                capi.check(L.aps_put(h, C.byref(k), C.c_void_p(buf), page), "put")
            t_put = (time.perf_counter() - t0) / n
            t0 = time.perf_counter()
            for k in keys:
                capi.check(L.aps_get(h, C.byref(k), 0, page, C.c_void_p(buf), page, C.byref(out_len), C.byref(out_crc)), "get")
            t_get = (time.perf_counter() - t0) / n
            assert out_len.value == page
            t0 = time.perf_counter()
            for k in keys:
                capi.check(L.aps_delete(h, C.byref(k)), "delete")
            t_del = (time.perf_counter() - t0) / n
            rows.append({"page_bytes": page, "caller_buffer": kind, "put_us": round(t_put * 1e6, 1), "get_us": round(t_get * 1e6, 1),
                         "delete_us": round(t_del * 1e6, 1), "put_GBps": round(page / t_put / 1e9, 2),
                         "get_GBps": round(page / t_get / 1e9, 2)})
        capi.check(L.aps_close(h))
    print(json.dumps({"metric": "single_page_call_latency", "unit": "us", "n_gpus": 1,
                      "config": {"workload": "512 single-page put / get / delete calls through aps_put, aps_get, aps_delete (ctypes "
                                             "overhead of about 1 us per call included)"}, "rows": rows}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
