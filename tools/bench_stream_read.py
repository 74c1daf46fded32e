#!/usr/bin/env python
"""SURVEY.md 8f-1 measured: sequential reads of a fully cached file through the LocalCacheFileInStream mirror, one
CacheManager.get per page (the reference's loop, CACHE/LocalCacheFileInStream.java:121-145) against one batched store
call per read.  Prints one JSON line.   python tools/bench_stream_read.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    from alluxio_b200.cache import LocalCacheManager
    rows = []
    for page, read in ((64 << 10, 8 << 20), (1 << 20, 8 << 20), (64 << 10, 1 << 20)):
        file_len = 256 << 20
        data = np.random.default_rng(3).integers(0, 256, file_len, dtype=np.uint8).tobytes()
        m = LocalCacheManager.create(page, 2 * file_len, evictor="LRU", slab_slots=file_len // page + 64, max_batch=4096)
        warm = m.open_file("9", file_len, lambda pos, n: data[pos:pos + n])
        buf = bytearray(read)
        while warm.read(buf) > 0:                           # fill the cache
            pass
        out = {}
        for mode in ("per_page", "batched"):
            s = m.open_file("9", file_len, lambda pos, n: data[pos:pos + n])
            if mode == "batched":
                s.set_batched()
            s.read(buf)                                     # first call allocates the staging buffer
            s.seek(0)
            t0 = time.perf_counter()
            total = 0
            while True:
                r = s.read(buf)
                if r <= 0:
                    break
                total += r
            dt = time.perf_counter() - t0
            assert total == file_len and bytes(buf[:1024]) == data[file_len - read:file_len - read + 1024]
            out[mode] = round(total / dt / 1e9, 2)
        rows.append({"page_bytes": page, "read_bytes": read, "per_page_GBps": out["per_page"], "batched_GBps": out["batched"],
                     "speedup": round(out["batched"] / out["per_page"], 2)})
        m.close()
    print(json.dumps({"metric": "cached_stream_read_GBps", "unit": "GB/s", "n_gpus": 1,
                      "config": {"workload": "sequential reads of a fully cached 256 MiB file into a pageable buffer through the "
                                             "stream object: one get per page against one batched store call per read"},
                      "rows": rows}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
