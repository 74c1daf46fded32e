#!/usr/bin/env python
"""BASELINE configs[2] with the policy ON THE DEVICE (opt-in APS_FLAG_DEVICE_CLOCK): 64 KiB pages,
Zipf(0.99) over 10 M PageIds, capacity 2,097,152 pages.  Per window of accesses: one batched get (hits are copied
and CRC-checked), the distinct misses are generated and put, and the store makes room by itself with a CLOCK
sweep.  Nothing goes through the host except the trace.  CLOCK is not the reference's evictor, so this run is NOT a
parity run (tools/bench_small_pages.py is): it shows what the same store does when the host policy (8 M
accesses/s) is taken out of the loop.  Prints one JSON line.
  python tools/bench_device_clock.py [--accesses 50000000] [--window 16384]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    import numpy as np
    import torch
    from alluxio_b200 import capi
    from bench_small_pages import zipf_trace
    ap = argparse.ArgumentParser()
    ap.add_argument("--accesses", type=int, default=50_000_000)
    ap.add_argument("--ids", type=int, default=10_000_000)
    ap.add_argument("--capacity-pages", type=int, default=2_097_152)
    ap.add_argument("--page-kib", type=int, default=64)
    ap.add_argument("--window", type=int, default=16384)
    args = ap.parse_args()
    ps, W = args.page_kib << 10, args.window
    L = capi.lib()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    t0 = time.time()
    ids_h = zipf_trace(args.ids, args.accesses, 0.99, 3)
    t_trace = time.time() - t0
    cfg = capi.aps_config(device=0, page_size=ps, cache_bytes=args.capacity_pages * ps, slab_slots=args.capacity_pages,
                          table_capacity=0, max_batch=W,
                          flags=capi.APS_FLAG_CRC | capi.APS_FLAG_DEVICE_CLOCK | capi.APS_FLAG_STABLE_REQS)
    h = C.c_void_p()
    capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "open")
    sink = torch.empty(W * ps, dtype=torch.uint8, device=dev)
    stage = torch.empty(W * ps, dtype=torch.uint8, device=dev)
    res = torch.zeros((W, 2), dtype=torch.int32, device=dev)
    slot_off = torch.arange(W, dtype=torch.int64, device=dev) * ps
    ids = torch.from_numpy(ids_h).to(dev)
    n = args.accesses
    hits = torch.zeros((), dtype=torch.int64, device=dev)
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    puts = 0
    launches0 = L.aps_total_kernel_launches()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record(stream)
    for w0 in range(0, n, W):
        w = ids[w0:w0 + W]
        m = w.numel()
        rq = torch.zeros((m, 5), dtype=torch.int64, device=dev)
        rq[:, 1], rq[:, 2] = w // 1000, w % 1000
        rq[:, 3] = ps << 32                                   # page_off 0, nbytes = page
        rq[:, 4] = sink.data_ptr() + slot_off[:m]
        capi.check(L.aps_get_batch_dev(h, m, rq.data_ptr(), res.data_ptr(), sp), "get")
        ln = res[:m, 0]
        miss = ln < 0
        hits += (ln == ps).sum()
        bad += ((ln >= 0) & (ln != ps)).sum() + (miss & (ln != -capi.APS_ENOTFOUND)).sum()
        new = torch.unique(w[miss])                           # (sizes come back to the host here)
        k = new.numel()
        if k:
            f, i = new // 1000, new % 1000
            capi.check(L.aps_fill_pages_dev(0, stage.data_ptr(), ps, ps, k, f.data_ptr(), i.data_ptr(), 3, sp), "fill")
            pq = torch.zeros((k, 5), dtype=torch.int64, device=dev)
            pq[:, 1], pq[:, 2], pq[:, 3] = f, i, ps
            pq[:, 4] = stage.data_ptr() + slot_off[:k]
            capi.check(L.aps_put_batch_dev(h, k, pq.data_ptr(), res.data_ptr(), sp), "put")
            bad += (res[:k, 0] != ps).sum()
            puts += k
    e1.record(stream)
    torch.cuda.synchronize()
    wall = time.time() - t0
    gpu_s = e0.elapsed_time(e1) * 1e-3
    st = capi.aps_stats_t()
    capi.check(L.aps_stats(h, C.byref(st)), "stats")
    line = {
        "metric": "small_page_stress_device_policy", "n_gpus": 1,
        "config": {"workload": "%d KiB pages, Zipf(0.99) over %d PageIds, %d accesses, capacity %d pages; device-side CLOCK "
                   "(opt-in, not the reference's evictor), windows of %d accesses: batched get, distinct misses put"
                   % (args.page_kib, args.ids, n, args.capacity_pages, W)},
        "accesses_per_s": round(n / wall), "seconds": {"trace": round(t_trace, 2), "wall": round(wall, 2), "gpu": round(gpu_s, 2)},
        "hit_ratio": round(int(hits.item()) / n, 4), "puts": puts, "clock_evictions": int(st.clock_evictions),
        "payload_GBps": round((int(hits.item()) + puts) * ps / wall / 1e9, 1),
        "resident_pages": int(st.pages), "tombstones": int(st.tombstones),
        "buckets_per_lookup": round(st.probe_buckets / max(1, st.lookups), 4),
        "bad_results": int(bad.item()), "crc_mismatches": int(st.crc_mismatches),
        "gpu_launches": int(L.aps_total_kernel_launches() - launches0),
        "compare": "host LRU/FIFO policy on the same trace: profiles/r1_bench_small_pages.json (hit ratio 0.8253, 7.7 M accesses/s)",
    }
    capi.check(L.aps_close(h))
    print(json.dumps(line))
    return 0 if line["bad_results"] == 0 and line["crc_mismatches"] == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
