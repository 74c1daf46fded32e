#!/usr/bin/env python
"""K1 on its own ("K1 hash ops"): Mops/s of batched lookups, inserts and erases on a table far
larger than L2, next to the measured random-access rate of HBM (tools/micro/random_sector_rate.cu, built and
run here when nvcc is present).  A probe reads one 256-byte bucket with eight lanes, so the bound is the rate
of random 256-byte reads; algorithmic bytes per op = 32 B entry x 8 entries x buckets probed.
Pages are 16 bytes so that the payload copy of a put is negligible beside its table work.
Prints one JSON line; run on a GPU box:  python tools/bench_table.py [--ids 8388608] [--batch 1048576]"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def micro():
    src = os.path.join(ROOT, "tools", "micro", "random_sector_rate.cu")
    exe = "/tmp/random_sector_rate"
    try:
        subprocess.run(["nvcc", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-o", exe, src], check=True,
                       capture_output=True, timeout=300)
        out = subprocess.run([exe, "8"], check=True, capture_output=True, timeout=120).stdout.decode()
        return [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    except Exception as e:  # the micro-benchmark is context, not the measurement
        return {"unavailable": str(e)[:200]}


def main():
    import numpy as np
    import torch
    import gpu_util as gu
    from alluxio_b200 import capi
    ap = argparse.ArgumentParser()
    ap.add_argument("--ids", type=int, default=8 << 20)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=8)
    args = ap.parse_args()
    peak = micro()
    L = capi.lib()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    n, nb = args.ids, args.batch
    page = 16
    cfg = capi.aps_config(device=0, page_size=page, cache_bytes=n * page, slab_slots=n, table_capacity=2 * n,
                          max_batch=nb, flags=0)
    st = C.c_void_p()
    capi.check(L.aps_open(C.byref(cfg), C.byref(st)), "aps_open")
    rng = np.random.default_rng(7)

    def keys_of(ids):
        k = np.zeros(len(ids), dtype=gu.KEY)
        k["file_lo"] = ids // 1000
        k["page_index"] = ids % 1000
        return k

    src = torch.zeros(nb * page, dtype=torch.uint8, device=dev)
    res = torch.zeros(nb * 8, dtype=torch.uint8, device=dev)
    lens = torch.zeros(nb, dtype=torch.int32, device=dev)

    def put_reqs(ids):
        r = np.zeros(len(ids), dtype=gu.PUT_REQ)
        r["file_lo"] = ids // 1000
        r["page_index"] = ids % 1000
        r["len"] = page
        r["src"] = src.data_ptr() + np.arange(len(ids), dtype=np.uint64) * page
        return gu.to_dev(r)

    def timed(fn, batches):
        ms = 0.0
        for b in batches:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn(b)
            e1.record(stream)
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return ms

    def do_put(d):
        capi.check(L.aps_put_batch_dev(st, nb, d.data_ptr(), res.data_ptr(), sp), "put")

    def do_lookup(d):
        capi.check(L.aps_lookup_batch_dev(st, nb, d.data_ptr(), lens.data_ptr(), sp), "lookup")

    def do_delete(d):
        capi.check(L.aps_delete_batch_dev(st, nb, d.data_ptr(), res.data_ptr(), sp), "delete")

    launches0 = L.aps_total_kernel_launches()
    # populate: ids 0..n-1 in batches
    pop = [put_reqs(np.arange(i, i + nb, dtype=np.uint64)) for i in range(0, n, nb)]
    ms_pop = timed(do_put, pop)
    del pop
    hit = [gu.to_dev(keys_of(rng.permutation(n)[:nb].astype(np.uint64))) for _ in range(args.reps)]
    miss = [gu.to_dev(keys_of((n + rng.permutation(n)[:nb]).astype(np.uint64))) for _ in range(args.reps)]
    timed(do_lookup, hit[:2])
    s0 = capi.aps_stats_t()
    L.aps_stats(st, C.byref(s0))
    ms_hit = timed(do_lookup, hit)
    ok_hit = int((lens == page).sum().item())
    s1 = capi.aps_stats_t()
    L.aps_stats(st, C.byref(s1))
    ms_miss = timed(do_lookup, miss)
    ok_miss = int((lens == -1).sum().item())
    s2 = capi.aps_stats_t()
    L.aps_stats(st, C.byref(s2))
    # churn: erase a random batch of resident ids, insert as many new ones, reps times
    ms_del = ms_ins = 0.0
    resident = np.arange(n, dtype=np.uint64)
    fresh = 2 * n
    for r in range(args.reps):
        pick = rng.permutation(n)[:nb]
        ms_del += timed(do_delete, [gu.to_dev(keys_of(resident[pick]))])
        new = np.arange(fresh, fresh + nb, dtype=np.uint64)
        fresh += nb
        ms_ins += timed(do_put, [put_reqs(new)])
        resident[pick] = new
    s3 = capi.aps_stats_t()
    L.aps_stats(st, C.byref(s3))
    final = [gu.to_dev(keys_of(resident[rng.permutation(n)[:nb]]))]
    timed(do_lookup, final)
    ok_final = int((lens == page).sum().item())
    # the table side of an insert alone: empty pages (nothing to copy), fresh ids into the places of a deleted batch
    ms_ins0 = 0.0
    for r in range(args.reps):
        pick = rng.permutation(n)[:nb]
        timed(do_delete, [gu.to_dev(keys_of(resident[pick]))])
        new = np.arange(fresh, fresh + nb, dtype=np.uint64)
        fresh += nb
        rq = np.zeros(nb, dtype=gu.PUT_REQ)
        rq["file_lo"], rq["page_index"], rq["len"] = new // 1000, new % 1000, 0
        rq["src"] = src.data_ptr()
        ms_ins0 += timed(do_put, [gu.to_dev(rq)])
        resident[pick] = new
    launches = int(L.aps_total_kernel_launches() - launches0)
    L.aps_close(st)

    def mops(count, ms):
        return round(count / (ms * 1e-3) / 1e6, 1)

    bph = (s1.probe_buckets - s0.probe_buckets) / max(1, s1.lookups - s0.lookups)
    bpm = (s2.probe_buckets - s1.probe_buckets) / max(1, s2.lookups - s1.lookups)
    line = {
        "metric": "hash_table_ops_Mops", "unit": "Mops/s", "n_gpus": 1,
        "config": {"workload": "K1 alone: %d resident ids in a %d-entry table (%d MiB, > L2), batches of %d, 16-byte pages"
                   % (n, 2 * n, 2 * n * 32 >> 20, nb)},
        "lookup_hit_Mops": mops(nb * args.reps, ms_hit), "lookup_miss_Mops": mops(nb * args.reps, ms_miss),
        "insert_populate_Mops": mops(n, ms_pop), "insert_churn_Mops": mops(nb * args.reps, ms_ins),
        "erase_Mops": mops(nb * args.reps, ms_del),
        "insert_empty_pages_Mops": mops(nb * args.reps, ms_ins0),
        "insert_note": "insert_* put 16-byte pages, so they include a copy-kernel pass in which a whole warp moves 16 bytes; "
                       "insert_empty_pages is the table side alone (claim an entry, pop a slot, publish)",
        "buckets_per_hit": round(bph, 4), "buckets_per_miss": round(bpm, 4),
        "tombstones_after_churn": int(s3.tombstones), "pages_after_churn": int(s3.pages),
        "checks": {"hits_found": ok_hit == nb, "misses_absent": ok_miss == nb, "resident_after_churn_found": ok_final == nb,
                   "pages_constant": int(s3.pages) == n},
        "gpu_launches": launches, "random_access_peak": peak,
    }
    if isinstance(peak, list) and len(peak) == 2:
        r256 = peak[1]["Mgranules_per_s"]
        line["roofline"] = {"bound": "hbm-random-256B", "peak": r256, "unit": "Mbuckets/s",
                            "achieved_hit": round(line["lookup_hit_Mops"] * bph, 1),
                            "frac_hit": round(line["lookup_hit_Mops"] * bph / r256, 3),
                            "frac_miss": round(line["lookup_miss_Mops"] * bpm / r256, 3)}
    print(json.dumps(line))
    return 0 if all(line["checks"].values()) else 1


if __name__ == "__main__":
    sys.exit(main())
