#!/usr/bin/env python
"""BASELINE configs[2]: small-page stress.  64 KiB pages, Zipf(0.99) accesses over 10 M PageIds
(10,000 files x 1,000 pages), capacity 2,097,152 pages (128 GiB), LRU evictor as the reference behaves
(a hit does not refresh recency), miss => put.  The evictor runs on the host and drives GPU
hash-table inserts and erases in batches (apc_replay_batched).  Prints one JSON line with the hit ratio, the
store-operation rates and, with --check, the comparison with the oracle's replay of the same trace.
  python tools/bench_small_pages.py [--accesses 10000000] [--capacity-pages 2097152] [--check]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def zipf_trace(n_ids, n_acc, theta, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    # rank -> id by a fixed bijective scramble; ranks drawn by inverse-CDF over the exact Zipf weights
    w = np.arange(1, n_ids + 1, dtype=np.float64) ** -theta
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    ranks = np.searchsorted(cdf, rng.random(n_acc))
    scramble = (np.arange(n_ids, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) >> np.uint64(11)) % np.uint64(n_ids)
    perm = np.argsort(scramble, kind="stable")
    return perm[ranks].astype(np.int64)


def main():
    import numpy as np
    from alluxio_b200.cache import LocalCacheManager
    ap = argparse.ArgumentParser()
    ap.add_argument("--accesses", type=int, default=10_000_000)
    ap.add_argument("--ids", type=int, default=10_000_000)
    ap.add_argument("--capacity-pages", type=int, default=2_097_152)
    ap.add_argument("--page-kib", type=int, default=64)
    ap.add_argument("--window", type=int, default=4096)
    ap.add_argument("--evictor", default="LRU")
    ap.add_argument("--check", action="store_true", help="also replay on the CPU oracle and compare everything")
    args = ap.parse_args()
    ps = args.page_kib << 10
    t0 = time.time()
    ids = zipf_trace(args.ids, args.accesses, 0.99, 3)
    files, pages = ids // 1000, ids % 1000
    lens = np.full(args.accesses, ps, np.int32)
    t_trace = time.time() - t0
    m = LocalCacheManager(ps, args.capacity_pages * ps, evictor=args.evictor, slab_slots=args.capacity_pages + 64,
                          max_batch=args.window)
    t0 = time.time()
    got = m.replay_batched(files, pages, lens, seed=3, window=args.window, victims_cap=1 if not args.check else None)
    wall = time.time() - t0
    st = m.store_stats()
    line = {
        "metric": "small_page_stress", "n_gpus": 1,
        "config": {"workload": "%d KiB pages, Zipf(0.99) over %d PageIds, %d accesses, capacity %d pages, %s evictor on the "
                   "host driving GPU table inserts/erases in batches of %d (BASELINE configs[2])"
                   % (args.page_kib, args.ids, args.accesses, args.capacity_pages, args.evictor, args.window)},
        "hit_ratio": round(float(got["hit"].mean()), 4), "victims": int(got["n_victims"]),
        "store_ops": {"gets": got["gets"], "puts": got["puts"], "deletes": got["deletes"], "flushes": got["flushes"]},
        "seconds": {"trace": round(t_trace, 2), "host_policy": round(got["seconds_host"], 2),
                    "store_batches": round(got["seconds_store"], 2), "wall": round(wall, 2)},
        "store_Mops_per_s": round((got["gets"] + got["puts"] + got["deletes"]) / got["seconds_store"] / 1e6, 3),
        "store_payload_GBps": round((got["gets"] + got["puts"]) * ps / got["seconds_store"] / 1e9, 1),
        "accesses_per_s_end_to_end": round(args.accesses / wall),
        "table": {"lookups": st["lookups"], "buckets_per_lookup": round(st["probe_buckets"] / max(1, st["lookups"]), 4),
                  "tombstones": st["tombstones"], "resident_pages": st["pages"], "capacity": st["table_capacity"]},
        "bad_gets": got["bad_gets"], "crc_mismatches": st["crc_mismatches"],
    }
    if args.check:
        from oracle import binding as orc
        t0 = time.time()
        want = orc.replay(args.evictor, ps, args.capacity_pages * ps, files, pages, lens)
        line["oracle"] = {"seconds": round(time.time() - t0, 2),
                          "hit_vector_equal": bool(np.array_equal(got["hit"], want["hit"])),
                          "put_results_equal": bool(np.array_equal(got["put_ok"], want["put_ok"])),
                          "victim_sequence_equal": bool(got["n_victims"] == want["n_victims"] and
                                                        np.array_equal(got["victims"], want["victims"])),
                          "final_residency_equal": (got["final_pages"], got["final_bytes"]) ==
                                                   (want["final_pages"], want["final_bytes"])}
    m.close()
    print(json.dumps(line))


if __name__ == "__main__":
    main()
