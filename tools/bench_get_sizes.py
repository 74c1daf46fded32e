#!/usr/bin/env python
"""Batched get with fused CRC32C across page sizes: every step moves the same 256 MiB (so every step is larger
than L2), split into pages of 4 KiB ... 1 MiB drawn uniformly from a 16 GiB resident set.  Shows how far the
copy kernel's task split holds up when pages are small (BASELINE configs[2] uses 64 KiB pages).
Prints one JSON line; run on a GPU box:  python tools/bench_get_sizes.py"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np
    import torch
    import gpu_util as gu
    from alluxio_b200 import capi
    from oracle import binding as orc
    L = capi.lib()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    step_bytes, resident_bytes = 256 << 20, 16 << 30
    peak = 6444.1
    pp = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pp):
        peak = float(json.load(open(pp))["hbm_gbs"])
    rows = []
    for page in (4096, 16384, 65536, 262144, 1048576):
        n_res, batch = resident_bytes // page, step_bytes // page
        fill_batch = min(n_res, max(batch, 4096))
        cfg = capi.aps_config(device=0, page_size=page, cache_bytes=n_res * page, slab_slots=n_res + 64, table_capacity=0,
                              max_batch=fill_batch, flags=(0 if os.environ.get('APS_SIZES_NOCRC') == '1' else capi.APS_FLAG_CRC) | capi.APS_FLAG_STABLE_REQS)
        h = C.c_void_p()
        capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "open")
        stage = torch.empty(fill_batch * page, dtype=torch.uint8, device=dev)
        res = torch.zeros(fill_batch * 8, dtype=torch.uint8, device=dev)
        ids = np.arange(n_res, dtype=np.uint64)
        for b0 in range(0, n_res, fill_batch):
            sel = ids[b0:b0 + fill_batch]
            f, i = sel // np.uint64(4096), sel % np.uint64(4096)
            df, di = torch.from_numpy(f.view(np.int64)).to(dev), torch.from_numpy(i.view(np.int64)).to(dev)
            capi.check(L.aps_fill_pages_dev(0, stage.data_ptr(), page, page, len(sel), df.data_ptr(), di.data_ptr(), 21, sp))
            r = np.zeros(len(sel), gu.PUT_REQ)
            r["file_lo"], r["page_index"], r["len"] = f, i, page
            r["src"] = stage.data_ptr() + np.arange(len(sel), dtype=np.uint64) * page
            capi.check(L.aps_put_batch_dev(h, len(sel), gu.to_dev(r).data_ptr(), res.data_ptr(), sp), "put")
        torch.cuda.synchronize()
        dst = torch.empty(step_bytes, dtype=torch.uint8, device=dev)
        rng = np.random.default_rng(5)
        pool = 24
        picks = rng.integers(0, n_res, size=(pool, batch)).astype(np.uint64)
        rq = np.zeros((pool, batch), gu.GET_REQ)
        rq["file_lo"], rq["page_index"] = picks // np.uint64(4096), picks % np.uint64(4096)
        rq["nbytes"] = page
        rq["dst"] = dst.data_ptr() + np.arange(batch, dtype=np.uint64)[None, :] * page
        d_rq = gu.to_dev(rq)
        d_res = torch.zeros(batch * 8, dtype=torch.uint8, device=dev)
        rb = batch * 40

        def step(k):
            capi.check(L.aps_get_batch_dev(h, batch, d_rq.data_ptr() + (k % pool) * rb, d_res.data_ptr(), sp), "get")

        for k in range(4):
            step(k)
        torch.cuda.synchronize()
        out = gu.from_dev(d_res, gu.RESULT)
        p = int(picks[3, 7])
        want = orc.fill_page(21, p // 4096, p % 4096, page)
        ok = bool((out["len"] == page).all()) and (not (cfg.flags & capi.APS_FLAG_CRC) or int(out["crc"][7]) == orc.crc32c(want)) and \
            np.array_equal(dst[7 * page:8 * page].cpu().numpy(), want)
        l0 = L.aps_total_kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        K = 100
        e0.record(stream)
        for k in range(K):
            step(4 + k)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        gbps = step_bytes / (ms * 1e-3) / 1e9
        rows.append({"page_bytes": page, "batch_pages": batch, "ms_per_step": round(ms, 4), "page_GBps": round(gbps, 1),
                     "frac_of_hbm_peak": round(2 * gbps / peak, 4), "launches_per_step": (L.aps_total_kernel_launches() - l0) / K,
                     "parity_ok": ok})
        capi.check(L.aps_close(h))
        del stage, dst, d_rq
        torch.cuda.empty_cache()
    print(json.dumps({"metric": "batched_page_get_fused_crc32c_GBps", "unit": "GB/s", "n_gpus": 1,
                      "config": {"workload": "256 MiB per step drawn from 16 GiB resident, page size swept"},
                      "hbm_peak_GBps": peak, "rows": rows}))
    return 0 if all(r["parity_ok"] for r in rows) else 1


if __name__ == "__main__":
    sys.exit(main())
