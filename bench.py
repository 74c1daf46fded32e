#!/usr/bin/env python
"""Benchmark of the page-cache hot path (BASELINE.json configs[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One step = one batched get of 256 x 1 MiB pages, drawn uniformly from a 64 GiB HBM-resident set,
with the CRC32C integrity check fused into the copy.  `value` times the device-resident form
(requests and destinations already in HBM) with CUDA events; `e2e` times the host-facing C-ABI call
aps_get_batch() with host request arrays and PINNED HOST destinations, so every page crosses PCIe
inside the timed region.  N > 1: one process per GPU (torchrun), one shard per GPU, no data-path
collective (weak scaling).  `--impl reference` times the CPU restatement of the reference's
LocalCacheManager + LocalPageStore on the host cores for the same workload.
"""
import argparse
import ctypes as C
import json
import os
import shutil
import statistics
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MB = 1 << 20
PAGE = MB
BATCH = 256
DATA_SEED = 0xA11C10
METRIC = "batched_page_get_fused_crc32c_GBps"
UNIT = "GB/s"


def workload_name(resident_pages):
    return ("GpuPageStore on 1 B200 per rank: 1 MiB pages, %d GiB resident in HBM, uniform random batched get of "
            "%d pages with fused CRC32C (BASELINE configs[1])" % (resident_pages * PAGE >> 30, BATCH))


# ------------------------------------------------------------------------------------------------ CPU legs

def cpu_sample_dir():
    for base in ("/dev/shm", tempfile.gettempdir()):
        try:
            if shutil.disk_usage(base).free > 3 << 30:
                return tempfile.mkdtemp(prefix="aps_cpu_", dir=base), base
        except OSError:
            pass
    return tempfile.mkdtemp(prefix="aps_cpu_"), tempfile.gettempdir()


def cpu_reference_run(steps, warmup, threads, sample_pages=2048, gets_per_step=BATCH):
    """Reference CPU path (oracle port of LocalCacheManager.get -> LocalPageStore.get, one file per
    page on a real directory) on `threads` host threads.  Returns (GB/s, ms_per_step, description)."""
    from oracle import binding as orc
    root, base = cpu_sample_dir()
    try:
        n_files = 16
        ppf = sample_pages // n_files
        orc.bench_local_get(root, PAGE, n_files, ppf, True, 64, DATA_SEED, 1, threads)  # populate + touch
        for w in range(warmup):
            orc.bench_local_get(root, PAGE, n_files, ppf, False, gets_per_step, DATA_SEED, 100 + w, threads)
        secs, nbytes = 0.0, 0
        for k in range(steps):
            s, b = orc.bench_local_get(root, PAGE, n_files, ppf, False, gets_per_step, DATA_SEED, 1000 + k, threads)
            secs += s
            nbytes += b
        desc = ("%d uniform random whole-page gets per step over a %d GiB file set (%d x 1 MiB page files under %s, "
                "resident in the OS page cache: the reference's best case)" %
                (gets_per_step, sample_pages * PAGE >> 30, sample_pages, base))
        return nbytes / secs / 1e9, secs / max(1, steps) * 1e3, desc
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import binding as orc
    orc.build()
    threads = os.cpu_count() or 1
    steps = max(1, min(args.steps, 200))
    gbps, ms, desc = cpu_reference_run(steps, min(args.warmup, 5), threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbps, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 5), "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(65536), "page_bytes": PAGE, "batch_pages": BATCH},
        "cpu_baseline": {"value": round(gbps, 3), "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": round(gbps, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is Java and cannot be built here (no JVM); this is the oracle's C++ port of "
                "LocalCacheManager+LocalPageStore, CRC not included because the reference has none",
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------ clocks

class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every ~10 ms through NVML while running."""

    def __init__(self, index, uuid=None):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode()) if uuid else None
            except Exception:  # noqa: BLE001
                self.h = None
            if self.h is None:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.01)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        busy = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------ GPU arm

def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from alluxio_b200 import capi
    from alluxio_b200 import build as build_mod

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the page store has no CPU fallback")
    if rank == 0:
        build_mod.build_libaps()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    L = capi.lib()
    dev = torch.device("cuda", local_rank)
    K, W = args.steps, max(3, args.warmup)

    resident = int(os.environ.get("APS_BENCH_PAGES", "65536"))  # pages per GPU
    free_b, _ = torch.cuda.mem_get_info(dev)
    while resident * PAGE + (8 << 30) > free_b and resident > 1024:
        resident //= 2  # a smaller GPU slice: stated in config.workload
    if world > 1:
        t = torch.tensor([resident], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        resident = int(t.item())
    # Global id space: world * resident pages; page g is (file g // 1024, index g % 1024).  With several
    # GPUs the cache is sharded: GPU aps_shard_of(key) owns the page (hash routing).
    from alluxio_b200 import sharding
    total_pages = world * resident
    gids = np.arange(total_pages, dtype=np.uint64)
    if world > 1:
        mine = sharding.owned_by(rank, world, gids // np.uint64(1024), gids % np.uint64(1024))
        owned = gids[mine]
    else:
        owned = gids
    slots = int(len(owned)) + 64
    cfg = capi.aps_config(device=local_rank, page_size=PAGE, cache_bytes=slots * PAGE, slab_slots=slots,
                          table_capacity=0, max_batch=BATCH,
                          flags=0 if os.environ.get("APS_BENCH_NOCRC") == "1" else capi.APS_FLAG_CRC)
    with_crc = cfg.flags != 0
    # the request pool below is written once, long before the timed steps: the promise APS_FLAG_STABLE_REQS asks for
    if os.environ.get("APS_BENCH_NO_EARLY") != "1":
        cfg.flags |= capi.APS_FLAG_STABLE_REQS
    h = C.c_void_p()
    capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "aps_open")
    # one explicit non-default stream for everything: torch copies, our kernels and the timing events
    # (a NULL aps_stream would mean "the store's own stream", which torch events cannot see)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    assert sp.value, "need a real stream handle"

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from gpu_util import GET_REQ, PUT_REQ, RESULT

    # ---- populate the pages this GPU owns: generate on the device, put 256 at a time
    stage = torch.empty(BATCH * PAGE, dtype=torch.uint8, device=dev)
    d_res = torch.zeros(BATCH * 8, dtype=torch.uint8, device=dev)
    def populate(store):
        first = None
        for b0 in range(0, len(owned), BATCH):
            ids = owned[b0:b0 + BATCH]
            nb = len(ids)
            files, idx = ids // np.uint64(1024), ids % np.uint64(1024)
            d_files = torch.from_numpy(files.view(np.int64)).to(dev)
            d_idx = torch.from_numpy(idx.view(np.int64)).to(dev)
            capi.check(L.aps_fill_pages_dev(local_rank, stage.data_ptr(), PAGE, PAGE, nb, d_files.data_ptr(),
                                            d_idx.data_ptr(), DATA_SEED, sp), "fill")
            reqs = np.zeros(nb, PUT_REQ)
            reqs["file_lo"], reqs["page_index"], reqs["len"] = files, idx, PAGE
            reqs["src"] = stage.data_ptr() + np.arange(nb, dtype=np.uint64) * PAGE
            d_reqs = torch.from_numpy(reqs.view(np.uint8).reshape(-1)).to(dev)
            capi.check(L.aps_put_batch_dev(store, nb, d_reqs.data_ptr(), d_res.data_ptr(), sp), "put")
            if b0 == 0:
                torch.cuda.synchronize(dev)
                first = d_res.cpu().numpy().view(RESULT).copy()
        torch.cuda.synchronize(dev)
        return first

    t_pop = time.time()
    first_put_res = populate(h)
    t_pop = time.time() - t_pop

    # ---- parity spot check against the oracle (the checker; never the thing measured)
    from oracle import binding as orc
    for i in (0, 1, 255):
        g = int(owned[i])
        want = orc.crc32c(orc.fill_page(DATA_SEED, g // 1024, g % 1024, PAGE))
        assert (not with_crc or int(first_put_res[i]["crc"]) == want) and int(first_put_res[i]["len"]) == PAGE, \
            "put CRC differs from oracle"

    # ---- several GPUs: map every peer's table and slab (CUDA IPC handles travel over torch.distributed)
    if world > 1:
        hnd = capi.aps_shard_handle()
        capi.check(L.aps_shard_export(h, C.byref(hnd)), "shard_export")
        blobs = [None] * world
        dist.all_gather_object(blobs, bytes(hnd.opaque))
        arr = (capi.aps_shard_handle * world)()
        for i, bl in enumerate(blobs):
            C.memmove(C.addressof(arr[i]), bl, 256)
        capi.check(L.aps_shard_attach(h, world, rank, arr), "shard_attach")
        dist.barrier()

    # ---- request pool: uniform random page ids over the whole (global) id space, seed 2 + rank
    rng = np.random.default_rng(2 + rank)
    pool = min(K + W, 512)
    dst = torch.empty(BATCH * PAGE, dtype=torch.uint8, device=dev)
    picks = rng.integers(0, total_pages, size=(pool, BATCH)).astype(np.uint64)
    reqs = np.zeros((pool, BATCH), GET_REQ)
    reqs["file_lo"], reqs["page_index"] = picks // np.uint64(1024), picks % np.uint64(1024)
    reqs["page_off"], reqs["nbytes"] = 0, PAGE
    reqs["dst"] = dst.data_ptr() + np.arange(BATCH, dtype=np.uint64)[None, :] * PAGE
    d_reqs = torch.from_numpy(reqs.view(np.uint8).reshape(-1)).to(dev)
    d_res = torch.zeros(BATCH * 8, dtype=torch.uint8, device=dev)
    req_bytes = BATCH * GET_REQ.itemsize
    if world > 1:
        pick_owner = sharding.shard_of_keys(np.zeros(picks.size, np.uint64), (picks // np.uint64(1024)).reshape(-1),
                                            (picks % np.uint64(1024)).reshape(-1), world)
        remote_frac = float((pick_owner != rank).mean())
    else:
        remote_frac = 0.0
    get_dev = L.aps_get_batch_sharded_dev if world > 1 else L.aps_get_batch_dev
    get_host = L.aps_get_batch_sharded if world > 1 else L.aps_get_batch

    def step_dev(k):
        capi.check(get_dev(h, BATCH, d_reqs.data_ptr() + (k % pool) * req_bytes, d_res.data_ptr(), sp), "get")

    for k in range(W):
        step_dev(k)
    torch.cuda.synchronize(dev)
    # correctness of the last warm-up step, against the oracle, before anything is timed
    res = d_res.cpu().numpy().view(RESULT)
    k_last = (W - 1) % pool
    for i in (0, 17, 255):
        p = int(picks[k_last, i])
        page = orc.fill_page(DATA_SEED, p // 1024, p % 1024, PAGE)
        assert int(res[i]["len"]) == PAGE and (not with_crc or int(res[i]["crc"]) == orc.crc32c(page)), \
            "get CRC differs from oracle"
        assert np.array_equal(dst[i * PAGE:(i + 1) * PAGE].cpu().numpy(), page), "get bytes differ from oracle"

    if os.environ.get("APS_BENCH_TIMELINE") == "1":  # diagnostic: per-CTA / per-warp timestamps of single launches
        none = np.uint64(0xFFFFFFFFFFFFFFFF)
        per_sm = []
        capi.check(L.aps_debug_timeline(h, 1, None, 0))
        for rep in range(4):
            for k in range(3):
                step_dev(rep * 3 + k)
            raw = np.zeros(1024 + 256 * 32 + 256, np.uint64)
            capi.check(L.aps_debug_timeline(h, 1, raw.ctypes.data, raw.size))
            tl = raw[:1024].reshape(256, 4)
            live = tl[:, 0] != none
            tl = tl[live].astype(np.int64)
            t0 = tl[:, 0].min()
            rel = (tl - t0) / 1e3
            smid = raw[1024 + 256 * 32:][live].astype(np.int64)
            per_sm.append(dict(zip(smid.tolist(), rel[:, 3].tolist())))
            if rep:
                continue
            print("timeline us: ctas=%d start[min/med/max]=%.1f/%.1f/%.1f tables=%.1f/%.1f/%.1f firstwarp=%.1f/%.1f/%.1f "
                  "end=%.1f/%.1f/%.1f" % ((len(rel),) + tuple(x for c in range(4) for x in
                                                              (rel[:, c].min(), np.median(rel[:, c]), rel[:, c].max()))),
                  file=sys.stderr)
            w = raw[1024:1024 + 256 * 32].reshape(256, 32)[live]
            w = np.where(w == none, 0, w).astype(np.int64)
            nw = int((w[0] != 0).sum())
            wrel = (w[:, :nw] - t0) / 1e3
            print("timeline warp end spread within CTA (max-min) median %.1f us; CTA end: p10 %.1f p50 %.1f p90 %.1f max %.1f"
                  % (np.median(wrel.max(axis=1) - wrel.min(axis=1)), *np.percentile(rel[:, 3], [10, 50, 90, 100])),
                  file=sys.stderr)
        common = sorted(set.intersection(*[set(d) for d in per_sm]))
        m = np.array([[d[sm] for sm in common] for d in per_sm])
        m = m - np.median(m, axis=1, keepdims=True)
        cc = np.corrcoef(m)
        print("timeline per-SM end offset: correlation between launches %s; SMs seen %d (ids %d..%d)" %
              (np.round(cc[0, 1:], 2).tolist(), len(common), common[0], common[-1]), file=sys.stderr)
        off = np.sort(m.mean(axis=0))
        print("timeline per-SM mean end offset us, sorted (every 6th): " + " ".join("%.1f" % x for x in off[::6]), file=sys.stderr)
        capi.check(L.aps_debug_timeline(h, 0, None, 0))
        slow = np.argsort(-m.mean(axis=0))[:8]
        print("timeline slowest SMs (id: mean offset us over 4 launches): " +
              ", ".join("%d: %+.1f" % (common[i], m[:, i].mean()) for i in slow), file=sys.stderr)

    sampler = ClockSampler(local_rank, getattr(torch.cuda.get_device_properties(dev), "uuid", None))
    sampler.start()

    # ---- timed region: device-resident batched get with fused CRC (value + roofline)
    launches0 = L.aps_total_kernel_launches()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for k in range(K):
        step_dev(W + k)
    ev1.record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches = L.aps_total_kernel_launches() - launches0
    assert launches == K, "a get step must be exactly one kernel launch (lookup + gather + CRC fused)"
    res = d_res.cpu().numpy().view(RESULT)
    assert (res["len"] == PAGE).all(), "a timed get did not return a full page"
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())

    # ---- end to end: host request arrays, pinned host destinations, results read back
    quick = os.environ.get("APS_BENCH_QUICK") == "1"  # kernel sweeps: skip the slow PCIe and CPU legs
    e2e_steps = 2 if quick else max(1, min(K, 100))
    pinned = torch.empty(BATCH * PAGE, dtype=torch.uint8).pin_memory()
    h_reqs = reqs.copy()
    h_reqs["dst"] = pinned.data_ptr() + np.arange(BATCH, dtype=np.uint64)[None, :] * PAGE
    h_res = np.zeros(BATCH, RESULT)

    def step_host(k):
        capi.check(get_host(h, BATCH, h_reqs[k % pool].ctypes.data, h_res.ctypes.data, sp), "get_batch")

    for k in range(3):
        step_host(k)
    p = int(picks[2 % pool, 5])
    page = orc.fill_page(DATA_SEED, p // 1024, p % 1024, PAGE)
    assert np.array_equal(pinned.numpy()[5 * PAGE:6 * PAGE], page)
    assert not with_crc or int(h_res[5]["crc"]) == orc.crc32c(page)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        step_host(3 + k)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert (h_res["len"] == PAGE).all()
    # what the link allows: a plain copy-engine transfer of the same 256 MiB from HBM to the pinned buffer
    pcie_gbps = None
    if not quick:
        dsrc = dst[:BATCH * PAGE]
        for _ in range(2):
            pinned.copy_(dsrc, non_blocking=True)
        torch.cuda.synchronize(dev)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        for _ in range(5):
            pinned.copy_(dsrc, non_blocking=True)
        c1.record(stream)
        torch.cuda.synchronize(dev)
        pcie_gbps = 5 * BATCH * PAGE / (c0.elapsed_time(c1) * 1e-3) / 1e9

    # ---- auxiliary numbers (not the headline): put with CRC
    aux = {}
    aux_steps = max(1, min(K, 50))
    put_reqs = np.zeros(BATCH, PUT_REQ)
    ids = owned[:BATCH]
    put_reqs["file_lo"], put_reqs["page_index"], put_reqs["len"] = ids // np.uint64(1024), ids % np.uint64(1024), PAGE
    put_reqs["src"] = stage.data_ptr() + np.arange(BATCH, dtype=np.uint64) * PAGE  # re-put this GPU's first 256 pages
    files0 = torch.from_numpy((ids // np.uint64(1024)).view(np.int64)).to(dev)
    idx0 = torch.from_numpy((ids % np.uint64(1024)).view(np.int64)).to(dev)
    capi.check(L.aps_fill_pages_dev(local_rank, stage.data_ptr(), PAGE, PAGE, BATCH, files0.data_ptr(), idx0.data_ptr(),
                                    DATA_SEED, sp), "fill")
    d_put = torch.from_numpy(put_reqs.view(np.uint8).reshape(-1)).to(dev)
    for _ in range(3):
        capi.check(L.aps_put_batch_dev(h, BATCH, d_put.data_ptr(), d_res.data_ptr(), sp), "put")
    torch.cuda.synchronize(dev)
    ev0.record(stream)
    for _ in range(aux_steps):
        capi.check(L.aps_put_batch_dev(h, BATCH, d_put.data_ptr(), d_res.data_ptr(), sp), "put")
    ev1.record(stream)
    torch.cuda.synchronize(dev)
    aux["put_fused_crc32c_GBps"] = round(aux_steps * BATCH * PAGE / (ev0.elapsed_time(ev1) * 1e-3) / 1e9, 1)
    aux["put_note"] = "re-puts one resident 256 MiB batch (source and slots cached in L2 between steps): upper bound"

    # ---- several GPUs: the same multi-get through NCCL (all-to-all of keys, owner-side gather into a send
    # buffer, all-to-all of pages): the plain-collective baseline that the fused P2P kernel is compared with
    if world > 1 and not quick:
        nccl_steps = max(1, min(K, 50))
        sendbuf = torch.empty(2 * BATCH * PAGE, dtype=torch.uint8, device=dev)   # pages this GPU serves per step
        recvbuf = torch.empty(BATCH * PAGE, dtype=torch.uint8, device=dev)
        srv_res = torch.zeros(2 * BATCH * 8, dtype=torch.uint8, device=dev)

        def step_nccl(k):
            pk = picks[k % pool]
            own = sharding.shard_of_keys(np.zeros(BATCH, np.uint64), pk // np.uint64(1024), pk % np.uint64(1024), world)
            order = np.argsort(own, kind="stable")
            counts = np.bincount(own, minlength=world)
            keys = torch.from_numpy(np.stack([pk[order] // np.uint64(1024), pk[order] % np.uint64(1024)], 1).astype(np.int64)).to(dev)
            cnt_out = torch.from_numpy(counts.astype(np.int64)).to(dev)
            cnt_in = torch.empty_like(cnt_out)
            dist.all_to_all_single(cnt_in, cnt_out)
            n_in = cnt_in.cpu().tolist()
            n_srv = int(sum(n_in))
            assert n_srv <= 2 * BATCH, "request skew beyond the serve buffer"
            keys_in = torch.empty((n_srv, 2), dtype=torch.int64, device=dev)
            dist.all_to_all_single(keys_in, keys, output_split_sizes=n_in, input_split_sizes=counts.tolist())
            # owner side: one local batched get straight into the send buffer
            rq = torch.zeros((n_srv, 5), dtype=torch.int64, device=dev)
            rq[:, 1], rq[:, 2] = keys_in[:, 0], keys_in[:, 1]
            rq[:, 3] = PAGE << 32                                   # page_off = 0, nbytes = PAGE
            rq[:, 4] = sendbuf.data_ptr() + torch.arange(n_srv, device=dev, dtype=torch.int64) * PAGE
            for lo in range(0, n_srv, BATCH):
                m = min(BATCH, n_srv - lo)
                capi.check(L.aps_get_batch_dev(h, m, rq.data_ptr() + lo * 40, srv_res.data_ptr() + lo * 8, sp), "serve")
            dist.all_to_all_single(recvbuf[:BATCH * PAGE], sendbuf[:n_srv * PAGE],
                                   output_split_sizes=[int(c) * PAGE for c in counts], input_split_sizes=[c * PAGE for c in n_in])
            return order

        order = step_nccl(0)
        torch.cuda.synchronize(dev)
        p0 = int(picks[0, order[3]])
        assert np.array_equal(recvbuf[3 * PAGE:4 * PAGE].cpu().numpy(), orc.fill_page(DATA_SEED, p0 // 1024, p0 % 1024, PAGE))
        dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for k in range(nccl_steps):
            step_nccl(1 + k)
        torch.cuda.synchronize(dev)
        dist.barrier()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        aux["nccl_all_to_all_GBps"] = round(world * nccl_steps * BATCH * PAGE / float(t.item()) / 1e9, 1)
        aux["nccl_note"] = "same multi-get via torch.distributed all_to_all_single (keys, then pages served by aps_get_batch_dev)"

    clocks = sampler.stop()
    stats = capi.aps_stats_t()
    capi.check(L.aps_stats(h, C.byref(stats)), "stats")
    assert stats.crc_mismatches == 0
    if world > 1:
        t = torch.tensor([remote_frac], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        remote_frac = float(t.item()) / world
        dist.barrier()  # peers may still be reading this GPU's slab
    capi.check(L.aps_close(h), "close")

    # ---- the same three paths with the checksum off (BASELINE configs[1]: "each with CRC off/on"): a second
    # store opened without APS_FLAG_CRC over the same resident set, after the first one released its HBM
    if world == 1 and with_crc and not quick:
        cfg.flags = cfg.flags & capi.APS_FLAG_STABLE_REQS
        h2 = C.c_void_p()
        capi.check(L.aps_open(C.byref(cfg), C.byref(h2)), "aps_open (no CRC)")
        populate(h2)
        n2 = max(1, min(K, 200))
        for k in range(5):
            capi.check(L.aps_get_batch_dev(h2, BATCH, d_reqs.data_ptr() + (k % pool) * req_bytes, d_res.data_ptr(), sp), "get")
        torch.cuda.synchronize(dev)
        ev0.record(stream)
        for k in range(n2):
            capi.check(L.aps_get_batch_dev(h2, BATCH, d_reqs.data_ptr() + ((5 + k) % pool) * req_bytes, d_res.data_ptr(), sp), "get")
        ev1.record(stream)
        torch.cuda.synchronize(dev)
        aux["get_nocrc_GBps"] = round(n2 * BATCH * PAGE / (ev0.elapsed_time(ev1) * 1e-3) / 1e9, 1)
        p = int(picks[(5 + n2 - 1) % pool, 9])
        assert np.array_equal(dst[9 * PAGE:10 * PAGE].cpu().numpy(), orc.fill_page(DATA_SEED, p // 1024, p % 1024, PAGE))
        for _ in range(3):
            capi.check(L.aps_put_batch_dev(h2, BATCH, d_put.data_ptr(), d_res.data_ptr(), sp), "put")
        torch.cuda.synchronize(dev)
        ev0.record(stream)
        for _ in range(aux_steps):
            capi.check(L.aps_put_batch_dev(h2, BATCH, d_put.data_ptr(), d_res.data_ptr(), sp), "put")
        ev1.record(stream)
        torch.cuda.synchronize(dev)
        aux["put_nocrc_GBps"] = round(aux_steps * BATCH * PAGE / (ev0.elapsed_time(ev1) * 1e-3) / 1e9, 1)
        n3 = max(1, min(K, 30))
        for k in range(2):
            capi.check(L.aps_get_batch(h2, BATCH, h_reqs[k % pool].ctypes.data, h_res.ctypes.data, sp), "get_batch")
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for k in range(n3):
            capi.check(L.aps_get_batch(h2, BATCH, h_reqs[(2 + k) % pool].ctypes.data, h_res.ctypes.data, sp), "get_batch")
        torch.cuda.synchronize(dev)
        aux["e2e_get_nocrc_GBps"] = round(n3 * BATCH * PAGE / (time.perf_counter() - t0) / 1e9, 2)
        capi.check(L.aps_close(h2), "close")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- CPU baseline on a bounded sample (rank 0, N = 1 only)
    cpu = None
    if world == 1 and not args.no_cpu_baseline and not quick:
        orc.build()
        threads = os.cpu_count() or 1
        gbps_c, _, desc = cpu_reference_run(steps=int(os.environ.get("APS_CPU_STEPS", "40")), warmup=2, threads=threads)
        cpu = {"value": round(gbps_c, 3), "unit": UNIT, "cores": threads, "kind": "port", "sample": desc}
        # SURVEY.md 8d asks for one thread as well, and for the CPU CRC32C rate (not reference-derived: the reference
        # has no CRC) so that the fused checksum has something on the host to stand beside
        g1, _, _ = cpu_reference_run(steps=4, warmup=1, threads=1)
        cpu["one_thread_value"] = round(g1, 3)
        blob = orc.fill_page(DATA_SEED, 1, 1, 64 << 20)
        rates = {}
        for impl in ("sb8", "hw"):
            try:
                t0 = time.perf_counter()
                orc.crc32c(blob, impl)
                rates[impl] = round(blob.nbytes / (time.perf_counter() - t0) / 1e9, 2)
            except Exception:
                pass
        cpu["crc32c_one_thread_GBps"] = rates
        cpu["crc32c_note"] = "slicing-by-8 / SSE4.2 CRC32C of 64 MiB on one host core; not part of the reference path"

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get("copy_kernel_crc_dram_bytes_per_launch")
    step_bytes = BATCH * PAGE
    # One step is ONE launch of the fused kernel, so the kernel's average launch duration over the timed
    # region is ms_total / K (CUDA events on the launching stream; launch gaps included: conservative).
    achieved = 2 * step_bytes / (ms_total / K * 1e-3) / 1e9
    value = world * K * step_bytes / (ms_total * 1e-3) / 1e9
    if world > 1:
        # cross-shard multi-get: the bytes that cross NVLink per GPU per step bound the step
        nv_peak = 770.0  # measured peer-copy bandwidth per direction per GPU on this pool (B200_PROFILING.md); 900 nominal
        nv_achieved = remote_frac * step_bytes / (ms_total / K * 1e-3) / 1e9
        roofline = {"bound": "nvlink", "achieved": round(nv_achieved, 1), "peak": nv_peak, "unit": "GB/s",
                    "frac": round(nv_achieved / nv_peak, 4), "traffic": None,
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                    "kernel": "aps::copy_kernel<crc,512,4,get> reading peer slabs over NVLink (lookup + gather + CRC, one launch per step)",
                    "algorithmic_bytes_per_launch": int(remote_frac * step_bytes),
                    "remote_fraction": round(remote_frac, 4), "avg_launch_ms": round(ms_total / K, 5),
                    "launches_timed": int(launches), "frac_of_nominal_900": round(nv_achieved / 900.0, 4),
                    "all_to_all_pull_ceiling_GBps": 630.0,  # tools/micro/p2p_all_bw.cu, every GPU reading all its peers
                    "frac_of_pull_ceiling": round(nv_achieved / 630.0, 4),
                    "hbm_traffic_GBps_per_gpu": round(achieved, 1)}
    else:
        roofline = None
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms_total / K, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(resident) if world == 1 else
                   ("hash-sharded cache over %d B200: 1 MiB pages, %d GiB per GPU (%d GiB aggregate), every GPU issues batched "
                    "gets of %d uniform ids over the whole id space, served from the owning GPU over NVLink P2P with fused "
                    "CRC32C (BASELINE configs[4])" % (world, resident * PAGE >> 30, world * resident * PAGE >> 30, BATCH)),
                   "page_bytes": PAGE, "resident_pages": resident,
                   "batch_pages": BATCH, "populate_s": round(t_pop, 2),
                   "l2": "inputs larger than L2: each step reads 256 MiB drawn from the resident set and writes 256 MiB",
                   "shards": "one store per GPU; owner = hash(PageId) mod n; gets read the owner's table and slab in place "
                             "(P2P), no data-path collective" if world > 1 else "single store"},
        "roofline": roofline if roofline is not None else {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                     "kernel": "aps::copy_kernel<crc, threads, group, get> (K1 lookup + K2 gather + K3 fused CRC32C, one launch per step)",
                     "algorithmic_bytes_per_launch": 2 * step_bytes,
                     "avg_launch_ms": round(ms_total / K, 5), "launches_timed": int(launches),
                     "frac_of_nominal_8000": round(achieved / 8000.0, 4)},
        "e2e": {"value": round(world * e2e_steps * step_bytes / e2e_s / 1e9, 2), "unit": UNIT,
                "h2d_bytes_per_step": BATCH * GET_REQ.itemsize, "d2h_bytes_per_step": step_bytes + BATCH * RESULT.itemsize,
                "steps": e2e_steps, "path": "aps_get_batch: host request arrays -> pinned host destinations (PCIe-bound; gathered in HBM by the fused "
                        "kernel, then moved by the copy engine)",
                "pcie_memcpy_d2h_GBps": None if pcie_gbps is None else round(pcie_gbps, 2),
                "frac_of_pcie_memcpy": None if pcie_gbps is None else round(e2e_steps * step_bytes / e2e_s / 1e9 / pcie_gbps, 4)},
        "gpu_launches": int(launches), "clocks": clocks, "aux": aux,
        "copy_variant": os.environ.get("APS_COPY_VARIANT", "default") + ("" if with_crc else " NOCRC"),
        "table": {"lookups": int(stats.lookups), "probe_buckets": int(stats.probe_buckets)},
        "slow_sms": int(stats.slow_sms),
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
