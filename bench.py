#!/usr/bin/env python
"""Benchmark of the page-cache hot path (BASELINE.json configs[1]; configs[4] when --gpus > 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One step = one batched get of 256 x 1 MiB pages, drawn uniformly from the HBM-resident set (64 GiB on one GPU,
128 GiB per GPU when sharded), with the CRC32C integrity check fused into the copy.  `value` times the
device-resident form (requests and destinations already in HBM) with CUDA events; `e2e` times the host-facing C-ABI
call aps_get_batch() with host request arrays and PINNED HOST destinations, so every page crosses PCIe inside the
timed region.  N > 1: one process per GPU (torchrun), one shard per GPU, every GPU asks for ids from the whole id
space (weak scaling, (N-1)/N of the bytes cross NVLink).  `--impl reference` times the CPU restatement of the
reference's LocalCacheManager + LocalPageStore (BASELINE configs[0]: 8 GiB = 64 files x 128 pages, uniform gets,
mt19937_64 seed 1) on the host cores, on tmpfs and on local disk, with all host threads and with one.
"""
import argparse
import ctypes as C
import json
import os
import shutil
import statistics
import subprocess
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
N_DST = 4  # destination buffers the timed steps rotate through (4 x 256 MiB = 1 GiB, far beyond the 126 MB L2);
           # with the push transport: as many as multi-gets are kept in flight, if that is more

# BASELINE configs[0] (SURVEY.md 8d cfg 1): the reference's own CPU-runnable case
CFG0_FILES, CFG0_PAGES_PER_FILE, CFG0_GETS, CFG0_SEED = 64, 128, 65536, 1


def config_block(world, resident, t_pop=None, transport=None):
    """The `config` object; both arms print the same keys."""
    if world == 1:
        workload = ("GpuPageStore on 1 B200: 1 MiB pages, %d GiB resident in HBM, uniform random batched get of %d pages "
                    "with fused CRC32C (BASELINE configs[1])" % (resident * PAGE >> 30, BATCH))
    else:
        workload = ("hash-sharded cache over %d B200: 1 MiB pages, %d GiB per GPU (%d GiB aggregate), every GPU issues batched "
                    "gets of %d uniform ids over the whole id space, served by the owning GPU over NVLink with fused CRC32C "
                    "(BASELINE configs[4])" % (world, resident * PAGE >> 30, world * resident * PAGE >> 30, BATCH))
    return {"workload": workload, "page_bytes": PAGE, "resident_pages": resident, "batch_pages": BATCH,
            "populate_s": t_pop,
            "l2": "inputs larger than L2: each step reads 256 MiB drawn at random from the resident set and writes 256 MiB into "
                  "one of %d destination buffers used in rotation (%d MiB in all)" % (N_DST, N_DST * BATCH),
            "request_pool": "one request batch per step (warm-up + timed), none reused",
            "shards": "single store" if world == 1 else
                      "one store per GPU; owner = hash(PageId) mod n; transport: %s" % (transport or "see roofline.kernel"),
            "reference_arm": "BASELINE configs[0]: LocalCacheManager + LocalPageStore port, %d files x %d pages of 1 MiB (8 GiB), "
                             "uniform whole-page gets, mt19937_64 seed %d, steps of %d gets"
                             % (CFG0_FILES, CFG0_PAGES_PER_FILE, CFG0_SEED, BATCH)}


# ------------------------------------------------------------------------------------------------ CPU legs

def fs_type(path):
    best, kind = "", "?"
    try:
        real = os.path.realpath(path)
        for line in open("/proc/mounts"):
            parts = line.split()
            if len(parts) >= 3 and (real == parts[1] or real.startswith(parts[1].rstrip("/") + "/")) and len(parts[1]) > len(best):
                best, kind = parts[1], parts[2]
    except OSError:
        pass
    return kind


def cpu_dirs():
    """(label, directory, fs) candidates: tmpfs, and the box's local disk (the first non-tmpfs writable place)."""
    out, seen = [], set()
    for base in ("/dev/shm", tempfile.gettempdir(), ROOT, os.path.expanduser("~")):
        try:
            kind = fs_type(base)
            label = "tmpfs" if kind in ("tmpfs", "ramfs") else "disk"
            if label in seen or not os.access(base, os.W_OK):
                continue
            seen.add(label)
            out.append((label, base, kind))
        except OSError:
            pass
    return out


def cpu_config0(base, threads_list, gets=CFG0_GETS, step_gets=None, steps=0, warmup=0):
    """BASELINE configs[0] on `base`: populate 64 x 128 pages of 1 MiB through the oracle's LocalPageStore, then uniform
    whole-page gets through LocalCacheManager.get.  Returns {"threads": {t: GB/s over `gets` gets}, "steps": (GB/s, ms)}.
    A smaller file set is used (and named) only when `base` lacks room."""
    from oracle import binding as orc
    n_files, ppf = CFG0_FILES, CFG0_PAGES_PER_FILE
    free = shutil.disk_usage(base).free
    while n_files * ppf * PAGE + (1 << 30) > free and ppf > 8:
        ppf //= 2
    root = tempfile.mkdtemp(prefix="aps_cfg0_", dir=base)
    try:
        t0 = time.time()
        orc.bench_local_get(root, PAGE, n_files, ppf, True, 64, DATA_SEED, CFG0_SEED, max(threads_list))  # populate + touch
        out = {"file_set_GiB": n_files * ppf * PAGE / 2**30, "files": n_files, "pages_per_file": ppf,
               "populate_s": round(time.time() - t0, 1), "threads": {}}
        for t in threads_list:
            n = gets if t > 1 else max(BATCH, gets // 8)  # one thread: an eighth of the gets keeps the run bounded
            s, b = orc.bench_local_get(root, PAGE, n_files, ppf, False, n, DATA_SEED, CFG0_SEED, t)
            out["threads"][t] = {"GBps": round(b / s / 1e9, 3), "gets": n}
        if steps:
            t = max(threads_list)
            for w in range(warmup):
                orc.bench_local_get(root, PAGE, n_files, ppf, False, step_gets, DATA_SEED, CFG0_SEED + 100000 + w, t)
            secs = nbytes = 0
            for k in range(steps):  # thread j of step k draws from mt19937_64(seed 1 + k * threads + j): disjoint streams
                s, b = orc.bench_local_get(root, PAGE, n_files, ppf, False, step_gets, DATA_SEED, CFG0_SEED + k * t, t)
                secs += s
                nbytes += b
            out["steps"] = (nbytes / secs / 1e9, secs / steps * 1e3)
        return out
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import binding as orc
    orc.build()
    threads = os.cpu_count() or 1
    steps, warmup = max(1, min(args.steps, 256)), min(args.warmup, 5)
    dirs = cpu_dirs()
    results = {}
    head = None  # the timed steps run on the first directory that works (tmpfs when the box has one)
    for label, base, kind in dirs:
        try:
            results[label] = cpu_config0(base, [threads, 1], step_gets=BATCH, steps=0 if head else steps, warmup=warmup)
            results[label]["dir"], results[label]["fs"] = base, kind
            head = head or results[label]
        except Exception as e:  # noqa: BLE001
            results[label] = {"unavailable": repr(e)[:200], "dir": base, "fs": kind}
    if head is None:
        raise SystemExit("reference arm: no writable directory could hold the file set: %r" % (results,))
    gbps, ms = head["steps"]
    desc = ("%d steps of %d uniform random whole-page gets over BASELINE configs[0]'s file set (%d files x %d pages of 1 MiB = "
            "%.0f GiB under %s [%s], resident in the OS page cache: the reference's best case), %d threads"
            % (steps, BATCH, head["files"], head["pages_per_file"], head["file_set_GiB"], head["dir"], head["fs"], threads))
    full = {label: ({"dir": r["dir"], "fs": r["fs"], "file_set_GiB": r["file_set_GiB"], "populate_s": r["populate_s"],
                     "all_threads_GBps": r["threads"][threads]["GBps"], "all_threads_gets": r["threads"][threads]["gets"],
                     "one_thread_GBps": r["threads"][1]["GBps"], "one_thread_gets": r["threads"][1]["gets"]}
                    if "threads" in r else r) for label, r in results.items()}
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbps, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_block(args.gpus, 131072 if args.gpus > 1 else 65536),
        "cpu_baseline": {"value": round(gbps, 3), "unit": UNIT, "cores": threads, "kind": "port", "sample": desc,
                         "configs0_65536_gets_seed1": full},
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

def run_tool(name, argv, timeout=900):
    """Runs tools/<name> in its own process (after the main store released its HBM) and returns its JSON line."""
    try:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", name)] + argv, capture_output=True, timeout=timeout)
        for line in reversed(out.stdout.decode().splitlines()):
            if line.startswith("{"):
                return json.loads(line)
        return {"unavailable": (out.stderr.decode() or "no output")[-300:]}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)[:300]}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from alluxio_b200 import capi
    from alluxio_b200 import build as build_mod
    from alluxio_b200 import sharding

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
    quick = os.environ.get("APS_BENCH_QUICK") == "1"  # kernel sweeps: skip the slow PCIe, CPU and auxiliary legs
    # several GPUs: "push" = the owners store the pages into the requester's landing buffer (aps_get_batch_push_*),
    # "pull" = the requester's kernel reads the owners' slabs in place (aps_get_batch_sharded_dev)
    transport = os.environ.get("APS_BENCH_TRANSPORT", "push") if world > 1 else "local"
    depth = int(os.environ.get("APS_BENCH_DEPTH", "6")) if transport == "push" else 1
    global N_DST
    N_DST = max(N_DST, depth)

    # pages per GPU: configs[1] = 64 GiB on one GPU; configs[4] = 128 GiB per GPU (1 TiB at 8)
    resident = int(os.environ.get("APS_BENCH_PAGES", "131072" if world > 1 else "65536"))
    free_b, _ = torch.cuda.mem_get_info(dev)
    while resident * PAGE + (10 << 30) > free_b and resident > 1024:
        resident //= 2  # a smaller GPU slice: stated in config.workload
    if world > 1:
        t = torch.tensor([resident], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        resident = int(t.item())
    # Global id space: world * resident pages; page g is (file g // 1024, index g % 1024).  With several
    # GPUs the cache is sharded: GPU aps_shard_of(key) owns the page (hash routing).
    total_pages = world * resident
    gids = np.arange(total_pages, dtype=np.uint64)
    if world > 1:
        owned = gids[sharding.owned_by(rank, world, gids // np.uint64(1024), gids % np.uint64(1024))]
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
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    assert sp.value, "need a real stream handle"

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from gpu_util import GET_REQ, KEY, PUT_REQ, RESULT, raw_tensor
    from oracle import binding as orc

    def ids_to_dev(ids):
        return (torch.from_numpy((ids // np.uint64(1024)).view(np.int64)).to(dev),
                torch.from_numpy((ids % np.uint64(1024)).view(np.int64)).to(dev))

    def put_reqs_for(ids, src_ptr):
        r = np.zeros(len(ids), PUT_REQ)
        r["file_lo"], r["page_index"], r["len"] = ids // np.uint64(1024), ids % np.uint64(1024), PAGE
        r["src"] = src_ptr + np.arange(len(ids), dtype=np.uint64) * PAGE
        return torch.from_numpy(r.view(np.uint8).reshape(-1)).to(dev)

    # ---- populate the pages this GPU owns: generate on the device, put 256 at a time
    stage = torch.empty(BATCH * PAGE, dtype=torch.uint8, device=dev)
    d_res = torch.zeros(BATCH * 8, dtype=torch.uint8, device=dev)

    def fill_stage(ids):
        d_files, d_idx = ids_to_dev(ids)
        capi.check(L.aps_fill_pages_dev(local_rank, stage.data_ptr(), PAGE, PAGE, len(ids), d_files.data_ptr(),
                                        d_idx.data_ptr(), DATA_SEED, sp), "fill")

    def populate(store):
        first = None
        for b0 in range(0, len(owned), BATCH):
            ids = owned[b0:b0 + BATCH]
            fill_stage(ids)
            d_reqs = put_reqs_for(ids, stage.data_ptr())
            capi.check(L.aps_put_batch_dev(store, len(ids), d_reqs.data_ptr(), d_res.data_ptr(), sp), "put")
            if b0 == 0:
                torch.cuda.synchronize(dev)
                first = d_res.cpu().numpy().view(RESULT).copy()
        torch.cuda.synchronize(dev)
        return first

    t_pop = time.time()
    first_put_res = populate(h)
    t_pop = time.time() - t_pop
    for i in (0, 1, 255):  # put CRCs against the oracle (the checker; never the thing measured)
        g = int(owned[i])
        want = orc.crc32c(orc.fill_page(DATA_SEED, g // 1024, g % 1024, PAGE))
        assert (not with_crc or int(first_put_res[i]["crc"]) == want) and int(first_put_res[i]["len"]) == PAGE, \
            "put CRC differs from oracle"

    # ---- destination buffers: N_DST x 256 MiB used in rotation.  With the push transport they are the store's landing
    # buffer (memory the peers can write); otherwise a plain device allocation.
    dst_bytes = N_DST * BATCH * PAGE
    if world > 1:
        # map every peer's table and slab (pull) and its landing buffer and mailboxes (push); CUDA IPC handles travel
        # over torch.distributed
        hnd = capi.aps_shard_handle()
        capi.check(L.aps_shard_export(h, C.byref(hnd)), "shard_export")
        blobs = [None] * world
        dist.all_gather_object(blobs, bytes(hnd.opaque))
        arr = (capi.aps_shard_handle * world)()
        for i, bl in enumerate(blobs):
            C.memmove(C.addressof(arr[i]), bl, 256)
        capi.check(L.aps_shard_attach(h, world, rank, arr), "shard_attach")

    def all_ranks_ok(ok):
        t = torch.tensor([1 if ok else 0], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    dst = None
    if transport == "push":
        # set-up is collective; if it fails on any rank (e.g. no room for the landing buffer) every rank falls back to the
        # pull form, and the line says which transport ran
        ph = capi.aps_push_handle()
        try:
            capi.check(L.aps_push_setup(h, dst_bytes, depth), "push_setup")
            capi.check(L.aps_push_export(h, C.byref(ph)), "push_export")
            ok = True
        except Exception as e:  # noqa: BLE001
            print("bench.py: push transport unavailable on rank %d: %r" % (rank, e), file=sys.stderr)
            ok = False
        if all_ranks_ok(ok):
            blobs = [None] * world
            dist.all_gather_object(blobs, bytes(ph.opaque))
            parr = (capi.aps_push_handle * world)()
            for i, bl in enumerate(blobs):
                C.memmove(C.addressof(parr[i]), bl, 192)
            try:
                capi.check(L.aps_push_attach(h, world, rank, parr), "push_attach")
                ok = True
            except Exception as e:  # noqa: BLE001
                print("bench.py: push attach failed on rank %d: %r" % (rank, e), file=sys.stderr)
                ok = False
            ok = all_ranks_ok(ok)
        else:
            ok = False
        if ok:
            nb = C.c_uint64()
            dst = raw_tensor(L.aps_push_landing(h, C.byref(nb)), dst_bytes, dev)
        else:
            transport, depth = "pull", 1
    if dst is None:
        dst = torch.empty(dst_bytes, dtype=torch.uint8, device=dev)
    if world > 1:
        dist.barrier()

    # ---- request pool: one batch per step (warm-up + timed), uniform random page ids over the whole (global) id space,
    # seed 2 + rank; batch p writes into destination buffer p % N_DST
    rng = np.random.default_rng(2 + rank)
    pool = W + K
    picks = rng.integers(0, total_pages, size=(pool, BATCH)).astype(np.uint64)
    reqs = np.zeros((pool, BATCH), GET_REQ)
    reqs["file_lo"], reqs["page_index"] = picks // np.uint64(1024), picks % np.uint64(1024)
    reqs["page_off"], reqs["nbytes"] = 0, PAGE
    reqs["dst"] = (dst.data_ptr() + (np.arange(pool, dtype=np.uint64)[:, None] % np.uint64(N_DST)) * np.uint64(BATCH * PAGE)
                   + np.arange(BATCH, dtype=np.uint64)[None, :] * np.uint64(PAGE))
    d_reqs = torch.from_numpy(reqs.view(np.uint8).reshape(-1)).to(dev)
    d_res_all = torch.zeros(pool * BATCH * 8, dtype=torch.uint8, device=dev)  # every step keeps its own results
    req_bytes, res_bytes = BATCH * GET_REQ.itemsize, BATCH * RESULT.itemsize
    if world > 1:
        pick_owner = sharding.shard_of_keys(np.zeros(picks.size, np.uint64), (picks // np.uint64(1024)).reshape(-1),
                                            (picks % np.uint64(1024)).reshape(-1), world)
        remote_frac = float((pick_owner != rank).mean())
    else:
        remote_frac = 0.0
    get_host = L.aps_get_batch_sharded if world > 1 else L.aps_get_batch

    def reqs_at(p):
        return d_reqs.data_ptr() + p * req_bytes

    def res_at(p):
        return d_res_all.data_ptr() + p * res_bytes

    def run_steps(p0, n):
        """Steps p0 .. p0+n-1 of the pool, back to back on `stream`."""
        if transport == "push":
            # `depth` multi-gets in flight: begin(p + depth - 1) is issued before end(p)
            for p in range(p0, min(p0 + depth - 1, p0 + n)):
                capi.check(L.aps_get_batch_push_begin_dev(h, BATCH, reqs_at(p), sp), "push begin")
            for p in range(p0, p0 + n):
                if p + depth - 1 < p0 + n:
                    capi.check(L.aps_get_batch_push_begin_dev(h, BATCH, reqs_at(p + depth - 1), sp), "push begin")
                capi.check(L.aps_get_batch_push_end_dev(h, res_at(p), sp), "push end")
        elif transport == "pull":
            for p in range(p0, p0 + n):
                capi.check(L.aps_get_batch_sharded_dev(h, BATCH, reqs_at(p), res_at(p), sp), "sharded get")
        else:
            for p in range(p0, p0 + n):
                capi.check(L.aps_get_batch_dev(h, BATCH, reqs_at(p), res_at(p), sp), "get")

    # The request pool above was built on the CPU while the GPU sat idle; W warm-up steps are half a millisecond of work.
    # Keep the GPU busy for ~50 ms first (synthetic-page fills into the staging buffer: not a step, nothing of the
    # workload) so that the warm-up and the timed steps run at steady clocks whatever K is.
    sampler = ClockSampler(local_rank, getattr(torch.cuda.get_device_properties(dev), "uuid", None))  # NVML set-up: before, not between
    warm_ids = owned[:BATCH]
    t_w = time.time()
    while time.time() - t_w < 0.05:
        for _ in range(20):
            fill_stage(warm_ids)
        torch.cuda.synchronize(dev)
    run_steps(0, W)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    # ---- parity of one FULL step (the last warm-up step): its 256 pages are set aside on the device now and compared with the
    # oracle right after the timed region (host copies and a second of CPU checking here would let the GPU clock down)
    p_chk = W - 1
    d_got_chk = dst[(p_chk % N_DST) * BATCH * PAGE:(p_chk % N_DST + 1) * BATCH * PAGE].clone()  # device copy now, host copy later

    def check_full_step():
        res_chk = d_res_all[p_chk * res_bytes:(p_chk + 1) * res_bytes].cpu().numpy().view(RESULT)
        got_chk = d_got_chk.cpu().numpy()
        for i in range(BATCH):
            p = int(picks[p_chk, i])
            page = orc.fill_page(DATA_SEED, p // 1024, p % 1024, PAGE)
            assert int(res_chk[i]["len"]) == PAGE, "get %d returned %d" % (i, int(res_chk[i]["len"]))
            assert not with_crc or int(res_chk[i]["crc"]) == orc.crc32c(page), "get CRC differs from oracle (request %d)" % i
            assert np.array_equal(got_chk[i * PAGE:(i + 1) * PAGE], page), "get bytes differ from oracle (request %d)" % i
        return BATCH

    sampler.start()

    # ---- timed region: device-resident batched get with fused CRC (value + roofline)
    launches0 = L.aps_total_kernel_launches()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    serve_tl = transport == "push" and os.environ.get("APS_BENCH_SERVE_TIMELINE") == "1"  # diagnostic, perturbs nothing
    if serve_tl:
        capi.check(L.aps_debug_timeline(h, 1, None, 0))
    ev0.record(stream)
    run_steps(W, K)
    ev1.record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    ms_total = ev0.elapsed_time(ev1)
    if serve_tl:
        raw = np.zeros(1024 + 256 * 32 + 256, np.uint64)
        capi.check(L.aps_debug_timeline(h, 1, raw.ctypes.data, raw.size))
        tl = raw[:1024].reshape(256, 4)
        live = (tl[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF)) & (tl[:, 0] != 0) & (tl[:, 3] != 0)
        tl = tl[live].astype(np.int64)
        t0 = tl[:, 1].min()  # the gate of the last round (= the round before it has completed)
        rel = (tl - t0) / 1e3
        wend = raw[1024:1024 + 256 * 32].reshape(256, 32)[live][:, :16].astype(np.int64)
        wrel = (wend - t0) / 1e3
        if rank == 0:
            pct = lambda a: " ".join("%.1f" % x for x in np.percentile(a, [0, 10, 50, 90, 100]))
            print("serve timeline (last round, us relative to its gate; p0 p10 p50 p90 p100): ctas=%d\n  cta resident: %s\n  gate passed: %s\n"
                  "  cta end: %s\n  warp end: %s\n  within-CTA warp end spread (max-min): %s"
                  % (len(rel), pct(rel[:, 0]), pct(rel[:, 1]), pct(rel[:, 3]), pct(wrel), pct(wrel.max(axis=1) - wrel.min(axis=1))), file=sys.stderr)
        capi.check(L.aps_debug_timeline(h, 0, None, 0))
    launches = L.aps_total_kernel_launches() - launches0
    per_step = 3 if transport == "push" else 1  # push: post + serve + wait; otherwise lookup + gather + CRC in ONE launch
    assert launches == per_step * K, "unexpected number of kernel launches per step: %d for %d steps" % (launches, K)
    parity_pages = check_full_step()
    res = d_res_all[W * res_bytes:].cpu().numpy().view(RESULT)
    assert (res["len"] == PAGE).all(), "a timed get did not return a full page"
    if with_crc:  # CRCs of one timed step against the oracle (bytes of a full step were checked above)
        for i in (0, 101, 255):
            p = int(picks[W + K - 1, i])
            assert int(res[(K - 1) * BATCH + i]["crc"]) == orc.crc32c(orc.fill_page(DATA_SEED, p // 1024, p % 1024, PAGE))
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())

    # ---- end to end: host request arrays, pinned host destinations, results read back
    e2e_steps = 2 if quick else max(1, min(K, 100))
    pinned = [torch.empty(BATCH * PAGE, dtype=torch.uint8).pin_memory() for _ in range(2)]
    n_h = e2e_steps + 3
    h_reqs = reqs[:n_h].copy() if n_h <= pool else np.resize(reqs, (n_h, BATCH)).copy()
    for p in range(n_h):
        h_reqs[p]["dst"] = pinned[p % 2].data_ptr() + np.arange(BATCH, dtype=np.uint64) * PAGE
    h_res = np.zeros(BATCH, RESULT)

    def step_host(p):
        capi.check(get_host(h, BATCH, h_reqs[p].ctypes.data, h_res.ctypes.data, sp), "get_batch")

    for p in range(3):
        step_host(p)
    pg = int(picks[2 % pool, 5])
    page = orc.fill_page(DATA_SEED, pg // 1024, pg % 1024, PAGE)
    assert np.array_equal(pinned[0].numpy()[5 * PAGE:6 * PAGE], page)
    assert not with_crc or int(h_res[5]["crc"]) == orc.crc32c(page)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for p in range(e2e_steps):
        step_host(3 + p)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert (h_res["len"] == PAGE).all()
    # what the box allows: a plain copy-engine transfer of the same 256 MiB from HBM to the pinned buffer, started on
    # every rank at the same moment (barrier), so that at N > 1 the GPUs compete for the host side as the e2e steps do
    pcie_gbps = None
    if not quick:
        dsrc = dst[:BATCH * PAGE]
        for _ in range(2):
            pinned[0].copy_(dsrc, non_blocking=True)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        for i in range(8):
            pinned[i % 2].copy_(dsrc, non_blocking=True)
        c1.record(stream)
        torch.cuda.synchronize(dev)
        ms_c = c0.elapsed_time(c1)
        if world > 1:
            t = torch.tensor([ms_c], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_c = float(t.item())
        pcie_gbps = 8 * BATCH * PAGE / (ms_c * 1e-3) / 1e9

    # ---- auxiliary numbers (not the headline): puts of FRESH pages with the fused CRC.  Each timed put follows an
    # untimed delete of the same 256 ids and an untimed refill of the source buffer, so every put inserts (claims a
    # table entry and a slab slot) and writes slots that are not in L2; only the put kernels are inside the event pairs.
    aux = {}
    aux_steps = max(1, min(K, 50, len(owned) // BATCH - 1))

    def fresh_puts(store, n_steps):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_steps)]
        first_crc = None
        for s_ in range(n_steps + 1):  # step 0 is a warm-up
            ids = owned[len(owned) - (s_ + 1) * BATCH:len(owned) - s_ * BATCH]
            keys = np.zeros(BATCH, KEY)
            keys["file_lo"], keys["page_index"] = ids // np.uint64(1024), ids % np.uint64(1024)
            d_keys = torch.from_numpy(keys.view(np.uint8).reshape(-1)).to(dev)
            d_put = put_reqs_for(ids, stage.data_ptr())
            capi.check(L.aps_delete_batch_dev(store, BATCH, d_keys.data_ptr(), d_res.data_ptr(), sp), "delete")
            fill_stage(ids)
            if s_:
                evs[s_ - 1][0].record(stream)
            capi.check(L.aps_put_batch_dev(store, BATCH, d_put.data_ptr(), d_res.data_ptr(), sp), "put")
            if s_:
                evs[s_ - 1][1].record(stream)
            if s_ == n_steps:
                torch.cuda.synchronize(dev)
                r = d_res.cpu().numpy().view(RESULT)
                assert (r["len"] == PAGE).all(), "a fresh put failed"
                first_crc = (int(ids[0]), int(r[0]["crc"]))
        torch.cuda.synchronize(dev)
        return n_steps * BATCH * PAGE / (sum(a.elapsed_time(b) for a, b in evs) * 1e-3) / 1e9, first_crc

    put_gbps, (gid0, crc0) = fresh_puts(h, aux_steps)
    assert not with_crc or crc0 == orc.crc32c(orc.fill_page(DATA_SEED, gid0 // 1024, gid0 % 1024, PAGE)), "put CRC differs from oracle"
    aux["put_fresh_fused_crc32c_GBps" if with_crc else "put_fresh_GBps"] = round(put_gbps, 1)

    clocks = sampler.stop()
    stats = capi.aps_stats_t()
    capi.check(L.aps_stats(h, C.byref(stats)), "stats")
    assert stats.crc_mismatches == 0
    if world > 1:
        t = torch.tensor([remote_frac], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        remote_frac = float(t.item()) / world
        dist.barrier()  # peers may still be reading this GPU's slab
    del dst
    capi.check(L.aps_close(h), "close")

    # ---- the same paths with the checksum off (BASELINE configs[1]: "each with CRC off/on"): a second store opened
    # without APS_FLAG_CRC over the same resident set, after the first one released its HBM
    if world == 1 and with_crc and not quick:
        cfg.flags = cfg.flags & capi.APS_FLAG_STABLE_REQS
        h2 = C.c_void_p()
        capi.check(L.aps_open(C.byref(cfg), C.byref(h2)), "aps_open (no CRC)")
        populate(h2)
        dst2 = torch.empty(dst_bytes, dtype=torch.uint8, device=dev)
        reqs2 = reqs.copy()
        reqs2["dst"] = (dst2.data_ptr() + (np.arange(pool, dtype=np.uint64)[:, None] % np.uint64(N_DST)) * np.uint64(BATCH * PAGE)
                        + np.arange(BATCH, dtype=np.uint64)[None, :] * np.uint64(PAGE))
        d_reqs2 = torch.from_numpy(reqs2.view(np.uint8).reshape(-1)).to(dev)
        n2 = max(1, min(K, 200))
        for p in range(W):
            capi.check(L.aps_get_batch_dev(h2, BATCH, d_reqs2.data_ptr() + p * req_bytes, res_at(p), sp), "get")
        torch.cuda.synchronize(dev)
        ev0.record(stream)
        for p in range(W, W + n2):
            capi.check(L.aps_get_batch_dev(h2, BATCH, d_reqs2.data_ptr() + p * req_bytes, res_at(p), sp), "get")
        ev1.record(stream)
        torch.cuda.synchronize(dev)
        aux["get_nocrc_GBps"] = round(n2 * BATCH * PAGE / (ev0.elapsed_time(ev1) * 1e-3) / 1e9, 1)
        p_last = W + n2 - 1
        pg = int(picks[p_last, 9])
        off = (p_last % N_DST) * BATCH * PAGE + 9 * PAGE
        assert np.array_equal(dst2[off:off + PAGE].cpu().numpy(), orc.fill_page(DATA_SEED, pg // 1024, pg % 1024, PAGE))
        aux["put_fresh_nocrc_GBps"] = round(fresh_puts(h2, aux_steps)[0], 1)
        n3 = max(1, min(K, 30))
        for p in range(2):
            capi.check(L.aps_get_batch(h2, BATCH, h_reqs[p].ctypes.data, h_res.ctypes.data, sp), "get_batch")
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for p in range(n3):
            capi.check(L.aps_get_batch(h2, BATCH, h_reqs[(2 + p) % n_h].ctypes.data, h_res.ctypes.data, sp), "get_batch")
        torch.cuda.synchronize(dev)
        aux["e2e_get_nocrc_GBps"] = round(n3 * BATCH * PAGE / (time.perf_counter() - t0) / 1e9, 2)
        del dst2
        capi.check(L.aps_close(h2), "close")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- the other BASELINE configs, measured by their own tools in separate processes once this one has released the
    # GPU memory (rank 0, N = 1): K1 rates, configs[2] (small-page replay, checked against the oracle), configs[3] (MD5)
    del stage, d_reqs, d_res_all
    torch.cuda.empty_cache()
    aux_level = os.environ.get("APS_BENCH_AUX", "full")
    if world == 1 and not quick and aux_level != "0":
        light = aux_level == "light"
        tb = run_tool("bench_table.py", ["--reps", "4"])
        aux["k1_table"] = {k: tb.get(k) for k in ("lookup_hit_Mops", "lookup_miss_Mops", "insert_churn_Mops", "erase_Mops",
                                                  "buckets_per_hit", "roofline", "checks", "unavailable") if k in tb}
        sm = run_tool("bench_small_pages.py", ["--accesses", "5000000" if light else "50000000"] + ([] if light else ["--check"]))
        aux["configs2_small_pages"] = {k: sm.get(k) for k in ("config", "hit_ratio", "victims", "accesses_per_s_end_to_end",
                                                              "store_Mops_per_s", "store_payload_GBps", "bad_gets", "crc_mismatches",
                                                              "oracle", "unavailable") if k in sm}
        md = run_tool("bench_md5.py", ["--files", "1024", "--check", "64"] if light else [])
        aux["configs3_md5"] = {k: md.get(k) for k in ("config", "value", "unit", "roofline", "cpu_baseline", "e2e", "files_checked",
                                                      "digest_mismatches", "md5sum_sample", "unavailable") if k in md}

    # ---- CPU baseline: BASELINE configs[0] on tmpfs, all host threads and one (rank 0, N = 1 only)
    cpu = None
    if world == 1 and not args.no_cpu_baseline and not quick:
        orc.build()
        threads = os.cpu_count() or 1
        label, base, kind = cpu_dirs()[0]
        c0 = cpu_config0(base, [threads, 1], gets=int(os.environ.get("APS_CPU_GETS", str(CFG0_GETS))))
        cpu = {"value": c0["threads"][threads]["GBps"], "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "%d uniform random whole-page gets (mt19937_64 seed %d) over BASELINE configs[0]'s file set: %d files x %d "
                         "pages of 1 MiB = %.0f GiB under %s [%s], resident in the OS page cache (the reference's best case)"
                         % (c0["threads"][threads]["gets"], CFG0_SEED, c0["files"], c0["pages_per_file"], c0["file_set_GiB"], base, kind),
               "one_thread_value": c0["threads"][1]["GBps"], "one_thread_gets": c0["threads"][1]["gets"]}
        # the CPU CRC32C rate (not reference-derived: the reference has no CRC), so that the fused checksum has something
        # on the host to stand beside
        blob = orc.fill_page(DATA_SEED, 1, 1, 64 << 20)
        rates = {}
        for impl in ("sb8", "hw"):
            try:
                t0 = time.perf_counter()
                orc.crc32c(blob, impl)
                rates[impl] = round(blob.nbytes / (time.perf_counter() - t0) / 1e9, 2)
            except Exception:  # noqa: BLE001
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
    # One step is ONE launch of the dominant kernel (the fused copy kernel; with the push transport the owner's serve
    # kernel), so its average launch duration over the timed region is ms_total / K (CUDA events on the launching
    # stream; launch gaps included: conservative).
    achieved = 2 * step_bytes / (ms_total / K * 1e-3) / 1e9
    value = world * K * step_bytes / (ms_total * 1e-3) / 1e9
    if world > 1:
        # cross-shard multi-get: the bytes that cross NVLink per GPU per step bound the step
        nv_achieved = remote_frac * step_bytes / (ms_total / K * 1e-3) / 1e9
        ceil_name, ceil = ("all_to_all_push_ceiling_GBps", 697.0) if transport == "push" else ("all_to_all_pull_ceiling_GBps", 626.0)
        kern = ("aps::copy_kernel<crc,512,4,get,serve>: each owner looks up, CRC-checks and STORES its pages into the requester's "
                "landing buffer over NVLink (one serve launch per GPU per step, %d multi-gets in flight)" % depth) if transport == "push" \
            else "aps::copy_kernel<crc,512,4,get> reading peer slabs over NVLink (lookup + gather + CRC, one launch per step)"
        roofline = {"bound": "nvlink", "achieved": round(nv_achieved, 1), "peak": 770.0, "unit": "GB/s",
                    "frac": round(nv_achieved / 770.0, 4), "traffic": None,
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                    "kernel": kern, "transport": transport, "pipeline_depth": depth,
                    "algorithmic_bytes_per_launch": int(remote_frac * step_bytes),
                    "remote_fraction": round(remote_frac, 4), "avg_launch_ms": round(ms_total / K, 5),
                    "launches_timed": int(launches), "frac_of_nominal_900": round(nv_achieved / 900.0, 4),
                    ceil_name: ceil,  # tools/micro/p2p_transports.cu, every GPU busy in both directions
                    "frac_of_transport_ceiling": round(nv_achieved / ceil, 4),
                    "hbm_traffic_GBps_per_gpu": round(achieved, 1)}
    else:
        roofline = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                    "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                    "kernel": "aps::copy_kernel<crc, threads, group, get> (K1 lookup + K2 gather + K3 fused CRC32C, one launch per step)",
                    "algorithmic_bytes_per_launch": 2 * step_bytes,
                    "avg_launch_ms": round(ms_total / K, 5), "launches_timed": int(launches),
                    "frac_of_nominal_8000": round(achieved / 8000.0, 4)}
    e2e_gbps = world * e2e_steps * step_bytes / e2e_s / 1e9
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms_total / K, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": config_block(world, resident, round(t_pop, 2), transport),
        "roofline": roofline,
        "parity_pages_checked": parity_pages,
        "e2e": {"value": round(e2e_gbps, 2), "unit": UNIT,
                "h2d_bytes_per_step": BATCH * GET_REQ.itemsize, "d2h_bytes_per_step": step_bytes + BATCH * RESULT.itemsize,
                "steps": e2e_steps,
                "path": ("aps_get_batch" if world == 1 else "aps_get_batch_sharded (requester-side kernel reads the owners' slabs)") +
                        ": host request arrays -> pinned host destinations (PCIe-bound; gathered in HBM by the fused kernel in two "
                        "pieces, the copy engine moving the first while the second is gathered)",
                "host_d2h_ceiling_GBps": None if pcie_gbps is None else round(world * pcie_gbps, 2),
                "host_d2h_ceiling_note": "plain cudaMemcpyAsync of 256 MiB HBM -> pinned host, all ranks at once (barrier), aggregate",
                "frac_of_host_d2h_ceiling": None if pcie_gbps is None else round(e2e_gbps / (world * pcie_gbps), 4)},
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
