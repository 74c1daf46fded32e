#!/usr/bin/env python
"""BASELINE configs[3]: `alluxio fs checksum` MD5 over 4,096 synthetic files of 64 MiB, one file per lane,
digests bit-exact with the oracle.  256 GiB does not fit one B200, so the files are fed in waves of `wave_mib`
per file (generated on the device between waves; only the MD5 kernels are timed).

Checked here, in the run that is timed: EVERY digest against the oracle's RFC 1321 MD5 of the same generator's
bytes (host cores, all threads), and a sample of files written to tmpfs and digested by the system's `md5sum`.
Reported beside the rate: the dependent-chain latency bound that holds for this kernel, the oracle's MD5 rate on
the host cores (`cpu_baseline`), and an end-to-end rate from pinned host buffers (`e2e`).
Prints one JSON line; run on a GPU box:  python tools/bench_md5.py [--files 4096] [--file-mib 64] [--wave-mib 16]"""
import argparse
import ctypes as C
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

IV = [0x67452301, 0xefcdab89, 0x98badcfe, 0x10325476]
SEED = 4


def main():
    import numpy as np
    import torch
    from alluxio_b200 import capi
    from oracle import binding as orc
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=4096)
    ap.add_argument("--file-mib", type=int, default=64)
    ap.add_argument("--wave-mib", type=int, default=16)
    ap.add_argument("--check", type=int, default=-1, help="files verified against the oracle (-1 = all)")
    ap.add_argument("--md5sum-files", type=int, default=64, help="files also written to tmpfs and digested by md5sum")
    ap.add_argument("--e2e-files", type=int, default=4096, help="files of the end-to-end leg (pinned host -> digest), 1 MiB each")
    args = ap.parse_args()
    L = capi.lib()
    orc.build()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    n, wave = args.files, args.wave_mib << 20
    waves = args.file_mib // args.wave_mib
    file_bytes = waves * wave
    buf = torch.empty(n * wave, dtype=torch.uint8, device=dev)
    ptrs = torch.from_numpy((buf.data_ptr() + np.arange(n, dtype=np.int64) * wave)).to(dev)
    lens = torch.full((n,), wave, dtype=torch.int64, device=dev)
    state = torch.from_numpy(np.tile(np.array(IV, dtype=np.uint32), (n, 1)).view(np.int32)).to(dev)
    prefix = torch.zeros(n, dtype=torch.int64, device=dev)
    out = torch.zeros(n * 16, dtype=torch.uint8, device=dev)
    files = torch.arange(n, dtype=torch.int64, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(waves)]
    launches0 = L.aps_total_kernel_launches()
    for w in range(waves):
        idx = torch.full((n,), w, dtype=torch.int64, device=dev)
        capi.check(L.aps_fill_pages_dev(0, buf.data_ptr(), wave, wave, n, files.data_ptr(), idx.data_ptr(), SEED, sp), "fill")
        prefix.fill_(w * wave)
        ev[w][0].record(stream)
        capi.check(L.aps_md5_stream_dev(0, n, ptrs.data_ptr(), lens.data_ptr(), state.data_ptr(), prefix.data_ptr(),
                                        int(w == waves - 1), out.data_ptr(), sp), "md5")
        ev[w][1].record(stream)
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    total = n * file_bytes
    digests = out.cpu().numpy().reshape(n, 16)
    md5_launches = int(L.aps_total_kernel_launches() - launches0) - waves
    try:  # the SM clock the kernel runs at: the maximum (no throttling seen on these boxes, see bench.py's clocks line)
        import pynvml
        pynvml.nvmlInit()
        sm_mhz = pynvml.nvmlDeviceGetMaxClockInfo(pynvml.nvmlDeviceGetHandleByIndex(0), pynvml.NVML_CLOCK_SM)
    except Exception:  # noqa: BLE001
        sm_mhz = 1965
    del buf

    # ---- every digest against the oracle (host cores; the generator is the oracle's own restatement)
    threads = os.cpu_count() or 1
    to_check = list(range(n)) if args.check < 0 or args.check >= n else sorted(set(list(range(args.check - 1)) + [n - 1]))

    def file_bytes_of(f):
        return np.concatenate([orc.fill_page(SEED, f, w, wave) for w in range(waves)])

    def one(f):
        data = file_bytes_of(f)
        t0 = time.perf_counter()
        d = orc.md5(data)
        return f, d, time.perf_counter() - t0

    t0 = time.time()
    md5_busy = 0.0
    bad = []
    with ThreadPoolExecutor(threads) as ex:
        for f, d, dt in ex.map(one, to_check):
            md5_busy += dt
            if d != digests[f].tobytes():
                bad.append(f)
    check_s = time.time() - t0
    cpu_rate = len(to_check) * file_bytes / (md5_busy / threads) / 1e9 if md5_busy else None  # all threads busy with MD5 alone

    # ---- a sample digested by the system's md5sum
    sample = None
    k = min(args.md5sum_files, n)
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > k * file_bytes + (1 << 30) else None
    if k and base and shutil.which("md5sum"):
        root = tempfile.mkdtemp(prefix="aps_md5_", dir=base)
        try:
            pick = sorted(set(np.linspace(0, n - 1, k).astype(int).tolist()))
            names = []
            for f in pick:
                path = os.path.join(root, "f%05d" % f)
                file_bytes_of(f).tofile(path)
                names.append(path)
            res = subprocess.run(["md5sum"] + names, capture_output=True, check=True).stdout.decode().splitlines()
            got = {os.path.basename(line.split()[1]): line.split()[0] for line in res}
            mism = [f for f in pick if got.get("f%05d" % f) != digests[f].tobytes().hex()]
            sample = {"files": len(pick), "mismatches": len(mism)}
        finally:
            shutil.rmtree(root, ignore_errors=True)

    # ---- end to end: files in PINNED HOST memory -> digests on the host (H2D copy + kernel + D2H of the digests timed)
    e2e = None
    ke = args.e2e_files
    if ke:
        # as many chains as the main run, 1 MiB per file (a few files would measure one chain's ~0.14 GB/s, not the link)
        esz = 1 << 20
        hbuf = torch.empty(ke * esz, dtype=torch.uint8).pin_memory()
        hnp = hbuf.numpy()
        for f in range(ke):
            hnp[f * esz:(f + 1) * esz] = orc.fill_page(SEED, f, 7, esz)
        dbuf = torch.empty(ke * esz, dtype=torch.uint8, device=dev)
        eptrs = torch.from_numpy((dbuf.data_ptr() + np.arange(ke, dtype=np.int64) * esz)).to(dev)
        elens = torch.full((ke,), esz, dtype=torch.int64, device=dev)
        eout = torch.zeros(ke * 16, dtype=torch.uint8, device=dev)
        hout = torch.zeros(ke * 16, dtype=torch.uint8).pin_memory()

        def once():
            dbuf.copy_(hbuf, non_blocking=True)
            capi.check(L.aps_md5_batch_dev(0, ke, eptrs.data_ptr(), elens.data_ptr(), eout.data_ptr(), sp), "md5")
            hout.copy_(eout, non_blocking=True)
            torch.cuda.synchronize()

        once()
        t0 = time.perf_counter()
        for _ in range(3):
            once()
        dt = (time.perf_counter() - t0) / 3
        ok = all(hout.numpy().reshape(ke, 16)[f].tobytes() == orc.md5(orc.fill_page(SEED, f, 7, esz)) for f in (0, ke // 2, ke - 1))
        e2e = {"value": round(ke * esz / dt / 1e9, 2), "unit": "GB/s", "h2d_bytes_per_step": ke * esz, "d2h_bytes_per_step": ke * 16,
               "files": ke, "file_mib": 1, "digests_equal_oracle": bool(ok),
               "note": "PCIe-bound: the files cross the link once, then one kernel; %d files of 1 MiB in pinned host memory per step" % ke}

    gbps = total / (ms * 1e-3) / 1e9
    # the chain inside a file is serial: 64 steps per 64-byte block, each step 3 dependent integer operations
    cycles_per_step = n * (sm_mhz * 1e6) * (ms * 1e-3) / (total / 64 * 64)
    model_cycles = 12.28  # the bare LOP3 -> IADD3 -> LEA.HI chain of one step on B200, measured (tools/micro/md5_chain_latency.cu)
    line = {
        "metric": "fs_checksum_md5_GBps", "value": round(gbps, 1), "unit": "GB/s", "n_gpus": 1,
        "config": {"workload": "%d files x %d MiB, one file per lane, %d waves of %d MiB per file (BASELINE configs[3])"
                   % (n, args.file_mib, waves, args.wave_mib)},
        "ms_total": round(ms, 2), "bytes": total, "dtype": "u32", "data": "synthetic",
        "gpu_launches": md5_launches,
        "files_checked": len(to_check), "digest_mismatches": len(bad), "check_s": round(check_s, 1),
        "md5sum_sample": sample,
        "roofline": {"bound": "alu-latency", "unit": "GB/s", "achieved": round(gbps, 1),
                     "peak": round(gbps * cycles_per_step / model_cycles, 1), "frac": round(model_cycles / cycles_per_step, 3),
                     "cycles_per_md5_step_measured": round(cycles_per_step, 2), "cycles_per_md5_step_model": model_cycles,
                     "sm_mhz": sm_mhz,
                     "note": "chains x f_clk x 64 B / (64 steps x cycles per step): %d independent chains (files); one file alone "
                             "runs at f_clk / cycles_per_step bytes/s = %.0f MB/s, so the kernel pays only with thousands of files"
                             % (n, (sm_mhz * 1e6) / cycles_per_step / 1e6)},
        "cpu_baseline": {"value": None if cpu_rate is None else round(cpu_rate, 2), "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": "oracle RFC 1321 MD5 of %d files x %d MiB on %d threads (time inside MD5 only)"
                                   % (len(to_check), args.file_mib, threads)},
        "e2e": e2e,
    }
    print(json.dumps(line))
    return 1 if bad or (sample and sample["mismatches"]) else 0


if __name__ == "__main__":
    sys.exit(main())
