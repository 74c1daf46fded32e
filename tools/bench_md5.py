#!/usr/bin/env python
"""BASELINE configs[3]: `alluxio fs checksum` MD5 over 4,096 synthetic files of 64 MiB, one file per lane,
digests bit-exact with the oracle / hashlib.  256 GiB does not fit one B200, so the files are fed in waves
of `wave_mib` per file (generated on the device between waves; only the MD5 kernels are timed).
Prints one JSON line; run on a GPU box:  python tools/bench_md5.py [--files 4096] [--file-mib 64] [--wave-mib 16]"""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np
    import torch
    from alluxio_b200 import capi
    from oracle import binding as orc
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=4096)
    ap.add_argument("--file-mib", type=int, default=64)
    ap.add_argument("--wave-mib", type=int, default=16)
    ap.add_argument("--check", type=int, default=6, help="files verified against hashlib")
    args = ap.parse_args()
    L = capi.lib()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    n, wave = args.files, args.wave_mib << 20
    waves = args.file_mib // args.wave_mib
    seed = 4
    buf = torch.empty(n * wave, dtype=torch.uint8, device=dev)
    ptrs = torch.from_numpy((buf.data_ptr() + np.arange(n, dtype=np.int64) * wave)).to(dev)
    lens = torch.full((n,), wave, dtype=torch.int64, device=dev)
    iv = np.tile(np.array([0x67452301, 0xefcdab89, 0x98badcfe, 0x10325476], dtype=np.uint32), (n, 1))
    state = torch.from_numpy(iv.view(np.int32)).to(dev)
    prefix = torch.zeros(n, dtype=torch.int64, device=dev)
    out = torch.zeros(n * 16, dtype=torch.uint8, device=dev)
    files = torch.arange(n, dtype=torch.int64, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(waves)]
    launches0 = L.aps_total_kernel_launches()
    for w in range(waves):
        idx = torch.full((n,), w, dtype=torch.int64, device=dev)
        capi.check(L.aps_fill_pages_dev(0, buf.data_ptr(), wave, wave, n, files.data_ptr(), idx.data_ptr(), seed, sp), "fill")
        prefix.fill_(w * wave)
        ev[w][0].record(stream)
        capi.check(L.aps_md5_stream_dev(0, n, ptrs.data_ptr(), lens.data_ptr(), state.data_ptr(), prefix.data_ptr(),
                                        int(w == waves - 1), out.data_ptr(), sp), "md5")
        ev[w][1].record(stream)
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    total = n * waves * wave
    digests = out.cpu().numpy().reshape(n, 16)
    # bit-exact check of a sample against hashlib over the same generator (the oracle's)
    t0 = time.time()
    bad = 0
    for f in list(range(min(args.check, n) - 1)) + [n - 1]:
        h = hashlib.md5()
        for w in range(waves):
            h.update(orc.fill_page(seed, f, w, wave).tobytes())
        bad += h.digest() != digests[f].tobytes()
    line = {
        "metric": "fs_checksum_md5_GBps", "value": round(total / (ms * 1e-3) / 1e9, 1), "unit": "GB/s", "n_gpus": 1,
        "config": {"workload": "%d files x %d MiB, one file per lane, %d waves of %d MiB per file (BASELINE configs[3])"
                   % (n, args.file_mib, waves, args.wave_mib)},
        "ms_total": round(ms, 2), "bytes": total, "dtype": "u32", "data": "synthetic",
        "gpu_launches": int(L.aps_total_kernel_launches() - launches0) - waves,
        "digest_mismatches_in_sample": int(bad), "sample_files": args.check, "check_s": round(time.time() - t0, 1),
        "roofline": {"bound": "alu-latency", "note": "serial 64-step chain per 64-byte block per file; %d chains" % n},
    }
    print(json.dumps(line))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
