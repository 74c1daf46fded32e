"""`alluxio fs checksum` for local files, digests computed on the GPU.

Mirrors shell/src/main/java/alluxio/cli/fs/command/ChecksumCommand.java:48-90: a path must be a file (a folder is
an error), the digest is the MD5 of the file's bytes in lower-case hex, and the command prints
"md5sum: <hex>" followed by an empty line.  The reference takes one path (wildcards expand to several); here any
number of paths is digested in ONE kernel launch per wave, one file per lane (aps_md5_batch).

    python -m alluxio_b200.checksum FILE [FILE ...]

Files are read into page-locked host memory in waves of at most `wave_bytes`; the kernel reads them over PCIe in
place.  There is no CPU fallback: without the CUDA library or a GPU this fails.
"""
import ctypes as C
import os
import sys

from . import capi


class FileDoesNotExistException(Exception):
    """alluxio.exception.FileDoesNotExistException with ExceptionMessage.PATH_MUST_BE_FILE."""


def md5_hex_of_files(paths, device=0, wave_bytes=1 << 30):
    """Returns the lower-case hex MD5 of every file, in order."""
    L = capi.lib()
    for p in paths:
        if os.path.isdir(p):
            raise FileDoesNotExistException("Path \"%s\" must be a file." % p)
        if not os.path.isfile(p):
            raise FileDoesNotExistException("Path \"%s\" does not exist." % p)
    sizes = [os.path.getsize(p) for p in paths]
    out = [None] * len(paths)
    i = 0
    while i < len(paths):
        # one wave: as many whole files as fit (always at least one)
        j, total = i, 0
        while j < len(paths) and (j == i or total + sizes[j] <= wave_bytes):
            total += (sizes[j] + 255) & ~255
            j += 1
        n = j - i
        buf = C.c_void_p()
        capi.check(L.aps_host_alloc_pinned(max(total, 256), C.byref(buf)), "aps_host_alloc_pinned")
        try:
            ptrs = (C.c_void_p * n)()
            lens = (C.c_uint64 * n)()
            at = 0
            for k in range(n):
                size = sizes[i + k]
                ptrs[k] = buf.value + at
                lens[k] = size
                if size:
                    dst = (C.c_uint8 * size).from_address(buf.value + at)
                    with open(paths[i + k], "rb") as f:
                        got = f.readinto(dst)
                    if got != size:
                        raise IOError("short read of %s" % paths[i + k])
                at += (size + 255) & ~255
            digests = ((C.c_uint8 * 16) * n)()
            capi.check(L.aps_md5_batch(device, n, ptrs, lens, digests), "aps_md5_batch")
            for k in range(n):
                out[i + k] = bytes(digests[k]).hex()
        finally:
            L.aps_host_free_pinned(buf)
        i = j
    return out


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if not argv:
        print("Usage: checksum <path> [<path> ...]", file=sys.stderr)
        return 2
    try:
        digests = md5_hex_of_files(argv)
    except FileDoesNotExistException as e:
        print(str(e), file=sys.stderr)
        return 1
    for d in digests:
        print("md5sum: " + d + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
