"""`alluxio fs checksum` mirror (alluxio_b200/checksum.py) against ChecksumCommand
(shell/src/main/java/alluxio/cli/fs/command/ChecksumCommand.java:48-90): MD5 hex of the file bytes, "md5sum: <hex>"
plus an empty line, a folder is an error.  The reference's own test uses bytes 0..9
(ChecksumCommandIntegrationTest.java:32-46)."""
import hashlib
import subprocess
import sys

import numpy as np
import pytest


def test_folder_and_missing_path_are_errors(tmp_path):
    from alluxio_b200.checksum import FileDoesNotExistException, md5_hex_of_files
    with pytest.raises(FileDoesNotExistException):
        md5_hex_of_files([str(tmp_path)])
    with pytest.raises(FileDoesNotExistException):
        md5_hex_of_files([str(tmp_path / "absent")])


@pytest.mark.gpu
def test_digests_and_output_format(tmp_path):
    from alluxio_b200.checksum import md5_hex_of_files
    rng = np.random.default_rng(12)
    blobs = [bytes(range(10)), b"", b"a", rng.integers(0, 256, 55, dtype=np.uint8).tobytes(),
             rng.integers(0, 256, 64, dtype=np.uint8).tobytes(), rng.integers(0, 256, 3 * 1024 * 1024 + 17, dtype=np.uint8).tobytes()]
    paths = []
    for k, b in enumerate(blobs):
        p = tmp_path / ("f%d" % k)
        p.write_bytes(b)
        paths.append(str(p))
    want = [hashlib.md5(b).hexdigest() for b in blobs]
    assert want[0] == "c56bd5480f6e5413cb62a0ad9666613a"      # the reference test's vector
    assert md5_hex_of_files(paths) == want
    assert md5_hex_of_files(paths, wave_bytes=1024) == want   # one file per wave
    out = subprocess.run([sys.executable, "-m", "alluxio_b200.checksum", paths[0], paths[5]], capture_output=True, timeout=300,
                         cwd=str(tmp_path.parent), env=dict(__import__("os").environ, PYTHONPATH=__import__("os").path.dirname(
                             __import__("os").path.dirname(__import__("os").path.abspath(__file__)))))
    assert out.returncode == 0
    assert out.stdout.decode() == "md5sum: %s\n\nmd5sum: %s\n\n" % (want[0], want[5])
