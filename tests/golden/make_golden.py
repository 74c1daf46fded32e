#!/usr/bin/env python
"""Writes tests/golden/*.json.  The vectors are literals from RFC 3720 B.4, RFC 1321 A.5 and the
reference's evictor tests; MD5 entries are cross-checked with hashlib before they are written."""
import hashlib
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))

crc = [
    {"name": "check", "hex": b"123456789".hex(), "crc32c": "e3069283"},
    {"name": "32 x 00 (RFC 3720 B.4)", "hex": (b"\x00" * 32).hex(), "crc32c": "8a9136aa"},
    {"name": "32 x ff (RFC 3720 B.4)", "hex": (b"\xff" * 32).hex(), "crc32c": "62a8ab43"},
    {"name": "0..31 (RFC 3720 B.4)", "hex": bytes(range(32)).hex(), "crc32c": "46dd794e"},
    {"name": "31..0 (RFC 3720 B.4)", "hex": bytes(range(31, -1, -1)).hex(), "crc32c": "113fdb5c"},
    {"name": "empty", "hex": "", "crc32c": "00000000"},
]
md5 = [
    ("", "d41d8cd98f00b204e9800998ecf8427e"), ("a", "0cc175b9c0f1b6a831c399e269772661"),
    ("abc", "900150983cd24fb0d6963f7d28e17f72"), ("message digest", "f96b697d7cb7938d525a2f31aaf161d0"),
    ("abcdefghijklmnopqrstuvwxyz", "c3fcd3d76192e4007dfb496cca67e13b"),
    ("ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789", "d174ab98d277d9f5a5611c2c9f419d9f"),
    ("1234567890" * 8, "57edf4a22be3c955ac49da2e2107b67a"),
]
md5_vectors = [{"name": "RFC 1321 A.5", "hex": s.encode().hex(), "md5": d} for s, d in md5]
md5_vectors.append({"name": "ChecksumCommandIntegrationTest bytes 0..9", "hex": bytes(range(10)).hex(),
                    "md5": "c56bd5480f6e5413cb62a0ad9666613a"})
for v in md5_vectors:
    assert hashlib.md5(bytes.fromhex(v["hex"])).hexdigest() == v["md5"], v["name"]
json.dump({"crc32c": crc, "md5": md5_vectors}, open(os.path.join(HERE, "checksum_kats.json"), "w"), indent=1)

# ops: g = updateOnGet, p = updateOnPut, d = updateOnDelete, e = evict() must return the id (null = None)
F, S, T = ["1L", 2], ["3L", 4], ["5L", 6]
ONE, TWO, THREE, FOUR = ["1L", 1], ["2L", 2], ["3L", 3], ["4L", 4]
lru = {
    "evictGetOrder": [["g", F], ["g", S], ["e", F], ["d", F], ["e", S]],
    "evictUpdatedGetOrder": [["g", F], ["g", S], ["g", T], ["g", F], ["e", S], ["d", S], ["e", T], ["d", T], ["e", F]],
    "evictPutOrder": [["p", F], ["p", S], ["e", F], ["d", F], ["e", S]],
    "evictUpdatedPutOrder": [["p", F], ["p", S], ["p", T], ["p", F], ["e", S], ["d", S], ["e", T], ["d", T], ["e", F]],
    "evictAfterDelete": [["p", F], ["p", S], ["p", T], ["d", S], ["e", F], ["d", F], ["e", T]],
    "evictEmpty": [["e", None]],
    "evictAllGone": [["p", F], ["p", S], ["p", T], ["d", F], ["d", S], ["d", T], ["e", None]],
}
lfu = {
    "evictGetOrder": [["g", ONE], ["g", TWO], ["g", TWO]] + [["g", FOUR]] * 4 +
                     [["e", ONE], ["d", ONE], ["e", TWO], ["d", TWO], ["e", FOUR], ["d", FOUR]],
    "evictGetMixedOrder": [["g", x] for x in (FOUR, TWO, FOUR, TWO, FOUR, ONE, FOUR)] +
                          [["e", ONE], ["d", ONE], ["e", TWO], ["d", TWO], ["e", FOUR], ["d", FOUR]],
    "evictGetRecentOrder": [["g", ONE], ["g", TWO], ["e", ONE], ["d", ONE], ["e", TWO]],
    "evictUpdatedGetRecentOrder": [["g", x] for x in (ONE, ONE, TWO, TWO, THREE, THREE, TWO, THREE, ONE)] +
                                  [["e", TWO], ["d", TWO], ["e", THREE], ["d", THREE], ["e", ONE]],
    "evictDeleteSkippedBucket": [["g", ONE]] + [["g", FOUR]] * 4 + [["e", ONE], ["d", ONE], ["e", FOUR], ["d", FOUR]],
    "evictDeleteTwice": [["g", ONE], ["g", ONE], ["e", ONE], ["d", ONE], ["g", ONE], ["e", ONE]],
    "evictAfterDelete": [["p", x] for x in (ONE, TWO, TWO, THREE, THREE, THREE)] +
                        [["d", TWO], ["e", ONE], ["d", ONE], ["e", THREE]],
    "evictEmpty": [["e", None]],
    "evictAllGone": [["p", x] for x in (ONE, TWO, TWO, THREE, THREE, THREE)] +
                    [["d", ONE], ["d", TWO], ["d", THREE], ["e", None]],
}
json.dump({"LRU": lru, "LFU": lfu}, open(os.path.join(HERE, "evictor_sequences.json"), "w"), indent=1)
print("golden vectors written")
