/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under alluxio_b200/ may link, import or execute
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may.
 *
 * CPU restatement of the two checksums on the hot path.
 *
 * MD5: the reference computes it in a third-party dependency that is NOT in /root/reference:
 *   shell/src/main/java/alluxio/cli/fs/command/ChecksumCommand.java:83-90 calls
 *   org.apache.commons.codec.digest.DigestUtils.md5Hex(InputStream) (commons-codec 1.13,
 *   pom.xml:270-272), which feeds the stream into JDK MessageDigest("MD5").  The published
 *   algorithm is RFC 1321; it is restated below and pinned by the RFC 1321 A.5 suite and by the
 *   reference's own test vector (bytes 0..9, ChecksumCommandIntegrationTest.java:32-46).
 *
 * CRC32C: the reference has NO CRC code at all.  PARITY UNPINNED by the reference;
 *   the spec is CRC-32C/iSCSI (RFC 3720 B.4: poly 0x1EDC6F41, reflected 0x82F63B78, init and
 *   xorout 0xFFFFFFFF), identical to java.util.zip.CRC32C.  Three independent implementations
 *   live here so they can check one another: bit-at-a-time (the definition), slicing-by-8
 *   (fast enough for GiB-sized inputs) and the x86 SSE4.2 crc32 instruction.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

/* ---------------------------------------------------------------- CRC-32C */

#define CRC32C_POLY_REFLECTED 0x82F63B78u

/* Definition: one bit per step, reflected form.  RFC 3720 B.4. */
uint32_t orc_crc32c_bitwise(const void *data, size_t n) {
  const uint8_t *p = (const uint8_t *)data;
  uint32_t crc = 0xFFFFFFFFu;
  for (size_t i = 0; i < n; i++) {
    crc ^= p[i];
    for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (CRC32C_POLY_REFLECTED & (0u - (crc & 1u)));
  }
  return crc ^ 0xFFFFFFFFu;
}

static uint32_t g_sb8[8][256];
static int g_sb8_ready;

static void sb8_init(void) {
  for (uint32_t i = 0; i < 256; i++) {
    uint32_t c = i;
    for (int k = 0; k < 8; k++) c = (c >> 1) ^ (CRC32C_POLY_REFLECTED & (0u - (c & 1u)));
    g_sb8[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; i++)
    for (int t = 1; t < 8; t++) g_sb8[t][i] = (g_sb8[t - 1][i] >> 8) ^ g_sb8[0][g_sb8[t - 1][i] & 0xFF];
  g_sb8_ready = 1;
}

/* Continue a raw (un-inverted) CRC register over more bytes; slicing-by-8. */
uint32_t orc_crc32c_update_raw(uint32_t crc, const void *data, size_t n) {
  const uint8_t *p = (const uint8_t *)data;
  if (!g_sb8_ready) sb8_init();
  while (n && ((uintptr_t)p & 7)) {
    crc = (crc >> 8) ^ g_sb8[0][(crc ^ *p++) & 0xFF];
    n--;
  }
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    uint32_t lo = (uint32_t)w ^ crc, hi = (uint32_t)(w >> 32);
    crc = g_sb8[7][lo & 0xFF] ^ g_sb8[6][(lo >> 8) & 0xFF] ^ g_sb8[5][(lo >> 16) & 0xFF] ^ g_sb8[4][lo >> 24] ^
          g_sb8[3][hi & 0xFF] ^ g_sb8[2][(hi >> 8) & 0xFF] ^ g_sb8[1][(hi >> 16) & 0xFF] ^ g_sb8[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) crc = (crc >> 8) ^ g_sb8[0][(crc ^ *p++) & 0xFF];
  return crc;
}

uint32_t orc_crc32c(const void *data, size_t n) {
  return orc_crc32c_update_raw(0xFFFFFFFFu, data, n) ^ 0xFFFFFFFFu;
}

/* Third implementation: the CPU's own CRC32C instruction.  Returns 0 and sets *ok = 0 when the
 * build has no SSE4.2. */
uint32_t orc_crc32c_hw(const void *data, size_t n, int *ok) {
#if defined(__SSE4_2__)
  const uint8_t *p = (const uint8_t *)data;
  uint64_t crc = 0xFFFFFFFFu;
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    crc = _mm_crc32_u64(crc, w);
    p += 8;
    n -= 8;
  }
  while (n--) crc = _mm_crc32_u8((uint32_t)crc, *p++);
  if (ok) *ok = 1;
  return (uint32_t)crc ^ 0xFFFFFFFFu;
#else
  (void)data; (void)n;
  if (ok) *ok = 0;
  return 0;
#endif
}

/* crc(A || B) from crc(A), crc(B), len(B): zlib's crc32_combine construction restated for the
 * CRC-32C polynomial (GF(2) matrix squaring).  Used by tests as a size-independent property. */
static uint32_t gf2_times(const uint32_t *mat, uint32_t vec) {
  uint32_t sum = 0;
  while (vec) {
    if (vec & 1) sum ^= *mat;
    vec >>= 1;
    mat++;
  }
  return sum;
}
static void gf2_square(uint32_t *sq, const uint32_t *mat) {
  for (int n = 0; n < 32; n++) sq[n] = gf2_times(mat, mat[n]);
}
uint32_t orc_crc32c_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) {
  uint32_t even[32], odd[32];
  if (len2 == 0) return crc1;
  odd[0] = CRC32C_POLY_REFLECTED;
  uint32_t row = 1;
  for (int n = 1; n < 32; n++) {
    odd[n] = row;
    row <<= 1;
  }
  gf2_square(even, odd);
  gf2_square(odd, even);
  do {
    gf2_square(even, odd);
    if (len2 & 1) crc1 = gf2_times(even, crc1);
    len2 >>= 1;
    if (len2 == 0) break;
    gf2_square(odd, even);
    if (len2 & 1) crc1 = gf2_times(odd, crc1);
    len2 >>= 1;
  } while (len2 != 0);
  return crc1 ^ crc2;
}

/* -------------------------------------------------------------------- MD5 */
/* RFC 1321.  State words a,b,c,d; 64 steps per 64-byte block; little-endian everywhere. */

typedef struct {
  uint32_t h[4];
  uint64_t nbytes;
  uint8_t buf[64];
} orc_md5_ctx;

static const uint32_t MD5_K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
static const uint8_t MD5_S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                                  5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                  6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};

static void md5_block(uint32_t h[4], const uint8_t *blk) {
  uint32_t m[16];
  for (int i = 0; i < 16; i++)
    m[i] = (uint32_t)blk[4 * i] | ((uint32_t)blk[4 * i + 1] << 8) | ((uint32_t)blk[4 * i + 2] << 16) |
           ((uint32_t)blk[4 * i + 3] << 24);
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
  for (int i = 0; i < 64; i++) {
    uint32_t f;
    int g;
    if (i < 16) {
      f = (b & c) | (~b & d);
      g = i;
    } else if (i < 32) {
      f = (d & b) | (~d & c);
      g = (5 * i + 1) & 15;
    } else if (i < 48) {
      f = b ^ c ^ d;
      g = (3 * i + 5) & 15;
    } else {
      f = c ^ (b | ~d);
      g = (7 * i) & 15;
    }
    uint32_t t = a + f + MD5_K[i] + m[g];
    a = d;
    d = c;
    c = b;
    b = b + ((t << MD5_S[i]) | (t >> (32 - MD5_S[i])));
  }
  h[0] += a;
  h[1] += b;
  h[2] += c;
  h[3] += d;
}

void orc_md5_init(orc_md5_ctx *c) {
  c->h[0] = 0x67452301u;
  c->h[1] = 0xefcdab89u;
  c->h[2] = 0x98badcfeu;
  c->h[3] = 0x10325476u;
  c->nbytes = 0;
}

void orc_md5_update(orc_md5_ctx *c, const void *data, size_t n) {
  const uint8_t *p = (const uint8_t *)data;
  size_t fill = (size_t)(c->nbytes & 63);
  c->nbytes += n;
  if (fill) {
    size_t take = 64 - fill < n ? 64 - fill : n;
    memcpy(c->buf + fill, p, take);
    p += take;
    n -= take;
    if (fill + take < 64) return;
    md5_block(c->h, c->buf);
  }
  while (n >= 64) {
    md5_block(c->h, p);
    p += 64;
    n -= 64;
  }
  if (n) memcpy(c->buf, p, n);
}

void orc_md5_final(orc_md5_ctx *c, uint8_t out[16]) {
  uint64_t bits = c->nbytes * 8;
  size_t fill = (size_t)(c->nbytes & 63);
  c->buf[fill++] = 0x80;
  if (fill > 56) {
    memset(c->buf + fill, 0, 64 - fill);
    md5_block(c->h, c->buf);
    fill = 0;
  }
  memset(c->buf + fill, 0, 56 - fill);
  for (int i = 0; i < 8; i++) c->buf[56 + i] = (uint8_t)(bits >> (8 * i));
  md5_block(c->h, c->buf);
  for (int i = 0; i < 4; i++)
    for (int k = 0; k < 4; k++) out[4 * i + k] = (uint8_t)(c->h[i] >> (8 * k));
}

/* One-shot digest, the shape of DigestUtils.md5(InputStream) (ChecksumCommand.java:88). */
void orc_md5(const void *data, size_t n, uint8_t out[16]) {
  orc_md5_ctx c;
  orc_md5_init(&c);
  orc_md5_update(&c, data, n);
  orc_md5_final(&c, out);
}

/* DigestUtils.md5Hex: lower-case hex (commons-codec Hex.encodeHexString default). */
void orc_md5_hex(const void *data, size_t n, char out[33]) {
  static const char hex[] = "0123456789abcdef";
  uint8_t d[16];
  orc_md5(data, n, d);
  for (int i = 0; i < 16; i++) {
    out[2 * i] = hex[d[i] >> 4];
    out[2 * i + 1] = hex[d[i] & 15];
  }
  out[32] = 0;
}
