/* TEST INFRASTRUCTURE ONLY.  Runs integration/jni/aps_jni.c -- the JNI shim under GpuPageStore.java -- without a JVM and
 * without a GPU: the natives are called the way the Java class calls them (integration/java/.../GpuPageStore.java), over
 * the mock JNIEnv of fake_env.c (copying array accessors) and the store double of tests/host_double/fake_aps.cc.  What is
 * under test is the shim's own code: argument unpacking, which array is written back and which is released with
 * JNI_ABORT, every Get paired with a Release, statuses and per-request results reaching the Java arrays, the direct-buffer
 * offsets of the batched calls.  Built with AddressSanitizer + UBSan by tests/test_jni_shim_runs.py.
 * aps_md5_batch is answered by the oracle's RFC 1321 MD5 here (tests/host_double/md5_via_oracle.c; the kernel is tested in
 * tests/test_gpu_md5.py). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "fake_env.h"
#include "../../integration/jni/aps_jni.c" /* the shim itself, compiled into this test */

void orc_md5(const void* data, size_t n, uint8_t out[16]); /* oracle/checksums.c; aps_md5_batch: tests/host_double/md5_via_oracle.c */

static int g_failures = 0;
#define CHECK(cond, ...)                                     \
  do {                                                       \
    if (!(cond)) {                                           \
      fprintf(stderr, "FAILED %s:%d: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);                          \
      fprintf(stderr, "\n");                                 \
      ++g_failures;                                          \
    }                                                        \
  } while (0)
#define BALANCED() CHECK(fake_env_outstanding() == 0, "%d array / string copies were not released", fake_env_outstanding())

enum { PAGE = 4096 };
static unsigned char byte_of(long file, long idx, int k) { return (unsigned char)(file * 37 + idx * 11 + k * 3 + (k >> 5)); }
static int len_of(long idx) { return idx % 4 == 3 ? 1000 + (int)idx : PAGE; }

static jlongArray longs(const jlong* v, jsize n) {
  jlongArray a = fake_new_array(FAKE_LONG, n);
  if (n) memcpy(fake_data(a), v, (size_t)n * sizeof(jlong));
  return a;
}
static jintArray ints(const jint* v, jsize n) {
  jintArray a = fake_new_array(FAKE_INT, n);
  if (n && v) memcpy(fake_data(a), v, (size_t)n * sizeof(jint));
  return a;
}

int main(void) {
  JNIEnv* env = fake_env();
  jlongArray handle_out = fake_new_array(FAKE_LONG, 1);
  jlong h;
  int i, k;

  /* ---- GpuPageStore(options): nativeOpen */
  CHECK(JFN(nativeOpen)(env, NULL, 0, PAGE, 64 * PAGE, 0, 0, 0, 1, handle_out) == APS_OK, "nativeOpen");
  h = ((jlong*)fake_data(handle_out))[0];
  CHECK(h != 0, "no handle");
  BALANCED();

  /* ---- put / get / delete, one page per call, through a pinned direct buffer (allocPinned) */
  {
    jobject pinned = JFN(nativeAllocPinned)(env, NULL, PAGE);
    unsigned char* p = (unsigned char*)fake_data(pinned);
    jintArray out_len = fake_new_array(FAKE_INT, 1);
    CHECK(pinned != NULL, "nativeAllocPinned");
    for (i = 0; i < 10; ++i) {
      for (k = 0; k < len_of(i); ++k) p[k] = byte_of(7, i, k);
      CHECK(JFN(nativePut)(env, NULL, h, 0, 7, i, pinned, len_of(i)) == APS_OK, "nativePut %d", i);
    }
    for (i = 0; i < 10; ++i) {
      const int off = i * 13, want = len_of(i) - off;
      memset(p, 0xEE, PAGE);
      CHECK(JFN(nativeGet)(env, NULL, h, 0, 7, i, off, PAGE, pinned, PAGE, out_len) == APS_OK, "nativeGet %d", i);
      CHECK(((jint*)fake_data(out_len))[0] == want, "nativeGet %d: %d bytes, expected %d", i, ((jint*)fake_data(out_len))[0], want);
      for (k = 0; k < want; ++k)
        if (p[k] != byte_of(7, i, off + k)) {
          CHECK(0, "nativeGet %d: byte %d", i, k);
          break;
        }
    }
    CHECK(JFN(nativeGet)(env, NULL, h, 0, 7, 99, 0, PAGE, pinned, PAGE, out_len) == APS_ENOTFOUND, "absent page");
    CHECK(JFN(nativeGet)(env, NULL, h, 0, 7, 3, PAGE, 1, pinned, PAGE, out_len) == APS_EINVAL, "offset beyond a short page");
    CHECK(JFN(nativeDelete)(env, NULL, h, 0, 7, 0) == APS_OK && JFN(nativeDelete)(env, NULL, h, 0, 7, 0) == APS_ENOTFOUND, "nativeDelete");
    JFN(nativeFreePinned)(env, NULL, pinned);
    fake_free(pinned);
    fake_free(out_len);
    BALANCED();
  }

  /* ---- putBatch / getBatch / deleteBatch over ONE direct buffer with per-request offsets */
  {
    enum { N = 16 };
    unsigned char* mem = (unsigned char*)malloc((size_t)(N + 1) * PAGE); /* ByteBuffer.allocateDirect */
    jobject buf = fake_new_direct_buffer(mem, (jlong)(N + 1) * PAGE);
    jlong keys[3 * N], offs[N];
    jint lens[N], page_off[N], nbytes[N];
    jlongArray jkeys, joffs;
    jintArray jlens, jstatus, jcrc, jpo, jnb, jout_len;
    for (i = 0; i < N; ++i) {
      keys[3 * i] = 0x1122334455667788LL;
      keys[3 * i + 1] = -5 - i; /* negative longs: the unsigned halves of a 32-hex id */
      keys[3 * i + 2] = 1000 + i;
      offs[i] = (jlong)(N - 1 - i) * PAGE + 3; /* not in request order, not aligned */
      lens[i] = len_of(i);
      for (k = 0; k < lens[i]; ++k) mem[offs[i] + k] = byte_of(9, i, k);
    }
    jkeys = longs(keys, 3 * N);
    joffs = longs(offs, N);
    jlens = ints(lens, N);
    jstatus = ints(NULL, N);
    jcrc = ints(NULL, N);
    memset(fake_data(jstatus), 0x7F, N * sizeof(jint));
    CHECK(JFN(nativePutBatch)(env, NULL, h, N, jkeys, buf, joffs, jlens, jstatus, jcrc) == APS_OK, "nativePutBatch");
    for (i = 0; i < N; ++i) CHECK(((jint*)fake_data(jstatus))[i] == APS_OK, "put status %d = %d", i, ((jint*)fake_data(jstatus))[i]);
    BALANCED();
    /* the same keys twice in one put batch: refused as a whole, statuses untouched */
    keys[3] = keys[0]; keys[4] = keys[1]; keys[5] = keys[2];
    {
      jlongArray dup = longs(keys, 3 * N);
      CHECK(JFN(nativePutBatch)(env, NULL, h, N, dup, buf, joffs, jlens, jstatus, jcrc) == APS_EINVAL, "duplicate ids in a put batch");
      fake_free(dup);
    }
    keys[3] = 0x1122334455667788LL; keys[4] = -6; keys[5] = 1001;
    BALANCED();

    /* get: every page at its own offset and length, request 5 absent, request 6 with an offset beyond its page */
    memset(mem, 0xEE, (size_t)(N + 1) * PAGE);
    for (i = 0; i < N; ++i) {
      page_off[i] = i * 7;
      nbytes[i] = PAGE;
    }
    keys[3 * 5 + 2] = 424242;
    page_off[6] = PAGE + 1;
    fake_free(jkeys);
    jkeys = longs(keys, 3 * N);
    jpo = ints(page_off, N);
    jnb = ints(nbytes, N);
    jout_len = ints(NULL, N);
    CHECK(JFN(nativeGetBatch)(env, NULL, h, N, jkeys, jpo, jnb, buf, joffs, jout_len, jcrc) == APS_OK, "nativeGetBatch");
    for (i = 0; i < N; ++i) {
      const jint got = ((jint*)fake_data(jout_len))[i];
      const jint want = i == 5 ? -APS_ENOTFOUND : i == 6 ? -APS_EINVAL : len_of(i) - page_off[i];
      CHECK(got == want, "getBatch request %d: %d, expected %d", i, got, want);
      for (k = 0; k < got; ++k)
        if (mem[offs[i] + k] != byte_of(9, i, page_off[i] + k)) {
          CHECK(0, "getBatch request %d byte %d", i, k);
          break;
        }
      if (got > 0) CHECK(mem[offs[i] + got] == 0xEE, "getBatch request %d wrote past its length", i);
    }
    BALANCED();

    /* getPages(): four longs per resident page */
    {
      jlongArray listed = JFN(nativeIterate)(env, NULL, h);
      CHECK(listed != NULL && ((fake_obj*)listed)->n == 4 * (9 + N), "nativeIterate: %d longs", listed ? ((fake_obj*)listed)->n : -1);
      if (listed) {
        long total = 0;
        for (i = 0; i < ((fake_obj*)listed)->n; i += 4) total += (long)((jlong*)fake_data(listed))[i + 3];
        CHECK(total > 0, "lengths");
        fake_free(listed);
      }
      BALANCED();
    }

    /* delete: request 5 is absent */
    memset(fake_data(jstatus), 0x7F, N * sizeof(jint));
    CHECK(JFN(nativeDeleteBatch)(env, NULL, h, N, jkeys, jstatus) == APS_OK, "nativeDeleteBatch");
    for (i = 0; i < N; ++i)
      CHECK(((jint*)fake_data(jstatus))[i] == (i == 5 ? APS_ENOTFOUND : APS_OK), "delete status %d = %d", i, ((jint*)fake_data(jstatus))[i]);
    BALANCED();
    fake_free(jkeys); fake_free(joffs); fake_free(jlens); fake_free(jstatus); fake_free(jcrc); fake_free(jpo); fake_free(jnb);
    fake_free(jout_len); fake_free(buf);
    free(mem);
  }

  /* ---- md5Batch: ranges of one buffer, digests into a byte[] */
  {
    enum { F = 5 };
    unsigned char* mem = (unsigned char*)malloc(20000);
    jobject buf = fake_new_direct_buffer(mem, 20000);
    jlong off[F] = {0, 10, 10, 5000, 19999}, len[F] = {10, 0, 4990, 12345, 1};
    jlongArray joff = longs(off, F), jlen = longs(len, F);
    jbyteArray digests = fake_new_array(FAKE_BYTE, 16 * F);
    for (i = 0; i < 20000; ++i) mem[i] = (unsigned char)(i < 10 ? i : i * 7 + 1); /* bytes 0..9: the reference's own vector */
    CHECK(JFN(nativeMd5Batch)(env, NULL, 0, F, buf, joff, jlen, digests) == APS_OK, "nativeMd5Batch");
    for (i = 0; i < F; ++i) {
      uint8_t want[16];
      orc_md5(mem + off[i], (size_t)len[i], want);
      CHECK(memcmp((unsigned char*)fake_data(digests) + 16 * i, want, 16) == 0, "digest %d", i);
    }
    {
      static const unsigned char ref[16] = {0xc5, 0x6b, 0xd5, 0x48, 0x0f, 0x6e, 0x54, 0x13, 0xcb, 0x62, 0xa0, 0xad, 0x96, 0x66, 0x61, 0x3a};
      CHECK(memcmp(fake_data(digests), ref, 16) == 0, "ChecksumCommandIntegrationTest's vector (bytes 0..9)");
    }
    BALANCED();
    fake_free(joff); fake_free(jlen); fake_free(digests); fake_free(buf);
    free(mem);
  }

  /* ---- strings: error text; snapshot calls hand their paths and names over and release them whatever the outcome */
  {
    jstring text = JFN(nativeStrError)(env, NULL, APS_ENOTFOUND);
    jstring path = fake_new_string("/nonexistent/aps.snapshot");
    jobjectArray names = fake_new_array(FAKE_OBJECT, 2);
    jlongArray counts = fake_new_array(FAKE_LONG, 2);
    jobjectArray listed;
    CHECK(text && strlen((char*)fake_data(text)) > 0, "nativeStrError");
    ((jobject*)fake_data(names))[0] = fake_new_string("uuid-like-id");
    ((jobject*)fake_data(names))[1] = fake_new_string("/a/raw/path");
    CHECK(JFN(nativeSnapshotSave)(env, NULL, h, path, names) == APS_EIO, "snapshot save over the double");
    BALANCED();
    listed = JFN(nativeSnapshotNames)(env, NULL, path);
    CHECK(listed == NULL, "names of a missing snapshot");
    BALANCED();
    CHECK(JFN(nativeSnapshotLoad)(env, NULL, h, path, counts) == APS_EIO, "snapshot load over the double");
    BALANCED();
    fake_free(text); fake_free(path); fake_free(names); fake_free(counts);
  }

  /* ---- gpu.device = -1: one cache over every GPU.  GpuPageStore opens a store per device, attaches them to each other,
   * routes put / delete by nativeShardOf and reads any page through ONE sharded batch on shard 0. */
  {
    enum { M = 40 };
    jint shards = JFN(nativeDeviceCount)(env, NULL);
    jlong hs[8];
    jlongArray jhandles;
    jobject pinned = JFN(nativeAllocPinned)(env, NULL, M * PAGE);
    unsigned char* p = (unsigned char*)fake_data(pinned);
    jlong keys[3 * M], offs[M];
    jint po[M], nb[M], owner_pages[8] = {0};
    jlongArray jkeys, joffs;
    jintArray jpo, jnb, jout_len, jcrc;
    CHECK(shards == 2, "nativeDeviceCount = %d", shards);
    for (i = 0; i < shards; ++i) {
      CHECK(JFN(nativeOpen)(env, NULL, i, PAGE, 64 * PAGE, 128, 0, 0, 1, handle_out) == APS_OK, "nativeOpen shard %d", i);
      hs[i] = ((jlong*)fake_data(handle_out))[0];
    }
    jhandles = longs(hs, shards);
    for (i = 0; i < shards; ++i) CHECK(JFN(nativeShardAttachLocal)(env, NULL, jhandles, i) == APS_OK, "attach shard %d", i);
    CHECK(JFN(nativeShardAttachLocal)(env, NULL, jhandles, shards) == APS_EINVAL, "attach beyond the set");
    BALANCED();
    for (i = 0; i < M; ++i) {                      /* put: to the owner */
      const jint owner = JFN(nativeShardOf)(env, NULL, 0, 31, i, shards);
      jobject one = fake_new_direct_buffer(p + (size_t)i * PAGE, PAGE);
      CHECK(owner >= 0 && owner < shards, "owner %d", owner);
      ++owner_pages[owner];
      for (k = 0; k < len_of(i); ++k) p[(size_t)i * PAGE + k] = byte_of(31, i, k);
      CHECK(JFN(nativePut)(env, NULL, hs[owner], 0, 31, i, one, len_of(i)) == APS_OK, "put to owner");
      fake_free(one);
      keys[3 * i] = 0; keys[3 * i + 1] = 31; keys[3 * i + 2] = i;
      offs[i] = (jlong)i * PAGE;
      po[i] = 0;
      nb[i] = PAGE;
    }
    CHECK(owner_pages[0] > 0 && owner_pages[1] > 0, "routing put everything on one shard");
    memset(p, 0xEE, (size_t)M * PAGE);
    jkeys = longs(keys, 3 * M);
    joffs = longs(offs, M);
    jpo = ints(po, M);
    jnb = ints(nb, M);
    jout_len = ints(NULL, M);
    jcrc = ints(NULL, M);
    CHECK(JFN(nativeGetBatchSharded)(env, NULL, hs[0], M, jkeys, jpo, jnb, pinned, joffs, jout_len, jcrc) == APS_OK, "sharded get");
    for (i = 0; i < M; ++i) {
      CHECK(((jint*)fake_data(jout_len))[i] == len_of(i), "sharded get %d: %d bytes", i, ((jint*)fake_data(jout_len))[i]);
      for (k = 0; k < len_of(i); ++k)
        if (p[(size_t)i * PAGE + k] != byte_of(31, i, k)) {
          CHECK(0, "sharded get %d byte %d", i, k);
          break;
        }
    }
    /* the plain batched get on shard 0 sees only shard 0's pages */
    CHECK(JFN(nativeGetBatch)(env, NULL, hs[0], M, jkeys, jpo, jnb, pinned, joffs, jout_len, jcrc) == APS_OK, "local get");
    for (i = 0, k = 0; i < M; ++i) k += ((jint*)fake_data(jout_len))[i] >= 0;
    CHECK(k == owner_pages[0], "shard 0 answered %d pages, owns %d", k, owner_pages[0]);
    BALANCED();
    for (i = 0; i < shards; ++i) {                 /* getPages(): the shards' listings add up */
      jlongArray listed = JFN(nativeIterate)(env, NULL, hs[i]);
      CHECK(listed != NULL && ((fake_obj*)listed)->n == 4 * owner_pages[i], "shard %d lists %d longs", i, listed ? ((fake_obj*)listed)->n : -1);
      fake_free(listed);
      CHECK(JFN(nativeClose)(env, NULL, hs[i]) == APS_OK, "close shard %d", i);
    }
    JFN(nativeFreePinned)(env, NULL, pinned);
    fake_free(pinned); fake_free(jhandles); fake_free(jkeys); fake_free(joffs); fake_free(jpo); fake_free(jnb);
    fake_free(jout_len); fake_free(jcrc);
    BALANCED();
  }

  /* ---- close() is idempotent; a closed store answers APS_ECLOSED, and getPages() reports the failure as null */
  CHECK(JFN(nativeClose)(env, NULL, h) == APS_OK && JFN(nativeClose)(env, NULL, h) == APS_OK, "nativeClose twice");
  CHECK(JFN(nativeDelete)(env, NULL, h, 0, 7, 1) == APS_ECLOSED, "delete on a closed store");
  CHECK(JFN(nativeIterate)(env, NULL, h) == NULL, "nativeIterate on a closed store");
  BALANCED();
  fake_free(handle_out);
  if (g_failures) {
    fprintf(stderr, "%d check(s) failed\n", g_failures);
    return 1;
  }
  printf("ok\n");
  return 0;
}
