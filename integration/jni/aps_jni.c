/*
 * JNI shim between alluxio.client.file.cache.store.GpuPageStore and libaps.so (include/aps.h).
 * Thin by design: unpack arguments, call one aps_* function, return its status.
 * Build (on a machine with a JDK):  gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux \
 *     -I../../include aps_jni.c -L../../alluxio_b200 -laps -lcudart -o libaps_jni.so
 * Not compiled in this repository: the build image has no jni.h.
 */
#include <jni.h>
#include <stdlib.h>
#include <string.h>
#include <cuda_runtime_api.h>

#include "aps.h"
#include "aps_marshal.h"

#define JFN(name) Java_alluxio_client_file_cache_store_GpuPageStore_##name

JNIEXPORT jint JNICALL JFN(nativeOpen)(JNIEnv* env, jclass cls, jint device, jint page_size, jlong cache_bytes,
                                       jlong slab_slots, jlong table_capacity, jint max_batch, jint crc,
                                       jlongArray handle_out) {
  aps_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.device = device;
  cfg.page_size = (uint32_t)page_size;
  cfg.cache_bytes = (uint64_t)cache_bytes;
  cfg.slab_slots = (uint64_t)slab_slots;
  cfg.table_capacity = (uint64_t)table_capacity;
  cfg.max_batch = (uint32_t)max_batch;
  cfg.flags = crc ? APS_FLAG_CRC : 0;
  aps_store* s = NULL;
  int rc = aps_open(&cfg, &s);
  jlong h = (jlong)(intptr_t)s;
  (*env)->SetLongArrayRegion(env, handle_out, 0, 1, &h);
  return rc;
}

JNIEXPORT jint JNICALL JFN(nativeClose)(JNIEnv* env, jclass cls, jlong handle) {
  return aps_close((aps_store*)(intptr_t)handle);
}

JNIEXPORT jint JNICALL JFN(nativePut)(JNIEnv* env, jclass cls, jlong handle, jlong hi, jlong lo, jlong idx,
                                      jobject pinned, jint len) {
  aps_key k = {(uint64_t)hi, (uint64_t)lo, (uint64_t)idx};
  return aps_put((aps_store*)(intptr_t)handle, &k, (*env)->GetDirectBufferAddress(env, pinned), (uint32_t)len);
}

JNIEXPORT jint JNICALL JFN(nativeGet)(JNIEnv* env, jclass cls, jlong handle, jlong hi, jlong lo, jlong idx,
                                      jint page_off, jint bytes_to_read, jobject pinned, jint room,
                                      jintArray out_len) {
  aps_key k = {(uint64_t)hi, (uint64_t)lo, (uint64_t)idx};
  uint32_t got = 0;
  int rc = aps_get((aps_store*)(intptr_t)handle, &k, (uint32_t)page_off, (uint32_t)bytes_to_read,
                   (*env)->GetDirectBufferAddress(env, pinned), (uint32_t)room, &got, NULL);
  jint g = (jint)got;
  (*env)->SetIntArrayRegion(env, out_len, 0, 1, &g);
  return rc;
}

JNIEXPORT jint JNICALL JFN(nativeDelete)(JNIEnv* env, jclass cls, jlong handle, jlong hi, jlong lo, jlong idx) {
  aps_key k = {(uint64_t)hi, (uint64_t)lo, (uint64_t)idx};
  return aps_delete((aps_store*)(intptr_t)handle, &k);
}

typedef struct { jlong* v; size_t n, cap; int failed; } vec;
static int collect(void* user, const aps_key* key, uint32_t len) {
  vec* o = (vec*)user;
  if (o->n + 4 > o->cap) {
    size_t cap = o->cap ? o->cap * 2 : 1024;
    jlong* grown = (jlong*)realloc(o->v, cap * sizeof(jlong));
    if (!grown) {  /* keep what we have (freed by the caller) and stop the walk */
      o->failed = 1;
      return 1;
    }
    o->v = grown;
    o->cap = cap;
  }
  o->v[o->n++] = (jlong)key->file_hi;
  o->v[o->n++] = (jlong)key->file_lo;
  o->v[o->n++] = (jlong)key->page_index;
  o->v[o->n++] = (jlong)len;
  return 0;
}
/* getPages(): (fileHi, fileLo, pageIndex, length) per resident page, or NULL when the store could not be listed
 * (closed, CUDA failure, out of memory) -- GpuPageStore.getPages turns NULL into an IOException, so that a failed
 * listing is never mistaken for an empty store by LocalCacheManager.restore. */
JNIEXPORT jlongArray JNICALL JFN(nativeIterate)(JNIEnv* env, jclass cls, jlong handle) {
  vec o = {NULL, 0, 0, 0};
  jlongArray arr = NULL;
  int rc = aps_iterate((aps_store*)(intptr_t)handle, collect, &o);
  (void)cls;
  if (rc == APS_OK && !o.failed) {
    arr = (*env)->NewLongArray(env, (jsize)o.n);
    if (arr && o.n) (*env)->SetLongArrayRegion(env, arr, 0, (jsize)o.n, o.v);
  }
  free(o.v);
  return arr;
}

/* ---- batched calls: ONE aps_*_batch call per JNI call; pages move through one direct ByteBuffer ------------------
 * (GetDirectBufferAddress: no copy, no pinning by the JVM).  The marshalling itself is in aps_marshal.h. */
static jint get_batch(JNIEnv* env, int sharded, jlong handle, jint n, jlongArray keys, jintArray page_off, jintArray nbytes,
                      jobject buffer, jlongArray off, jintArray out_len, jintArray out_crc) {
  jlong* k = (*env)->GetLongArrayElements(env, keys, NULL);
  jint* po = (*env)->GetIntArrayElements(env, page_off, NULL);
  jint* nb = (*env)->GetIntArrayElements(env, nbytes, NULL);
  jlong* of = (*env)->GetLongArrayElements(env, off, NULL);
  jint* ol = (*env)->GetIntArrayElements(env, out_len, NULL);
  jint* oc = (*env)->GetIntArrayElements(env, out_crc, NULL);
  void* base = (*env)->GetDirectBufferAddress(env, buffer);
  int rc = (k && po && nb && of && ol && oc && base)
               ? aps_marshal_get_batch_ex((aps_store*)(intptr_t)handle, sharded, n, (const int64_t*)k, (const int32_t*)po,
                                          (const int32_t*)nb, base, (const int64_t*)of, (int32_t*)ol, (int32_t*)oc)
               : APS_EINVAL;
  if (oc) (*env)->ReleaseIntArrayElements(env, out_crc, oc, 0);
  if (ol) (*env)->ReleaseIntArrayElements(env, out_len, ol, 0);
  if (of) (*env)->ReleaseLongArrayElements(env, off, of, JNI_ABORT);
  if (nb) (*env)->ReleaseIntArrayElements(env, nbytes, nb, JNI_ABORT);
  if (po) (*env)->ReleaseIntArrayElements(env, page_off, po, JNI_ABORT);
  if (k) (*env)->ReleaseLongArrayElements(env, keys, k, JNI_ABORT);
  return rc;
}
JNIEXPORT jint JNICALL JFN(nativeGetBatch)(JNIEnv* env, jclass cls, jlong handle, jint n, jlongArray keys, jintArray page_off,
                                           jintArray nbytes, jobject buffer, jlongArray off, jintArray out_len, jintArray out_crc) {
  (void)cls;
  return get_batch(env, 0, handle, n, keys, page_off, nbytes, buffer, off, out_len, out_crc);
}
/* The same over a cache sharded across the GPUs of the box: `handle` is shard 0, every page comes from its owner. */
JNIEXPORT jint JNICALL JFN(nativeGetBatchSharded)(JNIEnv* env, jclass cls, jlong handle, jint n, jlongArray keys,
                                                  jintArray page_off, jintArray nbytes, jobject buffer, jlongArray off,
                                                  jintArray out_len, jintArray out_crc) {
  (void)cls;
  return get_batch(env, 1, handle, n, keys, page_off, nbytes, buffer, off, out_len, out_crc);
}

/* ---- one cache over several GPUs: a page lives in the store of GPU aps_shard_of(key, shards) ---------------------- */
JNIEXPORT jint JNICALL JFN(nativeDeviceCount)(JNIEnv* env, jclass cls) {
  int n = 0;
  (void)env;
  (void)cls;
  return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}
JNIEXPORT jint JNICALL JFN(nativeShardOf)(JNIEnv* env, jclass cls, jlong hi, jlong lo, jlong idx, jint shards) {
  aps_key k = {(uint64_t)hi, (uint64_t)lo, (uint64_t)idx};
  (void)env;
  (void)cls;
  return (jint)aps_shard_of(&k, (uint32_t)shards);
}
/* handles = the stores of all shards (GpuPageStore.mHandles); lets shard my_shard read the others in place. */
JNIEXPORT jint JNICALL JFN(nativeShardAttachLocal)(JNIEnv* env, jclass cls, jlongArray handles, jint my_shard) {
  jsize n = (*env)->GetArrayLength(env, handles), i;
  jlong* h = (*env)->GetLongArrayElements(env, handles, NULL);
  aps_store** stores = h ? (aps_store**)calloc((size_t)n, sizeof(*stores)) : NULL;
  int rc = APS_ENOMEM;
  (void)cls;
  if (stores) {
    for (i = 0; i < n; ++i) stores[i] = (aps_store*)(intptr_t)h[i];
    rc = my_shard >= 0 && my_shard < n ? aps_shard_attach_local(stores[my_shard], (uint32_t)n, (uint32_t)my_shard, stores) : APS_EINVAL;
  }
  free(stores);
  if (h) (*env)->ReleaseLongArrayElements(env, handles, h, JNI_ABORT);
  return rc;
}

JNIEXPORT jint JNICALL JFN(nativePutBatch)(JNIEnv* env, jclass cls, jlong handle, jint n, jlongArray keys, jobject buffer,
                                           jlongArray off, jintArray len, jintArray out_status, jintArray out_crc) {
  jlong* k = (*env)->GetLongArrayElements(env, keys, NULL);
  jlong* of = (*env)->GetLongArrayElements(env, off, NULL);
  jint* ln = (*env)->GetIntArrayElements(env, len, NULL);
  jint* os = (*env)->GetIntArrayElements(env, out_status, NULL);
  jint* oc = (*env)->GetIntArrayElements(env, out_crc, NULL);
  void* base = (*env)->GetDirectBufferAddress(env, buffer);
  int rc = (k && of && ln && os && oc && base)
               ? aps_marshal_put_batch((aps_store*)(intptr_t)handle, n, (const int64_t*)k, base, (const int64_t*)of, (const int32_t*)ln,
                                       (int32_t*)os, (int32_t*)oc)
               : APS_EINVAL;
  (void)cls;
  if (oc) (*env)->ReleaseIntArrayElements(env, out_crc, oc, 0);
  if (os) (*env)->ReleaseIntArrayElements(env, out_status, os, 0);
  if (ln) (*env)->ReleaseIntArrayElements(env, len, ln, JNI_ABORT);
  if (of) (*env)->ReleaseLongArrayElements(env, off, of, JNI_ABORT);
  if (k) (*env)->ReleaseLongArrayElements(env, keys, k, JNI_ABORT);
  return rc;
}

JNIEXPORT jint JNICALL JFN(nativeDeleteBatch)(JNIEnv* env, jclass cls, jlong handle, jint n, jlongArray keys, jintArray out_status) {
  jlong* k = (*env)->GetLongArrayElements(env, keys, NULL);
  jint* os = (*env)->GetIntArrayElements(env, out_status, NULL);
  int rc = (k && os) ? aps_marshal_delete_batch((aps_store*)(intptr_t)handle, n, (const int64_t*)k, (int32_t*)os) : APS_EINVAL;
  (void)cls;
  if (os) (*env)->ReleaseIntArrayElements(env, out_status, os, 0);
  if (k) (*env)->ReleaseLongArrayElements(env, keys, k, JNI_ABORT);
  return rc;
}

/* `alluxio fs checksum` (ChecksumCommand.calculateChecksum): MD5 of n files whose bytes lie in one PINNED direct buffer
 * (nativeAllocPinned; the kernel reads it in place over PCIe).  digests = byte[16 * n]. */
JNIEXPORT jint JNICALL JFN(nativeMd5Batch)(JNIEnv* env, jclass cls, jint device, jint n, jobject pinned, jlongArray off,
                                           jlongArray len, jbyteArray digests) {
  jlong* of = (*env)->GetLongArrayElements(env, off, NULL);
  jlong* ln = (*env)->GetLongArrayElements(env, len, NULL);
  jbyte* out = (*env)->GetByteArrayElements(env, digests, NULL);
  void* base = (*env)->GetDirectBufferAddress(env, pinned);
  int rc = (of && ln && out && base)
               ? aps_marshal_md5_batch(device, n, base, (const int64_t*)of, (const int64_t*)ln, (unsigned char*)out)
               : APS_EINVAL;
  (void)cls;
  if (out) (*env)->ReleaseByteArrayElements(env, digests, out, 0);
  if (ln) (*env)->ReleaseLongArrayElements(env, len, ln, JNI_ABORT);
  if (of) (*env)->ReleaseLongArrayElements(env, off, of, JNI_ABORT);
  return rc;
}

/* A page-locked direct ByteBuffer: kernels read and write it over PCIe without a bounce copy. */
JNIEXPORT jobject JNICALL JFN(nativeAllocPinned)(JNIEnv* env, jclass cls, jint bytes) {
  void* p = NULL;
  if (cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) return NULL;
  return (*env)->NewDirectByteBuffer(env, p, bytes);
}

JNIEXPORT void JNICALL JFN(nativeFreePinned)(JNIEnv* env, jclass cls, jobject pinned) {
  void* p = pinned ? (*env)->GetDirectBufferAddress(env, pinned) : NULL;
  (void)cls;
  if (p) cudaFreeHost(p);
}

JNIEXPORT jstring JNICALL JFN(nativeStrError)(JNIEnv* env, jclass cls, jint code) {
  return (*env)->NewStringUTF(env, aps_strerror(code));
}

/* names = the store's table of interned file ids (GpuPageStore.mInternedNames), written into the snapshot so that the
 * ids survive a restart (aps_snapshot_save_names). */
JNIEXPORT jint JNICALL JFN(nativeSnapshotSave)(JNIEnv* env, jclass cls, jlong handle, jstring path, jobjectArray names) {
  const char* p = (*env)->GetStringUTFChars(env, path, NULL);
  jsize n = names ? (*env)->GetArrayLength(env, names) : 0, i;
  const char** c_names = n ? (const char**)calloc((size_t)n, sizeof(char*)) : NULL;
  int rc = APS_OK;
  (void)cls;
  if (!p || (n && !c_names)) rc = APS_ENOMEM;
  for (i = 0; i < n && rc == APS_OK; ++i) {
    jstring js = (jstring)(*env)->GetObjectArrayElement(env, names, i);
    c_names[i] = js ? (*env)->GetStringUTFChars(env, js, NULL) : NULL;
    if (!c_names[i]) rc = APS_ENOMEM;
  }
  if (rc == APS_OK) rc = aps_snapshot_save_names((aps_store*)(intptr_t)handle, p, (uint32_t)n, c_names, NULL);
  for (i = 0; i < n; ++i)
    if (c_names && c_names[i]) (*env)->ReleaseStringUTFChars(env, (jstring)(*env)->GetObjectArrayElement(env, names, i), c_names[i]);
  free((void*)c_names);
  if (p) (*env)->ReleaseStringUTFChars(env, path, p);
  return rc;
}

/* The table of interned file ids a snapshot carries, in index order (empty array when it has none, NULL on error). */
typedef struct { JNIEnv* env; jobjectArray arr; jsize cap; int failed; } name_sink;
static int count_name(void* user, uint32_t index, const char* name) {
  (void)name;
  *(uint32_t*)user = index + 1;
  return 0;
}
static int store_name(void* user, uint32_t index, const char* name) {
  name_sink* k = (name_sink*)user;
  jstring js = (*k->env)->NewStringUTF(k->env, name);
  if (!js || (jsize)index >= k->cap) {
    k->failed = 1;
    return 1;
  }
  (*k->env)->SetObjectArrayElement(k->env, k->arr, (jsize)index, js);
  return 0;
}
JNIEXPORT jobjectArray JNICALL JFN(nativeSnapshotNames)(JNIEnv* env, jclass cls, jstring path) {
  const char* p = (*env)->GetStringUTFChars(env, path, NULL);
  uint32_t n = 0;
  name_sink k;
  (void)cls;
  if (!p) return NULL;
  k.env = env;
  k.arr = NULL;
  k.cap = 0;
  k.failed = 0;
  if (aps_snapshot_names(p, count_name, &n) == APS_OK) {
    k.arr = (*env)->NewObjectArray(env, (jsize)n, (*env)->FindClass(env, "java/lang/String"), NULL);
    k.cap = (jsize)n;
    if (k.arr && n && (aps_snapshot_names(p, store_name, &k) != APS_OK || k.failed)) k.arr = NULL;
  }
  (*env)->ReleaseStringUTFChars(env, path, p);
  return k.arr;
}

JNIEXPORT jint JNICALL JFN(nativeSnapshotLoad)(JNIEnv* env, jclass cls, jlong handle, jstring path, jlongArray counts) {
  const char* p = (*env)->GetStringUTFChars(env, path, NULL);
  uint64_t loaded = 0, rejected = 0;
  jlong out[2];
  int rc;
  (void)cls;
  if (!p) return APS_ENOMEM;
  rc = aps_snapshot_load((aps_store*)(intptr_t)handle, p, &loaded, &rejected);
  (*env)->ReleaseStringUTFChars(env, path, p);
  out[0] = (jlong)loaded;
  out[1] = (jlong)rejected;
  (*env)->SetLongArrayRegion(env, counts, 0, 2, out);
  return rc;
}
