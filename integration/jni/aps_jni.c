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

typedef struct { jlong* v; size_t n, cap; } vec;
static int collect(void* user, const aps_key* key, uint32_t len) {
  vec* o = (vec*)user;
  if (o->n + 4 > o->cap) {
    o->cap = o->cap ? o->cap * 2 : 1024;
    o->v = (jlong*)realloc(o->v, o->cap * sizeof(jlong));
    if (!o->v) return 1;
  }
  o->v[o->n++] = (jlong)key->file_hi;
  o->v[o->n++] = (jlong)key->file_lo;
  o->v[o->n++] = (jlong)key->page_index;
  o->v[o->n++] = (jlong)len;
  return 0;
}
JNIEXPORT jlongArray JNICALL JFN(nativeIterate)(JNIEnv* env, jclass cls, jlong handle) {
  vec o = {NULL, 0, 0};
  aps_iterate((aps_store*)(intptr_t)handle, collect, &o);
  jlongArray arr = (*env)->NewLongArray(env, (jsize)o.n);
  if (arr && o.n) (*env)->SetLongArrayRegion(env, arr, 0, (jsize)o.n, o.v);
  free(o.v);
  return arr;
}

/* A page-locked direct ByteBuffer: kernels read and write it over PCIe without a bounce copy. */
JNIEXPORT jobject JNICALL JFN(nativeAllocPinned)(JNIEnv* env, jclass cls, jint bytes) {
  void* p = NULL;
  if (cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) return NULL;
  return (*env)->NewDirectByteBuffer(env, p, bytes);
}

JNIEXPORT jstring JNICALL JFN(nativeStrError)(JNIEnv* env, jclass cls, jint code) {
  return (*env)->NewStringUTF(env, aps_strerror(code));
}

JNIEXPORT jint JNICALL JFN(nativeSnapshotSave)(JNIEnv* env, jclass cls, jlong handle, jstring path) {
  const char* p = (*env)->GetStringUTFChars(env, path, NULL);
  int rc;
  (void)cls;
  if (!p) return APS_ENOMEM;
  rc = aps_snapshot_save((aps_store*)(intptr_t)handle, p, NULL);
  (*env)->ReleaseStringUTFChars(env, path, p);
  return rc;
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
