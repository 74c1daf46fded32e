/* TEST INFRASTRUCTURE ONLY.  A small JNIEnv for running integration/jni/aps_jni.c without a JVM: Java arrays, strings
 * and direct ByteBuffers are tagged C structs, and the array accessors behave like the STRICTEST JVM the JNI
 * specification allows -- Get<Type>ArrayElements always hands out a COPY, Release writes it back unless the mode is
 * JNI_ABORT, and every copy must be released exactly once (fake_env_outstanding() counts the ones that were not).  A shim
 * that stores results into an array it releases with JNI_ABORT, forgets a release, or touches an array after releasing it
 * fails here (the last under AddressSanitizer).  Declared in fake_env.h; the table layout is that of tests/jni_stub/jni.h,
 * which the shim is compiled against in the same build. */
#include "fake_env.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int g_outstanding = 0;
int fake_env_outstanding(void) { return g_outstanding; }

static fake_obj* new_obj(int kind, jsize n, size_t elem) {
  fake_obj* o = (fake_obj*)calloc(1, sizeof(*o));
  if (!o) abort();
  o->kind = kind;
  o->n = n;
  o->elem = elem;
  o->data = n && elem ? calloc((size_t)n, elem) : NULL;
  return o;
}
jarray fake_new_array(int kind, jsize n) {
  const size_t elem = kind == FAKE_INT ? sizeof(jint) : kind == FAKE_LONG ? sizeof(jlong) : kind == FAKE_BYTE ? 1 : sizeof(jobject);
  return new_obj(kind, n, elem);
}
jobject fake_new_direct_buffer(void* addr, jlong capacity) {
  fake_obj* o = new_obj(FAKE_BUFFER, 0, 0);
  o->data = addr;
  o->capacity = capacity;
  return o;
}
jstring fake_new_string(const char* utf) {
  fake_obj* o = new_obj(FAKE_STRING, (jsize)strlen(utf) + 1, 1);
  memcpy(o->data, utf, strlen(utf) + 1);
  return o;
}
void fake_free(jobject obj) {
  fake_obj* o = (fake_obj*)obj;
  if (!o) return;
  if (o->kind == FAKE_OBJECT)
    for (jsize i = 0; i < o->n; ++i) fake_free(((jobject*)o->data)[i]);
  if (o->kind != FAKE_BUFFER) free(o->data);
  free(o);
}
void* fake_data(jobject obj) { return ((fake_obj*)obj)->data; }

static fake_obj* expect(jobject obj, int kind, const char* what) {
  fake_obj* o = (fake_obj*)obj;
  if (!o || o->kind != kind) {
    fprintf(stderr, "fake JNIEnv: %s called on an object of kind %d\n", what, o ? o->kind : -1);
    abort();
  }
  return o;
}

static void* get_elements(jobject arr, int kind, jboolean* is_copy, const char* what) {
  fake_obj* o = expect(arr, kind, what);
  void* copy = malloc(o->n ? (size_t)o->n * o->elem : 1);
  if (o->n) memcpy(copy, o->data, (size_t)o->n * o->elem);
  if (is_copy) *is_copy = 1;
  ++g_outstanding;
  return copy;
}
static void release_elements(jobject arr, int kind, void* elems, jint mode, const char* what) {
  fake_obj* o = expect(arr, kind, what);
  if (mode != JNI_ABORT && o->n) memcpy(o->data, elems, (size_t)o->n * o->elem);
  free(elems);
  --g_outstanding;
}

static void* GetDirectBufferAddress(JNIEnv* env, jobject buf) { (void)env; return expect(buf, FAKE_BUFFER, "GetDirectBufferAddress")->data; }
static jlong GetDirectBufferCapacity(JNIEnv* env, jobject buf) { (void)env; return expect(buf, FAKE_BUFFER, "GetDirectBufferCapacity")->capacity; }
static void SetIntArrayRegion(JNIEnv* env, jintArray a, jsize start, jsize len, const jint* src) {
  fake_obj* o = expect(a, FAKE_INT, "SetIntArrayRegion");
  (void)env;
  if (start < 0 || len < 0 || start + len > o->n) { fprintf(stderr, "ArrayIndexOutOfBoundsException\n"); abort(); }
  memcpy((jint*)o->data + start, src, (size_t)len * sizeof(jint));
}
static void SetLongArrayRegion(JNIEnv* env, jlongArray a, jsize start, jsize len, const jlong* src) {
  fake_obj* o = expect(a, FAKE_LONG, "SetLongArrayRegion");
  (void)env;
  if (start < 0 || len < 0 || start + len > o->n) { fprintf(stderr, "ArrayIndexOutOfBoundsException\n"); abort(); }
  memcpy((jlong*)o->data + start, src, (size_t)len * sizeof(jlong));
}
static jlongArray NewLongArray(JNIEnv* env, jsize n) { (void)env; return fake_new_array(FAKE_LONG, n); }
static jobject NewDirectByteBuffer(JNIEnv* env, void* addr, jlong cap) { (void)env; return fake_new_direct_buffer(addr, cap); }
static jstring NewStringUTF(JNIEnv* env, const char* utf) { (void)env; return fake_new_string(utf); }
static const char* GetStringUTFChars(JNIEnv* env, jstring s, jboolean* is_copy) { (void)env; return (const char*)get_elements(s, FAKE_STRING, is_copy, "GetStringUTFChars"); }
static void ReleaseStringUTFChars(JNIEnv* env, jstring s, const char* utf) { (void)env; release_elements(s, FAKE_STRING, (void*)utf, JNI_ABORT, "ReleaseStringUTFChars"); }
static jlong* GetLongArrayElements(JNIEnv* env, jlongArray a, jboolean* c) { (void)env; return (jlong*)get_elements(a, FAKE_LONG, c, "GetLongArrayElements"); }
static void ReleaseLongArrayElements(JNIEnv* env, jlongArray a, jlong* e, jint mode) { (void)env; release_elements(a, FAKE_LONG, e, mode, "ReleaseLongArrayElements"); }
static jint* GetIntArrayElements(JNIEnv* env, jintArray a, jboolean* c) { (void)env; return (jint*)get_elements(a, FAKE_INT, c, "GetIntArrayElements"); }
static void ReleaseIntArrayElements(JNIEnv* env, jintArray a, jint* e, jint mode) { (void)env; release_elements(a, FAKE_INT, e, mode, "ReleaseIntArrayElements"); }
static jbyte* GetByteArrayElements(JNIEnv* env, jbyteArray a, jboolean* c) { (void)env; return (jbyte*)get_elements(a, FAKE_BYTE, c, "GetByteArrayElements"); }
static void ReleaseByteArrayElements(JNIEnv* env, jbyteArray a, jbyte* e, jint mode) { (void)env; release_elements(a, FAKE_BYTE, e, mode, "ReleaseByteArrayElements"); }
static jsize GetArrayLength(JNIEnv* env, jarray a) { (void)env; return ((fake_obj*)a)->n; }
static jobject GetObjectArrayElement(JNIEnv* env, jobjectArray a, jsize i) {
  fake_obj* o = expect(a, FAKE_OBJECT, "GetObjectArrayElement");
  (void)env;
  if (i < 0 || i >= o->n) { fprintf(stderr, "ArrayIndexOutOfBoundsException\n"); abort(); }
  return ((jobject*)o->data)[i];
}
static void SetObjectArrayElement(JNIEnv* env, jobjectArray a, jsize i, jobject v) {
  fake_obj* o = expect(a, FAKE_OBJECT, "SetObjectArrayElement");
  (void)env;
  if (i < 0 || i >= o->n) { fprintf(stderr, "ArrayIndexOutOfBoundsException\n"); abort(); }
  ((jobject*)o->data)[i] = v;
}
static jobjectArray NewObjectArray(JNIEnv* env, jsize n, jclass cls, jobject init) { (void)env; (void)cls; (void)init; return fake_new_array(FAKE_OBJECT, n); }
static jclass FindClass(JNIEnv* env, const char* name) { (void)env; return (jclass)name; }

static const struct JNINativeInterface_ g_table = {
    GetDirectBufferAddress, GetDirectBufferCapacity, SetIntArrayRegion, SetLongArrayRegion, NewLongArray, NewDirectByteBuffer,
    NewStringUTF, GetStringUTFChars, ReleaseStringUTFChars, GetLongArrayElements, ReleaseLongArrayElements, GetIntArrayElements,
    ReleaseIntArrayElements, GetByteArrayElements, ReleaseByteArrayElements, GetArrayLength, GetObjectArrayElement,
    SetObjectArrayElement, NewObjectArray, FindClass};
static JNIEnv g_env = &g_table;
JNIEnv* fake_env(void) { return &g_env; }
