/* Minimal stand-in for the JDK's jni.h (absent from this image), just enough to type-check integration/jni/aps_jni.c
 * against include/aps.h in tests/test_c_abi.py, and to RUN it over the mock JNIEnv of fake_env.c (tests/test_jni_shim_runs.py).
 * Declares only what the shim uses. */
#ifndef JNI_STUB_H_
#define JNI_STUB_H_
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef signed char jbyte; typedef unsigned char jboolean;
typedef void* jobject; typedef jobject jclass; typedef jobject jstring; typedef jobject jarray; typedef jarray jintArray;
typedef jarray jlongArray; typedef jarray jbyteArray; typedef jarray jobjectArray; typedef jint jsize;
#define JNI_ABORT 2
#define JNIEXPORT
#define JNICALL
struct JNINativeInterface_;
typedef const struct JNINativeInterface_* JNIEnv;
struct JNINativeInterface_ {
  void* (*GetDirectBufferAddress)(JNIEnv*, jobject);
  jlong (*GetDirectBufferCapacity)(JNIEnv*, jobject);
  void (*SetIntArrayRegion)(JNIEnv*, jintArray, jsize, jsize, const jint*);
  void (*SetLongArrayRegion)(JNIEnv*, jlongArray, jsize, jsize, const jlong*);
  jlongArray (*NewLongArray)(JNIEnv*, jsize);
  jobject (*NewDirectByteBuffer)(JNIEnv*, void*, jlong);
  jstring (*NewStringUTF)(JNIEnv*, const char*);
  const char* (*GetStringUTFChars)(JNIEnv*, jstring, jboolean*);
  void (*ReleaseStringUTFChars)(JNIEnv*, jstring, const char*);
  jlong* (*GetLongArrayElements)(JNIEnv*, jlongArray, jboolean*);
  void (*ReleaseLongArrayElements)(JNIEnv*, jlongArray, jlong*, jint);
  jint* (*GetIntArrayElements)(JNIEnv*, jintArray, jboolean*);
  void (*ReleaseIntArrayElements)(JNIEnv*, jintArray, jint*, jint);
  jbyte* (*GetByteArrayElements)(JNIEnv*, jbyteArray, jboolean*);
  void (*ReleaseByteArrayElements)(JNIEnv*, jbyteArray, jbyte*, jint);
  jsize (*GetArrayLength)(JNIEnv*, jarray);
  jobject (*GetObjectArrayElement)(JNIEnv*, jobjectArray, jsize);
  void (*SetObjectArrayElement)(JNIEnv*, jobjectArray, jsize, jobject);
  jobjectArray (*NewObjectArray)(JNIEnv*, jsize, jclass, jobject);
  jclass (*FindClass)(JNIEnv*, const char*);
};
#endif
