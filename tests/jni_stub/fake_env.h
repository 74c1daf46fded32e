/* TEST INFRASTRUCTURE ONLY: the objects and constructors of the mock JNIEnv in fake_env.c. */
#ifndef FAKE_ENV_H_
#define FAKE_ENV_H_
#include <stddef.h>

#include "jni.h"

enum { FAKE_INT = 1, FAKE_LONG, FAKE_BYTE, FAKE_OBJECT, FAKE_BUFFER, FAKE_STRING };
typedef struct fake_obj {
  int kind;
  jsize n;        /* elements (arrays), bytes including the terminator (strings) */
  size_t elem;    /* element size */
  void* data;     /* the Java heap copy of an array / string; the native address of a direct buffer */
  jlong capacity; /* direct buffers */
} fake_obj;

JNIEnv* fake_env(void);
int fake_env_outstanding(void); /* Get...Elements / GetStringUTFChars results not yet released */
jarray fake_new_array(int kind, jsize n);
jobject fake_new_direct_buffer(void* addr, jlong capacity);
jstring fake_new_string(const char* utf);
void fake_free(jobject obj);
void* fake_data(jobject obj); /* the "Java heap" contents, as the Java caller would see them after the call */
#endif
