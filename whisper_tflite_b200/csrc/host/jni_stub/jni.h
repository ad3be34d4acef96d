// Declaration-only stand-in for the JDK's <jni.h>: just the types and JNIEnv members jni_whisper.cpp touches,
// so the bridge is at least compiled on a box without a JDK. Never linked into anything that runs.
#pragma once
#include <cstdint>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL

using jlong = int64_t;
using jint = int32_t;
using jsize = jint;
using jfloat = float;
using jboolean = uint8_t;
class _jobject {};
using jobject = _jobject*;
using jstring = jobject;
using jfloatArray = jobject;

struct JNIEnv {
  const char* GetStringUTFChars(jstring s, jboolean* is_copy);
  void ReleaseStringUTFChars(jstring s, const char* chars);
  jsize GetArrayLength(jfloatArray a);
  jfloat* GetFloatArrayElements(jfloatArray a, jboolean* is_copy);
  void ReleaseFloatArrayElements(jfloatArray a, jfloat* elems, jint mode);
  jstring NewStringUTF(const char* utf);
};
