// JNI bridge for io.github.jerinphilip.whisper.EngineNative — same four natives, same handle convention
// (Engine* as jlong) as /root/reference/bindings/java/whisper.tflite.cpp:17-71, over the B200 engines.
// No JDK ships in this image: the build compiles this file against jni_stub/jni.h (declarations only) to keep
// it honest; with a real JDK pass -I$JAVA_HOME/include -I$JAVA_HOME/include/linux instead.
#include <jni.h>

#include <string>
#include <vector>

#include "whisper.h"

namespace {
struct Utf {
  JNIEnv* env;
  jstring s;
  const char* c;
  Utf(JNIEnv* e, jstring str) : env(e), s(str), c(e->GetStringUTFChars(str, nullptr)) {}
  ~Utf() { env->ReleaseStringUTFChars(s, c); }
};
whisper::Engine* engine_of(jlong handle) { return reinterpret_cast<whisper::Engine*>(handle); }  // NOLINT
}  // namespace

extern "C" {

JNIEXPORT jlong JNICALL Java_io_github_jerinphilip_whisper_EngineNative_create(JNIEnv* env, jobject /*thiz*/,
                                                                              jlong engineType, jstring modelPath,
                                                                              jstring vocabPath,
                                                                              jboolean isMultilingual) {
  Utf model(env, modelPath), vocab(env, vocabPath);
  fprintf(stderr, "model: %s\nvocab: %s\n", model.c, vocab.c);
  whisper::Engine* engine =
      whisper::create_engine(static_cast<whisper::EngineType>(engineType), model.c, vocab.c, isMultilingual != 0U);
  return reinterpret_cast<jlong>(engine);
}

JNIEXPORT void JNICALL Java_io_github_jerinphilip_whisper_EngineNative_destroy(JNIEnv* /*env*/, jobject /*thiz*/,
                                                                              jlong nativePtr) {
  delete engine_of(nativePtr);
}

JNIEXPORT jstring JNICALL Java_io_github_jerinphilip_whisper_EngineNative_transcribeBuffer(JNIEnv* env,
                                                                                          jobject /*thiz*/,
                                                                                          jlong nativePtr,
                                                                                          jfloatArray samples) {
  const jsize len = env->GetArrayLength(samples);
  jfloat* data = env->GetFloatArrayElements(samples, nullptr);
  std::vector<float> pcm(data, data + len);
  env->ReleaseFloatArrayElements(samples, data, 0);
  return env->NewStringUTF(engine_of(nativePtr)->transcribe(pcm).c_str());
}

JNIEXPORT jstring JNICALL Java_io_github_jerinphilip_whisper_EngineNative_transcribeFile(JNIEnv* env,
                                                                                        jobject /*thiz*/,
                                                                                        jlong nativePtr,
                                                                                        jstring waveFile) {
  Utf path(env, waveFile);
  return env->NewStringUTF(engine_of(nativePtr)->transcribe(path.c).c_str());
}

}  // extern "C"
