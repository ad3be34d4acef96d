// Declaration-only stand-in (see c/c_api_types.h in this stub tree). The real header pulls
// FlatBufferModel / InterpreterBuilder in transitively, so they are declared here too.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>
#include "tensorflow/lite/c/c_api_types.h"
#include "tensorflow/lite/kernels/register.h"
namespace tflite {
[[noreturn]] inline void stub_unavailable(const char* what) {
  std::fprintf(stderr, "tflite stub: %s called, but TFLite is unavailable on this box\n", what);
  std::abort();
}
class Interpreter {
 public:
  TfLiteStatus AllocateTensors() { stub_unavailable("AllocateTensors"); }
  TfLiteStatus Invoke() { stub_unavailable("Invoke"); }
  TfLiteStatus ResizeInputTensor(int, const std::vector<int>&) { stub_unavailable("ResizeInputTensor"); }
  TfLiteStatus SetNumThreads(int) { stub_unavailable("SetNumThreads"); }
  template <class T> T* typed_input_tensor(int) { stub_unavailable("typed_input_tensor"); }
  template <class T> T* typed_output_tensor(int) { stub_unavailable("typed_output_tensor"); }
  TfLiteTensor* input_tensor(size_t) { stub_unavailable("input_tensor"); }
  TfLiteTensor* output_tensor(size_t) { stub_unavailable("output_tensor"); }
  TfLiteTensor* tensor(int) { stub_unavailable("tensor"); }
  const std::vector<int>& outputs() const { stub_unavailable("outputs"); }
};
class FlatBufferModel {
 public:
  static std::unique_ptr<FlatBufferModel> BuildFromFile(const char*) { stub_unavailable("BuildFromFile"); }
};
class InterpreterBuilder {
 public:
  InterpreterBuilder(const FlatBufferModel&, const ops::builtin::BuiltinOpResolver&) {}
  TfLiteStatus operator()(std::unique_ptr<Interpreter>*) { stub_unavailable("InterpreterBuilder"); }
};
}  // namespace tflite
