// Minimal declaration-only stand-in for tensorflow/lite/c/c_api_types.h.
// TEST INFRASTRUCTURE: exists only so the reference's own whisper.cpp can be
// compiled where it lies (/root/reference) for the mel/Reader/decode oracle.
// The TFLite interpreter itself is absent on this box; any call into it aborts.
#pragma once
#include <cstddef>
#include <cstdint>
typedef enum {
  kTfLiteNoType = 0, kTfLiteFloat32 = 1, kTfLiteInt32 = 2, kTfLiteUInt8 = 3,
  kTfLiteInt64 = 4, kTfLiteString = 5, kTfLiteBool = 6, kTfLiteInt16 = 7,
  kTfLiteComplex64 = 8, kTfLiteInt8 = 9, kTfLiteFloat16 = 10, kTfLiteFloat64 = 11,
  kTfLiteComplex128 = 12, kTfLiteUInt64 = 13, kTfLiteResource = 14,
  kTfLiteVariant = 15, kTfLiteUInt32 = 16, kTfLiteUInt16 = 17, kTfLiteInt4 = 18,
  kTfLiteBFloat16 = 19,
} TfLiteType;
typedef enum { kTfLiteOk = 0, kTfLiteError = 1 } TfLiteStatus;
typedef struct TfLiteIntArray { int size; int data[8]; } TfLiteIntArray;
typedef struct TfLiteTensor {
  TfLiteType type; void* data_raw; TfLiteIntArray* dims; size_t bytes; const char* name;
} TfLiteTensor;
