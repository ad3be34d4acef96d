// Declaration-only stand-in (see c/c_api_types.h in this stub tree).
#pragma once
namespace tflite { namespace ops { namespace builtin {
class BuiltinOpResolver {};
}}}  // namespace tflite::ops::builtin
