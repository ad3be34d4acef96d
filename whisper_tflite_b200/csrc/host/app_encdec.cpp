// app/encdec of the reference (/root/reference/app/encdec.cpp:24-56) on the B200 engine:
//   encdec --model-prefix <prefix> --vocab <filters_vocab_multilingual.bin> --input <wav>
// loads <prefix>.b200w (in place of <prefix>.encoder.tflite + <prefix>.decoder.tflite), multilingual = true,
// prints the transcript. The three required string flags are parsed by hand (the reference pulls in CLI11 for them).
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>

#include "whisper.h"

namespace {
struct Options {
  std::string model_prefix, vocab, input;
};

bool parse(int argc, char* argv[], Options& o) {
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i], v;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) {
      v = a.substr(eq + 1);
      a = a.substr(0, eq);
    } else if (i + 1 < argc) {
      v = argv[++i];
    } else {
      return false;
    }
    if (a == "--model-prefix") o.model_prefix = v;
    else if (a == "--vocab") o.vocab = v;
    else if (a == "--input") o.input = v;
    else return false;
  }
  return !o.model_prefix.empty() && !o.vocab.empty() && !o.input.empty();
}
}  // namespace

int main(int argc, char* argv[]) {
  Options options;
  if (!parse(argc, argv, options)) {
    std::fprintf(stderr, "Usage: encdec --model-prefix <prefix> --vocab <vocab.bin> --input <wav>\n");
    return 106;  // CLI11's RequiredError exit code
  }
  using namespace whisper;  // NOLINT
  const bool multilingual = true;
  EncDec encdec(options.model_prefix, options.vocab, multilingual);
  const std::string text = encdec.transcribe(options.input.c_str());
  std::cout << text << "\n";
  return 0;
}
