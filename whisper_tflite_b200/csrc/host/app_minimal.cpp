// app/minimal of the reference (/root/reference/app/minimal.cpp:21-46) on the B200 engine:
//   minimal <model prefix> <vocab file> <wav file>      (multilingual = false, Monolith engine)
#include <cstdio>
#include <string>

#include "whisper.h"

int main(int argc, char* argv[]) {
  if (argc != 4) {
    fprintf(stderr, "Usage: minimal <model prefix> <vocab file> <pcm_file name>\n");
    return 1;
  }
  using namespace whisper;  // NOLINT
  const bool multilingual = false;
  Monolith monolith(argv[1], argv[2], multilingual);
  std::string text = monolith.transcribe(argv[3]);
  text = remove_extra_spaces(text);
  printf("\n%s\n\n", text.c_str());
  return 0;
}
