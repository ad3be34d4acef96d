// Read-only memory map of a file (vocab+filters blob). Same surface as the reference's
// whisper::MmapFile (/root/reference/whisper.tflite/mmap_file.h:6-30): move-only RAII,
// data()/size(), std::runtime_error on open/stat/mmap failure (mmap_file.cpp:15-30).
#pragma once
#include <cstddef>
#include <string>

namespace whisper {

class MmapFile {
 public:
  MmapFile() = default;
  explicit MmapFile(const std::string& filepath);
  ~MmapFile();
  MmapFile(const MmapFile&) = delete;
  MmapFile& operator=(const MmapFile&) = delete;
  MmapFile(MmapFile&& from) noexcept;
  MmapFile& operator=(MmapFile&& from) noexcept;

  void* data() const { return data_; }
  size_t size() const { return size_; }

 private:
  void release() noexcept;
  int fd_ = -1;
  void* data_ = nullptr;
  size_t size_ = 0;
};

}  // namespace whisper
