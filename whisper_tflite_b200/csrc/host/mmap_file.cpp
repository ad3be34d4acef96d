#include "mmap_file.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <stdexcept>
#include <utility>

namespace whisper {

MmapFile::MmapFile(const std::string& filepath) {
  fd_ = ::open(filepath.c_str(), O_RDONLY);
  if (fd_ < 0) throw std::runtime_error("Failed to open file: " + filepath);
  struct stat st {};
  if (::fstat(fd_, &st) != 0) {
    ::close(fd_);
    fd_ = -1;
    throw std::runtime_error("Failed to get file size: " + filepath);
  }
  size_ = static_cast<size_t>(st.st_size);
  void* p = ::mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
  if (p == MAP_FAILED) {
    ::close(fd_);
    fd_ = -1;
    size_ = 0;
    throw std::runtime_error("Failed to mmap file: " + filepath);
  }
  data_ = p;
}

void MmapFile::release() noexcept {
  if (data_) ::munmap(data_, size_);
  if (fd_ >= 0) ::close(fd_);
  fd_ = -1;
  data_ = nullptr;
  size_ = 0;
}

MmapFile::~MmapFile() { release(); }

MmapFile::MmapFile(MmapFile&& from) noexcept
    : fd_(std::exchange(from.fd_, -1)), data_(std::exchange(from.data_, nullptr)), size_(std::exchange(from.size_, 0)) {}

MmapFile& MmapFile::operator=(MmapFile&& from) noexcept {
  if (this != &from) {
    release();
    fd_ = std::exchange(from.fd_, -1);
    data_ = std::exchange(from.data_, nullptr);
    size_ = std::exchange(from.size_, 0);
  }
  return *this;
}

}  // namespace whisper
