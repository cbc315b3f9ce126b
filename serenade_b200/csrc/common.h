// Host-side helpers: error reporting across the C ABI, device buffers, TMA tensor-map encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace srn {

void set_last_error(const std::string& msg);

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

[[noreturn]] inline void fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  throw Error(buf);
}

#define SRN_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) ::srn::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void alloc(size_t n, bool zero = true) {
    release();
    bytes = n ? n : 16;
    SRN_CUDA(cudaMalloc(&p, bytes));
    if (zero) SRN_CUDA(cudaMemset(p, 0, bytes));
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency).
CUtensorMap make_tmap_3d(const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t batch,
                         uint64_t ld_elems, uint64_t bstride_elems, uint32_t box_inner, uint32_t box_rows,
                         bool swizzle128 = true);

}  // namespace srn
