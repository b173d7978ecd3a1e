// Shared helpers for libegpt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/egpt_b200.h"

namespace egpt {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

// optional per-kernel-family event timing (egpt_profile_enable / egpt_profile_collect)
enum { PROF_GEMM = 0, PROF_ATTN = 1, PROF_ELEMENTWISE = 2, PROF_RASTER = 3, PROF_PREPROCESS = 4, PROF_POOL = 5,
       PROF_SPLIT = 6 };
void prof_begin(int category, double work, cudaStream_t stream);
void prof_end(cudaStream_t stream);
struct ProfScope {
  cudaStream_t st;
  ProfScope(int category, double work, cudaStream_t s) : st(s) { prof_begin(category, work, s); }
  ~ProfScope() { prof_end(st); }
};

#define EGPT_CHECK_ARG(cond, ...)                      \
  do {                                                 \
    if (!(cond)) {                                     \
      ::egpt::set_error(__VA_ARGS__);                  \
      return EGPT_ERR_INVALID;                         \
    }                                                  \
  } while (0)

#define EGPT_CUDA(call)                                                                     \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      ::egpt::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,  \
                        __LINE__);                                                          \
      return EGPT_ERR_CUDA;                                                                 \
    }                                                                                       \
  } while (0)

#define EGPT_LAUNCH_CHECK()                                                                 \
  do {                                                                                      \
    ::egpt::count_launch();                                                                 \
    cudaError_t e__ = cudaGetLastError();                                                   \
    if (e__ != cudaSuccess) {                                                               \
      ::egpt::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__),        \
                        __FILE__, __LINE__);                                                \
      return EGPT_ERR_CUDA;                                                                 \
    }                                                                                       \
  } while (0)

#define EGPT_TRY(call)              \
  do {                              \
    int rc__ = (call);              \
    if (rc__ != EGPT_OK) return rc__; \
  } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__host__ __device__ inline int dtype_size(int dt) {
  switch (dt) {
    case EGPT_U8: case EGPT_I8: return 1;
    case EGPT_U16: case EGPT_I16: return 2;
    case EGPT_U32: case EGPT_I32: return 4;
    case EGPT_U64: case EGPT_I64: return 8;
  }
  return 0;
}

// bump allocator over the caller's workspace
struct Arena {
  char* base;
  size_t cap, off;
  Arena(void* p, size_t n) : base((char*)p), cap(n), off(0) {}
  void* take(size_t bytes) {
    size_t o = align_up(off, 1024);
    if (o + bytes > cap) { off = cap + 1; return nullptr; }
    off = o + bytes;
    return base + o;
  }
  bool ok() const { return off <= cap; }
};

// generic integer load -> int64 for the event arrays
__device__ __forceinline__ long long load_int(const void* p, int dt, long long i) {
  switch (dt) {
    case EGPT_U8: return ((const uint8_t*)p)[i];
    case EGPT_I8: return ((const int8_t*)p)[i];
    case EGPT_U16: return ((const uint16_t*)p)[i];
    case EGPT_I16: return ((const int16_t*)p)[i];
    case EGPT_U32: return ((const uint32_t*)p)[i];
    case EGPT_I32: return ((const int32_t*)p)[i];
    case EGPT_U64: return (long long)((const uint64_t*)p)[i];
    default: return ((const long long*)p)[i];
  }
}

constexpr int kNumSMs = 148;

}  // namespace egpt
