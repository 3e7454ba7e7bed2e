// Shared host/device helpers for libdlas_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/dlas_b200.h"

namespace dlas {

void set_error(const char* fmt, ...);
int device_ok();    // DLAS_OK or DLAS_ERR_UNSUPPORTED_DEVICE (cached per device)
int sm_count();
bool prof_on();
int prof_begin(int slot, cudaStream_t st);
void prof_end(int idx, cudaStream_t st);

#define DLAS_CHECK_DEVICE()                     \
  do {                                          \
    int _rc = ::dlas::device_ok();              \
    if (_rc != DLAS_OK) return _rc;             \
  } while (0)

#define DLAS_CUDA(call)                                                                    \
  do {                                                                                     \
    cudaError_t _e = (call);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::dlas::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
      return DLAS_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

#define DLAS_REQUIRE(cond, ...)        \
  do {                                 \
    if (!(cond)) {                     \
      ::dlas::set_error(__VA_ARGS__);  \
      return DLAS_ERR_INVALID;         \
    }                                  \
  } while (0)

#define DLAS_LAUNCH_CHECK() DLAS_CUDA(cudaGetLastError())

// ---- counter hash shared by synth.cu (mirror of oracle/dlas_oracle.py splitmix64) -------------
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint64_t stream) {
  return splitmix64(seed ^ splitmix64(stream * 0xD1342543DE82EF95ull));
}
__host__ __device__ __forceinline__ uint32_t hash_u24(uint64_t key, uint64_t idx) {
  return (uint32_t)(splitmix64(key + idx) >> 40);
}

}  // namespace dlas
