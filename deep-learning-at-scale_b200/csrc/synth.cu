// Synthetic inputs generated in HBM: the reference's rows come out of a Spark CSV
// ingest (cnn.py:49-103) that is out of scope; its datasets were never published (Readme.md:25).
// Counter hash = splitmix64, bit-identical to oracle/dlas_oracle.py, so any slice of a
// device-generated shard can be regenerated on the host.  Also: the 64/16/20 split codes
// (cnn.py:68) and the lag-window gather of the dynamic ANN.
#include "common.cuh"

namespace dlas {
namespace {

constexpr int kThreads = 256;
constexpr uint64_t STREAM_GILBERT_P = 1, STREAM_GILBERT_S = 2, STREAM_GILBERT_R = 3;
constexpr uint64_t STREAM_X = 10, STREAM_NOISE = 11, STREAM_TEACHER = 12, STREAM_SPLIT = 20;

__device__ __forceinline__ float u01(uint32_t n24) { return (float)n24 * 5.9604644775390625e-8f; }

__global__ void __launch_bounds__(kThreads)
synth_gilbert_kernel(float* __restrict__ P, float* __restrict__ S, float* __restrict__ R,
                     uint64_t kP, uint64_t kS, uint64_t kR, uint64_t start, size_t n) {
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    uint64_t idx = start + i;
    P[i] = __fmaf_rn(4950.0f, u01(hash_u24(kP, idx)), 50.0f);
    S[i] = __fmaf_rn(124.0f, u01(hash_u24(kS, idx)), 4.0f);
    uint32_t nr = hash_u24(kR, idx) * 9u;  // < 2^28
    int j = (int)(nr >> 24);
    float v = (float)((nr & 0xFFFFFFu) >> 1) * 1.1920928955078125e-7f;  // 2^-23, exact
    R[i] = __fmul_rn(0.02f, ldexpf(1.0f + v, j));
  }
}

// One thread per row: X[row][k] = 2u-1, y = softsign(sum_k fma(x_k, w_k)) + 0.05*(2u-1).
__global__ void __launch_bounds__(kThreads)
synth_rows_kernel(float* __restrict__ X, float* __restrict__ y, uint64_t kX, uint64_t kN,
                  uint64_t kT, uint64_t start, size_t n_rows, int F) {
  extern __shared__ float wt[];
  for (int k = threadIdx.x; k < F; k += kThreads) wt[k] = 2.0f * u01(hash_u24(kT, (uint64_t)k)) - 1.0f;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n_rows; i += stride) {
    uint64_t row = start + i;
    float s = 0.0f;
    float* xr = X + i * (size_t)F;
    for (int k = 0; k < F; ++k) {
      float x = 2.0f * u01(hash_u24(kX, row * (uint64_t)F + (uint64_t)k)) - 1.0f;
      xr[k] = x;
      s = __fmaf_rn(x, wt[k], s);
    }
    float base = __fdiv_rn(s, 1.0f + fabsf(s));
    float noise = 2.0f * u01(hash_u24(kN, row)) - 1.0f;
    y[i] = __fmaf_rn(0.05f, noise, base);
  }
}

// One thread per sequence: AR(1)-smoothed noise, teacher on the last 8 steps.
__global__ void __launch_bounds__(kThreads)
synth_series_kernel(float* __restrict__ X, float* __restrict__ y, uint64_t kX, uint64_t kT,
                    uint64_t start, size_t n_seq, int T, int D) {
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n_seq; i += stride) {
    uint64_t seq = start + i;
    float s = 0.0f;
    float* xs = X + i * (size_t)T * D;
    for (int t = 0; t < T; ++t) {
      for (int k = 0; k < D; ++k) {
        float e = 2.0f * u01(hash_u24(kX, (seq * (uint64_t)T + (uint64_t)t) * (uint64_t)D + (uint64_t)k)) - 1.0f;
        float prev = t ? xs[(size_t)(t - 1) * D + k] : 0.0f;
        float x = __fmaf_rn(0.75f, prev, __fmul_rn(0.25f, e));
        xs[(size_t)t * D + k] = x;
      }
    }
    for (int t = T - 8 < 0 ? 0 : T - 8; t < T; ++t)
      for (int k = 0; k < D; ++k) {
        float w = 2.0f * u01(hash_u24(kT, (uint64_t)k)) - 1.0f;
        s = __fmaf_rn(xs[(size_t)t * D + k], w, s);
      }
    y[i] = __fdiv_rn(s, 1.0f + fabsf(s));
  }
}

__global__ void __launch_bounds__(kThreads)
split_kernel(int8_t* __restrict__ code, uint64_t key, uint64_t start, size_t n) {
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    uint32_t h = hash_u24(key, start + i) % 100u;
    code[i] = h < 64u ? 0 : (h < 80u ? 1 : 2);
  }
}

// out[i][l*F + k] = series[i + l][k]; one thread per output element, coalesced writes.
__global__ void __launch_bounds__(kThreads)
lag_windows_kernel(const float* __restrict__ series, float* __restrict__ out, size_t n_out, int F,
                   int n_lags) {
  const size_t W = (size_t)F * n_lags;
  const size_t total = n_out * W;
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t e = (size_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += stride) {
    size_t i = e / W, c = e - i * W;
    out[e] = series[i * F + c];  // windows of consecutive rows are contiguous in the row-major series
  }
}

unsigned grid_for(size_t n) {
  size_t need = (n + kThreads - 1) / kThreads;
  size_t cap = (size_t)sm_count() * 16;
  if (need < 1) need = 1;
  return (unsigned)(need < cap ? need : cap);
}

}  // namespace
}  // namespace dlas

using namespace dlas;

extern "C" int dlas_synth_gilbert_f32(float* P, float* S, float* R, uint64_t seed, uint64_t start,
                                      size_t n, void* stream) {
  DLAS_CHECK_DEVICE();
  if (n == 0) return DLAS_OK;
  DLAS_REQUIRE(P && S && R, "dlas_synth_gilbert_f32: null pointer");
  synth_gilbert_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(
      P, S, R, stream_key(seed, STREAM_GILBERT_P), stream_key(seed, STREAM_GILBERT_S),
      stream_key(seed, STREAM_GILBERT_R), start, n);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_synth_rows_f32(float* X, float* y, uint64_t seed, uint64_t start, size_t n_rows,
                                   int n_features, void* stream) {
  DLAS_CHECK_DEVICE();
  if (n_rows == 0) return DLAS_OK;
  DLAS_REQUIRE(X && y, "dlas_synth_rows_f32: null pointer");
  DLAS_REQUIRE(n_features > 0 && n_features <= 8192, "dlas_synth_rows_f32: n_features %d out of range", n_features);
  synth_rows_kernel<<<grid_for(n_rows), kThreads, n_features * sizeof(float), (cudaStream_t)stream>>>(
      X, y, stream_key(seed, STREAM_X), stream_key(seed, STREAM_NOISE), stream_key(seed, STREAM_TEACHER),
      start, n_rows, n_features);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_synth_series_f32(float* X, float* y, uint64_t seed, uint64_t start, size_t n_seq,
                                     int T, int d_in, void* stream) {
  DLAS_CHECK_DEVICE();
  if (n_seq == 0) return DLAS_OK;
  DLAS_REQUIRE(X && y, "dlas_synth_series_f32: null pointer");
  DLAS_REQUIRE(T > 0 && d_in > 0, "dlas_synth_series_f32: bad shape T=%d d_in=%d", T, d_in);
  synth_series_kernel<<<grid_for(n_seq), kThreads, 0, (cudaStream_t)stream>>>(
      X, y, stream_key(seed, STREAM_X), stream_key(seed, STREAM_TEACHER), start, n_seq, T, d_in);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_split_codes(int8_t* code, uint64_t seed, uint64_t start, size_t n, void* stream) {
  DLAS_CHECK_DEVICE();
  if (n == 0) return DLAS_OK;
  DLAS_REQUIRE(code, "dlas_split_codes: null pointer");
  split_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(code, stream_key(seed, STREAM_SPLIT), start, n);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_lag_windows_f32(const float* series, size_t n_rows, int n_features, int n_lags,
                                    float* out, void* stream) {
  DLAS_CHECK_DEVICE();
  DLAS_REQUIRE(n_features > 0 && n_lags > 0, "dlas_lag_windows_f32: bad shape");
  if (n_rows < (size_t)n_lags) return DLAS_OK;
  DLAS_REQUIRE(series && out, "dlas_lag_windows_f32: null pointer");
  size_t n_out = n_rows - n_lags + 1;
  lag_windows_kernel<<<grid_for(n_out * n_features * n_lags), kThreads, 0, (cudaStream_t)stream>>>(
      series, out, n_out, n_features, n_lags);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}
