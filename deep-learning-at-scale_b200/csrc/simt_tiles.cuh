// Shared-memory tile helpers of the fp32 CUDA-core kernels (MLP and LSTM).  Tiles are stored
// TRANSPOSED: tile[feature][RS] with the R rows of the batch tile contiguous (RS >= R, a multiple
// of 4 so float4 row loads stay aligned, chosen so that column reads spread over banks).
#pragma once
#include "common.cuh"

namespace dlas {

// out(j, r0..r0+RT) = sum_k inT[k][r] * Wm[k*ldw + j]; one (column j, row group) item per thread.
template <int R, int RS, int NT, int RT, class Epi>
__device__ __forceinline__ void gemm_rows(const float* __restrict__ inT, const float* __restrict__ Wm,
                                          int ldw, int K, int N, Epi epi) {
  constexpr int G = R / RT;
  const int items = N * G;
  for (int item = threadIdx.x; item < items; item += NT) {
    const int j = item % N, r0 = (item / N) * RT;
    float acc[RT];
#pragma unroll
    for (int i = 0; i < RT; ++i) acc[i] = 0.0f;
    const float* a = inT + r0;
    const float* w = Wm + j;
    int k = 0;
    for (; k + 4 <= K; k += 4) {
      float wv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) wv[u] = __ldg(w + (size_t)(k + u) * ldw);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* ar = a + (k + u) * RS;
        if (RT >= 4) {
#pragma unroll
          for (int i = 0; i < RT; i += 4) {
            float4 v = *reinterpret_cast<const float4*>(ar + i);
            acc[i] = fmaf(v.x, wv[u], acc[i]);
            acc[i + 1] = fmaf(v.y, wv[u], acc[i + 1]);
            acc[i + 2] = fmaf(v.z, wv[u], acc[i + 2]);
            acc[i + 3] = fmaf(v.w, wv[u], acc[i + 3]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < RT; ++i) acc[i] = fmaf(ar[i], wv[u], acc[i]);
        }
      }
    }
    for (; k < K; ++k) {
      float wv = __ldg(w + (size_t)k * ldw);
      const float* ar = a + k * RS;
#pragma unroll
      for (int i = 0; i < RT; ++i) acc[i] = fmaf(ar[i], wv, acc[i]);
    }
    epi(j, r0, acc);
  }
}

template <int R, int RS, int NT, class Epi>
__device__ __forceinline__ void gemm_dispatch(const float* inT, const float* Wm, int ldw, int K, int N, Epi epi) {
  // as many row groups as the thread count allows: RT = R / min(R, NT / N) rounded to a power of 2
  int g = NT / N;
  if (g > R) g = R;
  if (g < 1) g = 1;
  // rows per thread RT = R / g, g = number of row groups (a power of two <= R)
  if (g >= R) gemm_rows<R, RS, NT, 1>(inT, Wm, ldw, K, N, epi);
  else if (2 * g >= R) gemm_rows<R, RS, NT, (R >= 2 ? 2 : 1)>(inT, Wm, ldw, K, N, epi);
  else if (4 * g >= R) gemm_rows<R, RS, NT, (R >= 4 ? 4 : 1)>(inT, Wm, ldw, K, N, epi);
  else if (8 * g >= R) gemm_rows<R, RS, NT, (R >= 8 ? 8 : 1)>(inT, Wm, ldw, K, N, epi);
  else if (16 * g >= R) gemm_rows<R, RS, NT, (R >= 16 ? 16 : 1)>(inT, Wm, ldw, K, N, epi);
  else gemm_rows<R, RS, NT, (R >= 32 ? 32 : 1)>(inT, Wm, ldw, K, N, epi);
}

// dW[k][j] (+)= sum_r aT[k][r] * dzT[j][r];  db[j] (+)= sum_r dzT[j][r]
template <int R, int RS, int NT>
__device__ __forceinline__ void accumulate_dw(const float* __restrict__ aT, const float* __restrict__ dzT,
                                              int K, int N, float* __restrict__ pw, float* __restrict__ pb,
                                              bool first) {
  int kgn = NT / N;
  if (kgn < 1) kgn = 1;
  if (kgn > K) kgn = K;
  const int items = N * kgn;
  for (int item = threadIdx.x; item < items; item += NT) {
    const int j = item % N, kg = item / N;
    float dz[R];
#pragma unroll
    for (int i = 0; i < R; i += 4) {
      float4 v = *reinterpret_cast<const float4*>(dzT + j * RS + i);
      dz[i] = v.x; dz[i + 1] = v.y; dz[i + 2] = v.z; dz[i + 3] = v.w;
    }
    if (kg == 0) {
      float s = 0.0f;
#pragma unroll
      for (int i = 0; i < R; ++i) s += dz[i];
      pb[j] = first ? s : pb[j] + s;
    }
    for (int k = kg; k < K; k += kgn) {
      const float* ar = aT + k * RS;
      float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
#pragma unroll
      for (int i = 0; i < R; i += 4) {
        float4 v = *reinterpret_cast<const float4*>(ar + i);
        s0 = fmaf(v.x, dz[i], s0);
        s1 = fmaf(v.y, dz[i + 1], s1);
        s2 = fmaf(v.z, dz[i + 2], s2);
        s3 = fmaf(v.w, dz[i + 3], s3);
      }
      float s = (s0 + s1) + (s2 + s3);
      float* p = pw + (size_t)k * N + j;
      *p = first ? s : *p + s;
    }
  }
}


}  // namespace dlas
