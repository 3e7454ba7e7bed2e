// Tuning harness (not part of the library): times launch configurations of the Gilbert streaming
// kernel -- threads, float4 groups per thread, load policy, looping vs one-shot grid -- with and
// without the arithmetic, to separate the memory pattern from the instruction budget.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/gilbert_tune tools/gilbert_tune.cu && /tmp/gilbert_tune
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
#define GILBERT_TUNE 1
#include "../deep-learning-at-scale_b200/csrc/gilbert.cu"

using namespace dlas;

template <int POL> __device__ __forceinline__ float4 ld4(const float4* p) {
  if (POL == 0) return __ldg(p);
  if (POL == 1) return __ldcs(p);
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
template <int POL> __device__ __forceinline__ void st4(float4* p, float4 v) {
  if (POL == 0) *p = v;
  else if (POL == 1) __stcs(p, v);
  else asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
}

template <int TH, int U, int POL, int MATH, bool LOOP>
__global__ void __launch_bounds__(TH) k(const float4* __restrict__ P, const float4* __restrict__ S,
                                        const float4* __restrict__ R, float4* __restrict__ q, size_t n4) {
  const size_t stride = LOOP ? (size_t)gridDim.x * TH * U : n4;
  for (size_t base = (size_t)blockIdx.x * TH * U + threadIdx.x; base < n4; base += stride) {
    float4 p[U], s[U], r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      size_t i = base + (size_t)u * TH;
      if (i < n4) { p[u] = ld4<POL>(P + i); s[u] = ld4<POL>(S + i); r[u] = ld4<POL>(R + i); }
      else p[u] = s[u] = r[u] = make_float4(1, 1, 1, 1);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      size_t i = base + (size_t)u * TH;
      float4 o;
      if (MATH < 0) { o.x = p[u].x + s[u].x + r[u].x; o.y = p[u].y + s[u].y + r[u].y; o.z = p[u].z + s[u].z + r[u].z; o.w = p[u].w + s[u].w + r[u].w; }
      else {
        float n0, n1, n2, n3;
        o.x = gilbert_fast<(MATH < 0 ? 0 : MATH)>(p[u].x, s[u].x, r[u].x, n0);
        o.y = gilbert_fast<(MATH < 0 ? 0 : MATH)>(p[u].y, s[u].y, r[u].y, n1);
        o.z = gilbert_fast<(MATH < 0 ? 0 : MATH)>(p[u].z, s[u].z, r[u].z, n2);
        o.w = gilbert_fast<(MATH < 0 ? 0 : MATH)>(p[u].w, s[u].w, r[u].w, n3);
        float nm = fmaxf(fmaxf(fabsf(n0), fabsf(n1)), fmaxf(fabsf(n2), fabsf(n3)));
        if (nm > 125.0f) o.x = 0.f;
      }
      if (i < n4) st4<POL>(q + i, o);
    }
    if (!LOOP) break;
  }
}

struct f8 { float v[8]; };
template <int EV> __device__ __forceinline__ f8 ld8(const float* p) {
  f8 r; uint32_t* u = reinterpret_cast<uint32_t*>(r.v);
  if (EV) asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]) : "l"(p));
  else asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]) : "l"(p));
  return r;
}
template <int EV> __device__ __forceinline__ void st8(float* p, const f8& r) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(r.v);
  if (EV) asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]));
  else asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]));
}
// 256-bit loads: one f8 group (8 rows) per array per thread
template <int TH, int EV, int MATH>
__global__ void __launch_bounds__(TH) k8(const float* __restrict__ P, const float* __restrict__ S,
                                         const float* __restrict__ R, float* __restrict__ q, size_t n8) {
  size_t i = (size_t)blockIdx.x * TH + threadIdx.x;
  if (i >= n8) return;
  f8 p = ld8<EV>(P + i * 8), s = ld8<EV>(S + i * 8), r = ld8<EV>(R + i * 8), o;
  float nm = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (MATH < 0) o.v[j] = p.v[j] + s.v[j] + r.v[j];
    else { float n; o.v[j] = gilbert_fast<(MATH < 0 ? 0 : MATH)>(p.v[j], s.v[j], r.v[j], n); nm = fmaxf(nm, fabsf(n)); }
  }
  if (nm > 125.0f) o.v[0] = 0.f;
  st8<EV>(q + i * 8, o);
}
template <int TH, int EV, int MATH>
void run8(const float* P, const float* S, const float* R, float* q, size_t n) {
  size_t n8 = n / 8;
  unsigned grid = (unsigned)((n8 + TH - 1) / TH);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) k8<TH, EV, MATH><<<grid, TH>>>(P, S, R, q, n8);
  cudaEventRecord(a);
  for (int i = 0; i < 10; ++i) k8<TH, EV, MATH><<<grid, TH>>>(P, S, R, q, n8);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
  printf("k8 256-bit TH=%d EV=%d MATH=%d grid=%u  %.4f ms  %.0f GB/s  %s\n", TH, EV, MATH, grid, ms, n * 16.0 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

__global__ void copyk(const float4* __restrict__ a, float4* __restrict__ b, size_t n4) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n4) b[i] = a[i];
}

template <int TH, int U, int POL, int MATH, bool LOOP>
void run(const char* name, const float4* P, const float4* S, const float4* R, float4* q, size_t n4, int sms, int gm) {
  size_t need = (n4 + (size_t)TH * U - 1) / ((size_t)TH * U);
  unsigned grid = LOOP ? (unsigned)std::min<size_t>(need, (size_t)sms * gm) : (unsigned)need;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) k<TH, U, POL, MATH, LOOP><<<grid, TH>>>(P, S, R, q, n4);
  cudaEventRecord(a);
  const int reps = 10;
  for (int i = 0; i < reps; ++i) k<TH, U, POL, MATH, LOOP><<<grid, TH>>>(P, S, R, q, n4);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); ms /= reps;
  printf("%-34s TH=%d U=%d POL=%d MATH=%d LOOP=%d gm=%d grid=%u  %.4f ms  %.0f GB/s  %s\n", name, TH, U, POL, MATH, (int)LOOP, gm, grid, ms,
         n4 * 64.0 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  size_t n = 200000000, n4 = n / 4;
  float *P, *S, *R, *q;
  cudaMalloc(&P, n * 4); cudaMalloc(&S, n * 4); cudaMalloc(&R, n * 4); cudaMalloc(&q, n * 4);
  std::vector<float> h(1 << 20);
  for (size_t i = 0; i < h.size(); ++i) h[i] = 50.0f + (float)(i % 4000);
  for (size_t o = 0; o < n; o += h.size()) { size_t c = std::min(h.size(), n - o); cudaMemcpy(P + o, h.data(), c * 4, cudaMemcpyHostToDevice); cudaMemcpy(S + o, h.data(), c * 4, cudaMemcpyHostToDevice); cudaMemcpy(R + o, h.data(), c * 4, cudaMemcpyHostToDevice); }
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const float4 *P4 = (const float4*)P, *S4 = (const float4*)S, *R4 = (const float4*)R; float4* q4 = (float4*)q;
  { // plain copy reference: 2 x 800 MB
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) copyk<<<(unsigned)((n4 + 255) / 256), 256>>>(P4, q4, n4);
    cudaEventRecord(a); for (int i = 0; i < 10; ++i) copyk<<<(unsigned)((n4 + 255) / 256), 256>>>(P4, q4, n4);
    cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
    printf("copy 1r1w: %.4f ms %.0f GB/s\n", ms, n4 * 32.0 / ms / 1e6);
  }
#define RUN(TH, U, POL, MATH, LOOP, GM) run<TH, U, POL, MATH, LOOP>(#TH "," #U "," #POL "," #MATH "," #LOOP, P4, S4, R4, q4, n4, sms, GM)
  // memory pattern only (MATH=-1)
  RUN(256, 1, 0, -1, false, 0); RUN(256, 2, 0, -1, false, 0); RUN(256, 4, 0, -1, false, 0);
  RUN(128, 1, 0, -1, false, 0); RUN(128, 2, 0, -1, false, 0); RUN(512, 1, 0, -1, false, 0); RUN(512, 2, 0, -1, false, 0);
  RUN(256, 1, 1, -1, false, 0); RUN(256, 2, 1, -1, false, 0); RUN(256, 1, 2, -1, false, 0); RUN(256, 2, 2, -1, false, 0);
  RUN(256, 2, 0, -1, true, 8); RUN(256, 2, 0, -1, true, 16); RUN(256, 2, 0, -1, true, 32); RUN(256, 2, 1, -1, true, 16);
  RUN(256, 4, 0, -1, true, 8); RUN(512, 2, 0, -1, true, 4); RUN(1024, 1, 0, -1, false, 0); RUN(1024, 1, 0, -1, true, 2);
  // with math
  RUN(256, 1, 0, 0, false, 0); RUN(256, 2, 0, 0, false, 0); RUN(256, 4, 0, 0, false, 0); RUN(128, 2, 0, 0, false, 0);
  RUN(512, 1, 0, 0, false, 0); RUN(256, 2, 1, 0, false, 0); RUN(256, 2, 2, 0, false, 0); RUN(256, 2, 0, 0, true, 16);
  RUN(256, 1, 0, 7, false, 0); RUN(256, 2, 0, 7, false, 0); RUN(256, 4, 0, 7, false, 0); RUN(256, 2, 0, 5, false, 0);
  RUN(256, 2, 1, 7, false, 0); RUN(256, 2, 0, 7, true, 16); RUN(512, 2, 0, 7, false, 0); RUN(128, 4, 0, 7, false, 0);
  run8<256, 0, -1>(P, S, R, q, n); run8<256, 1, -1>(P, S, R, q, n); run8<128, 1, -1>(P, S, R, q, n); run8<512, 1, -1>(P, S, R, q, n);
  run8<256, 0, 0>(P, S, R, q, n); run8<256, 1, 0>(P, S, R, q, n); run8<128, 1, 0>(P, S, R, q, n);
  run8<256, 0, 7>(P, S, R, q, n); run8<256, 1, 7>(P, S, R, q, n); run8<128, 1, 7>(P, S, R, q, n); run8<512, 1, 7>(P, S, R, q, n);
  return 0;
}
