// Probe (not part of the library): checks every tcgen05.mma operand flavour the MLP kernel uses --
// K-major / MN-major views of the column-blocked SWIZZLE_128B fp16 tile layout, the instruction
// descriptor, TMEM accumulate and tcgen05.ld -- against a host reference.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/tc_probe tools/tc_probe.cu && /tmp/tc_probe
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <math.h>
#include "../deep-learning-at-scale_b200/csrc/tc_common.cuh"

using namespace dlas::tc;

struct Probe {
  int a_mn, b_mn;          // operand majors
  int a_rows, a_cols;      // physical tile dims (rows x cols) of A's tile
  int b_rows, b_cols;
  int a_off, b_off;        // A: row0 (K-major) or mn0 (MN-major); B likewise
  int K, N;                // M = 128
};

// tileA/tileB: logical row-major fp16 matrices of the PHYSICAL tile shape
__global__ void __launch_bounds__(128) probe_kernel(Probe p, const __half* __restrict__ tileA,
                                                    const __half* __restrict__ tileB, float* __restrict__ D,
                                                    int* __restrict__ err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  uint8_t* sA = smem;
  uint8_t* sB = smem + p.a_rows * ((p.a_cols + 63) / 64) * 128;
  const int tid = threadIdx.x, warp = tid >> 5;
  // fill tiles (8-element chunks)
  for (int idx = tid; idx < p.a_rows * (p.a_cols / 8); idx += 128) {
    int r = idx / (p.a_cols / 8), c = (idx % (p.a_cols / 8)) * 8;
    *reinterpret_cast<uint4*>(sA + tile_chunk_off(p.a_rows, r, c)) = *reinterpret_cast<const uint4*>(tileA + (size_t)r * p.a_cols + c);
  }
  for (int idx = tid; idx < p.b_rows * (p.b_cols / 8); idx += 128) {
    int r = idx / (p.b_cols / 8), c = (idx % (p.b_cols / 8)) * 8;
    *reinterpret_cast<uint4*>(sB + tile_chunk_off(p.b_rows, r, c)) = *reinterpret_cast<const uint4*>(tileB + (size_t)r * p.b_cols + c);
  }
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(&tmem_base, 256);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tm = tmem_base;
  if (tid == 0) {
    const uint32_t idesc = make_idesc(128, p.N, p.a_mn, p.b_mn);
    const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
    for (int k = 0; k < p.K; k += 16) {
      uint64_t ad = p.a_mn ? desc_mnmajor(a_base, p.a_rows, p.a_off, k) : desc_kmajor(a_base, p.a_rows, p.a_off, k);
      uint64_t bd = p.b_mn ? desc_mnmajor(b_base, p.b_rows, p.b_off, k) : desc_kmajor(b_base, p.b_rows, p.b_off, k);
      mma_f16(tm, ad, bd, idesc, k > 0);
    }
    mma_commit(&bar);
  }
  if (!mbar_wait(&bar, 0)) { if (tid == 0) *err = 1; }
  fence_after();
  for (int c0 = 0; c0 < p.N; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tm + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) D[(size_t)tid * p.N + c0 + j] = __uint_as_float(v[j]);
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 256);
}

static float hv(int i) { return (float)((i * 37 + 11) % 17 - 8) / 8.0f; }

int run(const char* name, Probe p) {
  // logical A[128][K], B[K][N]
  std::vector<float> A(128 * p.K), B((size_t)p.K * p.N);
  for (size_t i = 0; i < A.size(); ++i) A[i] = hv((int)i * 3 + 1);
  for (size_t i = 0; i < B.size(); ++i) B[i] = hv((int)i * 5 + 2);
  std::vector<__half> tA((size_t)p.a_rows * p.a_cols, __float2half(0.f)), tB((size_t)p.b_rows * p.b_cols, __float2half(0.f));
  // junk elsewhere so that wrong offsets show up
  for (size_t i = 0; i < tA.size(); ++i) tA[i] = __float2half(hv((int)i + 7) * 3.0f);
  for (size_t i = 0; i < tB.size(); ++i) tB[i] = __float2half(hv((int)i + 9) * 3.0f);
  for (int m = 0; m < 128; ++m)
    for (int k = 0; k < p.K; ++k) {
      if (p.a_mn) tA[(size_t)k * p.a_cols + p.a_off + m] = __float2half(A[m * p.K + k]);      // rows = K, cols = M
      else tA[(size_t)(p.a_off + m) * p.a_cols + k] = __float2half(A[m * p.K + k]);            // rows = M, cols = K
    }
  for (int k = 0; k < p.K; ++k)
    for (int n = 0; n < p.N; ++n) {
      if (p.b_mn) tB[(size_t)k * p.b_cols + p.b_off + n] = __float2half(B[(size_t)k * p.N + n]);  // rows = K, cols = N
      else tB[(size_t)(p.b_off + n) * p.b_cols + k] = __float2half(B[(size_t)k * p.N + n]);        // rows = N, cols = K
    }
  __half *dA, *dB; float* dD; int* dErr;
  cudaMalloc(&dA, tA.size() * 2); cudaMalloc(&dB, tB.size() * 2); cudaMalloc(&dD, 128 * p.N * 4); cudaMalloc(&dErr, 4);
  cudaMemcpy(dA, tA.data(), tA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, tB.data(), tB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, 128 * p.N * 4); cudaMemset(dErr, 0, 4);
  size_t smem = (size_t)p.a_rows * ((p.a_cols + 63) / 64) * 128 + (size_t)p.b_rows * ((p.b_cols + 63) / 64) * 128 + 1024;
  cudaError_t ea = cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); if (ea != cudaSuccess) printf("attr: %s\n", cudaGetErrorString(ea));
  probe_kernel<<<1, 128, smem>>>(p, dA, dB, dD, dErr);
  cudaError_t el = cudaGetLastError(); if (el != cudaSuccess) printf("launch: %s\n", cudaGetErrorString(el));
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> D(128 * p.N); int herr = 0;
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost); cudaMemcpy(&herr, dErr, 4, cudaMemcpyDeviceToHost);
  double maxerr = 0; int bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < p.N; ++n) {
      double s = 0;
      for (int k = 0; k < p.K; ++k) s += (double)A[m * p.K + k] * B[(size_t)k * p.N + n];
      double d = fabs(s - D[(size_t)m * p.N + n]);
      if (d > maxerr) maxerr = d;
      if (d > 1e-3) ++bad;
    }
  printf("%-28s a_mn=%d b_mn=%d K=%d N=%d  cuda=%s timeout=%d  max|err|=%.4g bad=%d/%d  %s\n", name, p.a_mn, p.b_mn, p.K, p.N,
         cudaGetErrorString(e), herr, maxerr, bad, 128 * p.N, (e == cudaSuccess && !herr && bad == 0) ? "OK" : "FAIL");
  if (bad && e == cudaSuccess) {
    printf("   D[0][0..7] ="); for (int n = 0; n < 8; ++n) printf(" %.3f", D[n]);
    printf("\n");
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dErr);
  return (e == cudaSuccess && !herr && bad == 0) ? 0 : 1;
}

int main() {
  int fails = 0;
  //                          a_mn b_mn a_rows a_cols b_rows b_cols a_off b_off  K    N
  fails += run("gemm1 (X.W1 chunk b)",   Probe{0, 1, 128, 64, 32, 256, 0, 128, 32, 128});
  fails += run("gemm1 N=256",            Probe{0, 1, 128, 64, 32, 256, 0, 0, 32, 256});
  fails += run("gemm2 (H1.W2h)",         Probe{0, 1, 128, 256, 256, 128, 0, 0, 256, 128});
  fails += run("dW2 (H1^T.dZ2h) M-hi",   Probe{1, 1, 128, 256, 128, 128, 128, 0, 128, 128});
  fails += run("dH1 (dZ2h.W2h^T) N-hi",  Probe{0, 0, 128, 128, 256, 128, 0, 128, 128, 128});
  fails += run("dW1 (dZ1^T.Xaug) N=48",  Probe{1, 1, 128, 256, 128, 64, 0, 0, 128, 48});
  fails += run("db2 (dZ2^T.ones) N=16",  Probe{1, 0, 128, 128, 16, 64, 0, 0, 64, 16});
  printf("%d failures\n", fails);
  return fails;
}
