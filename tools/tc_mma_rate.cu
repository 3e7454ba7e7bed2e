// Microbenchmark (not part of the library): cycles per tcgen05.mma (M=128, K=16, fp16) for the operand
// flavours and N the MLP kernel uses, issued back to back by one thread on one SM, accumulating in TMEM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/tc_mma_rate tools/tc_mma_rate.cu && /tmp/tc_mma_rate
#include <cuda_runtime.h>
#include <stdio.h>
#include "../deep-learning-at-scale_b200/csrc/tc_common.cuh"
using namespace dlas::tc;

template <int N, bool A_MN, bool B_MN, int REPS>
__global__ void __launch_bounds__(128) rate_kernel(long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u);
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(&slot, 512);
  fence_async_smem(); fence_before(); __syncthreads(); fence_after();
  const uint32_t tm = slot, sA = smem_u32(smem), sB = sA + 64 * 1024;
  if (warp == 0) {
    const bool leader = elect_one();
    const uint64_t a0 = A_MN ? desc_mnmajor(sA, 128, 0, 0) : desc_kmajor(sA, 128, 0, 0);
    const uint64_t b0 = B_MN ? desc_mnmajor(sB, 128, 0, 0) : desc_kmajor(sB, 256, 0, 0);
    constexpr uint32_t idesc = make_idesc(128, N, A_MN, B_MN);
    long long t0 = clock64();
    if (leader) {
#pragma unroll 1
      for (int r = 0; r < REPS; ++r) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          mma_f16(tm, a0 + (A_MN ? koff_mnmajor(i) : koff_kmajor(128, i)), b0 + (B_MN ? koff_mnmajor(i) : koff_kmajor(256, i)), idesc, 1);
      }
      mma_commit(&bar);
    }
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

template <int N, bool A_MN, bool B_MN>
void run(const char* name, long long* d) {
  constexpr int REPS = 64;
  auto k = rate_kernel<N, A_MN, B_MN, REPS>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  long long h[2];
  for (int it = 0; it < 2; ++it) { k<<<1, 128, 160 * 1024>>>(d); cudaDeviceSynchronize(); }
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-22s N=%3d a_mn=%d b_mn=%d  issue %.1f cyc/mma  complete %.1f cyc/mma  (floor N/2 = %d; bytes/mma = %d) %s\n", name, N, A_MN, B_MN,
         h[0] / (REPS * 8.0), h[1] / (REPS * 8.0), N / 2, 4096 + N * 32, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  run<64, false, true>("gemm1/2 (K,MN)", d);  run<128, false, true>("gemm1/2 (K,MN)", d);  run<256, false, true>("gemm1/2 (K,MN)", d);
  run<64, false, false>("dH1 (K,K)", d);      run<128, false, false>("dH1 (K,K)", d);      run<256, false, false>("dH1 (K,K)", d);
  run<128, true, true>("dW2 (MN,MN)", d);     run<256, true, true>("dW2 (MN,MN)", d);      run<48, true, true>("dW1 (MN,MN)", d);
  run<64, true, true>("dW (MN,MN)", d);       run<16, true, false>("db2 (MN,K)", d);
  return 0;
}
