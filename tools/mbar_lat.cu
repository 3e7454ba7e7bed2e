// Microbenchmark (sm_100a): cost of mbarrier waits.  (1) a wait on an already completed phase, (2) wake-up latency of a
// polling warp after another warp's arrive -- for try_wait vs test_wait loops, polled by 32 lanes or by one lane.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mbar_lat tools/mbar_lat.cu && ./mbar_lat
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool try_wait(uint32_t a, uint32_t par) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(a), "r"(par) : "memory");
  return ok;
}
__device__ __forceinline__ bool test_wait(uint32_t a, uint32_t par) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(a), "r"(par) : "memory");
  return ok;
}
__device__ __forceinline__ void arrive(uint32_t a) { asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(a) : "memory"); }

// mode: 0 try_wait all lanes, 1 try_wait lane 0, 2 test_wait all lanes, 3 test_wait lane 0
__global__ void k(int mode, int n_pollers, long long* out) {
  __shared__ uint64_t bar;
  __shared__ long long t_arrive;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a = smem_u32(&bar);
  if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(a)); }
  __syncthreads();
  long long sum_done = 0, sum_wake = 0;
  const int R = 64;
  for (int r = 0; r < R; ++r) {
    const uint32_t par = r & 1;
    if (warp == 0) {   // arriver: wait a while so the pollers are parked, then arrive
      long long t0 = clock64();
      while (clock64() - t0 < 3000) {}
      if (lane == 0) { t_arrive = clock64(); __threadfence_block(); arrive(a); }
    } else if (warp <= n_pollers) {
      const bool me = (mode & 1) ? lane == 0 : true;
      if (me) {
        if (mode < 2) while (!try_wait(a, par)) {}
        else while (!test_wait(a, par)) {}
      }
      __syncwarp();
      long long t1 = clock64();
      if (warp == 1 && lane == 0) sum_wake += t1 - *(volatile long long*)&t_arrive;
      // cost of a wait on the completed phase
      long long t2 = clock64();
      if (me) {
        if (mode < 2) while (!try_wait(a, par)) {}
        else while (!test_wait(a, par)) {}
      }
      __syncwarp();
      long long t3 = clock64();
      if (warp == 1 && lane == 0) sum_done += t3 - t2;
    }
    __syncthreads();
  }
  if (threadIdx.x == 32) { out[0] = sum_wake / R; out[1] = sum_done / R; }
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  const char* names[4] = {"try_wait, 32 lanes", "try_wait, lane 0  ", "test_wait, 32 lanes", "test_wait, lane 0  "};
  for (int np : {1, 8}) for (int m = 0; m < 4; ++m) {
    k<<<1, 32 * 9>>>(m, np, d);
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("%d polling warps, %s: wake-up after arrive %lld cycles, wait on a completed phase %lld cycles\n", np, names[m], h[0], h[1]);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
