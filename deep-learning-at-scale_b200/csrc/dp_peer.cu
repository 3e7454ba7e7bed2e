// Gradient all-reduce fused with the optimiser update over NVLink peer memory (one kernel per step and
// rank; no NCCL call, no host work).  BASELINE.json asks for the Spark driver's gradient averaging to
// be replaced by a collective over NVLink; the reference snapshot has no averaging code at all
// (SURVEY.md section 0), so the semantics are the ones pinned in SURVEY.md 8e: synchronous mean-gradient
// averaging every step, identical on every rank.
//
// Each rank's bucket (gradient sums + loss sum, double-buffered by iteration parity, plus `world`
// arrival flags) lives in memory that its peers map (the caller sets that up, e.g. with torch's
// symmetric memory).  Per step: block 0 release-stores (iteration+1) into flag[rank] of every peer;
// every block spins (bounded) on its OWN rank's flags until all peers have announced this iteration;
// then each thread loads the same four floats from all `world` buckets (L1-bypassing loads straight over
// NVLink; 298 KB per peer for the 32-256-256-1 net), adds them in rank order -- so every rank gets the
// same bits -- and applies the update.  The 8 x 298 KB a rank pulls at world = 8 cost ~4 us of NVLink
// time; what the kernel removes is the ~20 us launch + protocol latency of a small NCCL all-reduce and
// two extra kernels.  Double buffering makes one flag exchange per step enough: a rank can only
// overwrite parity p again after every peer has announced the step in between, i.e. finished reading.
#include "common.cuh"
#include "optim_math.cuh"
#include <math.h>

namespace dlas {
namespace {

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {   // never served from a stale L1 line
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

constexpr int MAXW = 8;     // one NVSwitch box

__global__ void __launch_bounds__(256)
dp_allreduce_update_kernel(const dlas_opt_cfg cfg, float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                           const unsigned long long* __restrict__ peer_bases, const int rank, const int world,
                           const size_t n, const size_t parity_stride, const size_t flags_offset, const float grad_scale,
                           const dlas_step_state* __restrict__ st, float* __restrict__ loss_ring, const long long ring_len) {
  __shared__ OptScalars o;
  __shared__ unsigned long long bases[MAXW];
  __shared__ int failed;
  const long long it = st->iteration;
  const unsigned epoch = (unsigned)(it + 1);
  if (threadIdx.x == 0) {
    double lr_t = (double)cfg.lr / (1.0 + (double)cfg.decay * (double)it);
    if (cfg.kind == DLAS_OPT_ADAM) {
      const double t = (double)it + 1.0;
      lr_t = lr_t * sqrt(1.0 - pow((double)cfg.beta2, t)) / (1.0 - pow((double)cfg.beta1, t));
    }
    o.kind = cfg.kind; o.lr_t = (float)lr_t; o.momentum = cfg.momentum; o.nesterov = cfg.nesterov;
    o.beta1 = cfg.beta1; o.beta2 = cfg.beta2; o.eps = cfg.eps;
    failed = 0;
  }
  if (threadIdx.x < world) bases[threadIdx.x] = peer_bases[threadIdx.x];
  __syncthreads();
  // announce: my bucket of this iteration is complete (the kernel that filled it precedes this one in the stream)
  if (blockIdx.x == 0 && threadIdx.x < world) {
    unsigned* f = reinterpret_cast<unsigned*>(reinterpret_cast<float*>(bases[threadIdx.x]) + flags_offset) + rank;
    __threadfence_system();
    st_release_sys(f, epoch);
  }
  // wait until every peer has announced this iteration (flags are monotonic: a peer may already be one step ahead)
  if (threadIdx.x < world) {
    const unsigned* f = reinterpret_cast<const unsigned*>(reinterpret_cast<const float*>(bases[rank]) + flags_offset) + threadIdx.x;
    bool seen = false;
    for (unsigned spin = 0; spin < (1u << 22); ++spin) {
      if ((int)(ld_acquire_sys(f) - epoch) >= 0) { seen = true; break; }
      __nanosleep(40);
    }
    if (!seen) failed = 1;
  }
  __syncthreads();
  const size_t off = (it & 1) ? parity_stride : 0;
  const size_t i0 = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i0 > n) return;                      // element n is the loss sum
  // all peer loads first (each is a ~2 us NVLink round trip: they must be in flight together), then the adds
  float4 q[MAXW];
#pragma unroll
  for (int r = 0; r < MAXW; ++r)
    q[r] = (r < world) ? ld_peer4(reinterpret_cast<const float*>(bases[r]) + off + i0) : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < MAXW; ++r) {         // rank order: bit-identical on every rank (absent ranks add +0)
    s.x += q[r].x; s.y += q[r].y; s.z += q[r].z; s.w += q[r].w;
  }
  // a peer never announced this iteration: its bucket may be stale.  Poison every sum (parameters and loss become NaN on
  // this rank) instead of applying a partial average -- the same failure the push/poll path shows (tc_reduce_kernel)
  if (failed) s.x = s.y = s.z = s.w = __int_as_float(0x7fc00000);
  const float sv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const size_t i = i0 + e;
    if (i < n) {
      float pi = p[i], mi = m[i], vi = (o.kind == 0) ? 0.f : v[i];
      opt_update(o, sv[e] * grad_scale, pi, mi, vi);
      p[i] = pi;
      m[i] = mi;
      if (o.kind != 0) v[i] = vi;
    } else if (i == n && loss_ring) {
      loss_ring[it % ring_len] = sv[e] * grad_scale;
    }
  }
}

}  // namespace
}  // namespace dlas

extern "C" int dlas_dp_allreduce_update(const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                                        const uint64_t* peer_bases_dev, int rank, int world, size_t n,
                                        size_t parity_stride, size_t flags_offset, float grad_scale,
                                        const dlas_step_state* state_dev, float* loss_ring, int64_t ring_len,
                                        void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  DLAS_REQUIRE(cfg && params && m && peer_bases_dev && state_dev, "dlas_dp_allreduce_update: null pointer");
  DLAS_REQUIRE(cfg->kind == DLAS_OPT_SGD || (cfg->kind == DLAS_OPT_ADAM && v), "dlas_dp_allreduce_update: bad optimiser / missing v");
  DLAS_REQUIRE(world >= 1 && world <= MAXW && rank >= 0 && rank < world, "dlas_dp_allreduce_update: rank %d / world %d", rank, world);
  DLAS_REQUIRE(parity_stride % 4 == 0 && parity_stride >= n + 1 && flags_offset >= 2 * parity_stride,
               "dlas_dp_allreduce_update: bucket layout (parity_stride %zu, flags_offset %zu) does not fit n = %zu", parity_stride,
               flags_offset, n);
  DLAS_REQUIRE(loss_ring == nullptr || ring_len >= 1, "dlas_dp_allreduce_update: ring_len < 1");
  const size_t n4 = (n + 1 + 3) / 4;
  dp_allreduce_update_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      *cfg, params, m, v, reinterpret_cast<const unsigned long long*>(peer_bases_dev), rank, world, n, parity_stride, flags_offset,
      grad_scale, state_dev, loss_ring, (long long)ring_len);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}
