// Shared by lstm.cu (fp32 CUDA-core kernels) and lstm_tc.cu (tcgen05 kernels): parameter layout and BPTT stash indexing of
// the LSTM regressor (Readme.md:21; BASELINE.json configs[4]).
#pragma once
#include "common.cuh"

namespace dlas {
namespace lstm {

constexpr int MAXL = 4;

struct LstmDesc {
  int d_in, H, L, T;
  int k_off[MAXL], b_off[MAXL], d_l[MAXL];
  int hw_off, hb_off, n_params;
  int fast;            // stash layout: 0 = [seq][l][t][comp][H] (the fp32 kernels'), 1 = the tensor-core kernels' own (below)
  size_t n_pad;        // sequences rounded up to 32 (fast layout)
};

inline int make_desc(int d_in, int H, int L, int T, LstmDesc& d) {
  DLAS_REQUIRE(L >= 1 && L <= MAXL, "lstm: n_layers %d not in 1..%d", L, MAXL);
  DLAS_REQUIRE(H >= 1 && H <= 128 && d_in >= 1 && d_in <= 128, "lstm: hidden %d / d_in %d not in 1..128", H, d_in);
  DLAS_REQUIRE(T >= 1, "lstm: T %d < 1", T);
  d.d_in = d_in; d.H = H; d.L = L; d.T = T;
  d.fast = 0; d.n_pad = 0;
  int o = 0, dl = d_in;
  for (int l = 0; l < L; ++l) {
    d.k_off[l] = o; o += (dl + H) * 4 * H;
    d.b_off[l] = o; o += 4 * H;
    d.d_l[l] = dl;
    dl = H;
  }
  d.hw_off = o; o += H;
  d.hb_off = o; o += 1;
  d.n_params = o;
  return DLAS_OK;
}

__host__ __device__ inline size_t pstride(int n_params) { return ((size_t)n_params + 1 + 3) & ~(size_t)3; }
// stash[seq][l][t][comp][H], comp: 0 i, 1 f, 2 g, 3 o, 4 c, 5 h (0..3 become dz after the backward pass)
__device__ __forceinline__ size_t stash_idx(const LstmDesc& d, size_t seq, int l, int t, int comp) {
  return ((((seq * d.L + l) * d.T + t) * 6) + comp) * (size_t)d.H;
}

// Stash access of the tensor-core kernels.  Their epilogue threads own one sequence (TMEM lane) and 16 consecutive units; in
// the fp32 layout a warp's 16-byte accesses would touch 32 different lines.  The fast layout (used when every stage of the
// step runs on the tensor cores) is [l][t][comp][16-unit group][sequence block of 32][4 x float4][32 sequences][4]: the
// warp's lanes then read / write 512 contiguous bytes per instruction.  `unit` is a multiple of 4 for the float4 form.
struct StashPos {     // the per-sequence part of a fast-layout address, computed once per thread and tile / chunk
  size_t inner, ugs, inner16;
};
__device__ __forceinline__ StashPos stash_pos(const LstmDesc& d, size_t seq) {
  StashPos p;
  p.inner = (((seq >> 5) * 4) * 32 + (seq & 31)) * 4;
  p.ugs = d.n_pad * 16;
  p.inner16 = (seq >> 5) * 512 + (seq & 31) * 8;
  return p;
}
// fp16 side arrays of the all-tensor-core step (fast layout only): h of every (layer, step) and dz of every (layer, step,
// gate) as the operands the next consumer needs -- the layer above's A tile, the weight-gradient GEMM -- so that nobody
// re-reads fp32 to round it again.  Slab (one (l, t[, gate])) = [16-unit group][sequence block of 32][2 x 16 B][32 sequences]
// [8 halves]: a warp's lanes touch 512 contiguous bytes per 16-byte access.  Index in halves; `unit` a multiple of 8.
__device__ __forceinline__ size_t idx16(const StashPos& p, size_t slab, int unit) {
  return (slab * 8 + (size_t)(unit >> 4)) * p.ugs + p.inner16 + (size_t)((((unit >> 3) & 1) << 8) + (unit & 7));
}
template <bool FAST>
__device__ __forceinline__ size_t sidx(const LstmDesc& d, const StashPos& p, size_t seq, int l, int t, int comp, int unit) {
  if (!FAST) return stash_idx(d, seq, l, t, comp) + unit;
  return (size_t)(((l * d.T + t) * 6 + comp) * (d.H >> 4) + (unit >> 4)) * p.ugs + p.inner + (size_t)((((unit >> 2) & 3) << 7) + (unit & 3));
}
template <bool FAST>
__device__ __forceinline__ size_t didx(const LstmDesc& d, const StashPos& p, size_t seq, int t, int unit) {
  if (!FAST) return (seq * d.T + t) * (size_t)d.H + unit;
  return (size_t)(t * (d.H >> 4) + (unit >> 4)) * p.ugs + p.inner + (size_t)((((unit >> 2) & 3) << 7) + (unit & 3));
}
__device__ __forceinline__ size_t sidx(const LstmDesc& d, size_t seq, int l, int t, int comp, int unit) {
  if (!d.fast) return stash_idx(d, seq, l, t, comp) + unit;
  const size_t slab = ((((size_t)l * d.T + t) * 6 + comp) * (d.H >> 4) + (unit >> 4)) * (d.n_pad * 16);
  return slab + (((seq >> 5) * 4 + ((unit >> 2) & 3)) * 32 + (seq & 31)) * 4 + (unit & 3);
}
// partial dx of a layer ([rank-partial][sequence][t][unit] in the fp32-compatible form): same idea
__device__ __forceinline__ size_t didx(const LstmDesc& d, size_t seq, int t, int unit) {
  if (!d.fast) return (seq * d.T + t) * (size_t)d.H + unit;
  const size_t slab = ((size_t)t * (d.H >> 4) + (unit >> 4)) * (d.n_pad * 16);
  return slab + (((seq >> 5) * 4 + ((unit >> 2) & 3)) * 32 + (seq & 31)) * 4 + (unit & 3);
}

// Layer pipeline of the tensor-core recurrences: the CTA pair of layer l publishes how far it is (one arrival per epilogue
// warp and step, release at gpu scope after the warp's stores), the pair of the layer that consumes its sequence polls the
// counter (acquire) a round before it needs the data.  Producers have the lower block indices, so a waiting CTA's producer
// has always been scheduled; the poll is bounded all the same (false = timed out, the caller raises the error word).
__device__ __forceinline__ void prog_arrive(unsigned* p) { asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory"); }
__device__ __forceinline__ bool prog_wait(const unsigned* p, unsigned target) {
  for (unsigned spin = 0; spin < (1u << 23); ++spin) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    if (v >= target) return true;
    __nanosleep(64);
  }
  return false;
}
constexpr unsigned PROG_WARPS = 32;   // epilogue warps of a pair = arrivals per step

}  // namespace lstm
}  // namespace dlas
