// tcgen05 / TMEM / mbarrier primitives (inline PTX, sm_100a) and the shared-memory operand layout
// used by the tensor-core MLP kernel.
//
// Operand tiles are fp16, stored "column-blocked with 128-byte swizzle":
//   element (row r, col c) of a tile with `rows` rows lives at byte
//     (c / 64) * rows * 128  +  (r / 8) * 1024  +  (r % 8) * 128  +  ((((c % 64) / 8) ^ (r % 8)) * 16)  +  (c % 8) * 2
// i.e. per block of 64 columns, rows are 128 bytes apart and the eight 16-byte chunks of a row are
// XOR-swizzled with the row index (the hardware SWIZZLE_128B pattern; tile bases are 1024-byte
// aligned).  ONE physical layout serves both views tcgen05.mma can take of it:
//   K-major  (K = col): 8-row groups SBO = 1024 B apart; a K=16 step is 32 B inside the 128-B row;
//                       every 4th step moves to the next 64-column block (+rows*128 B)
//   MN-major (MN = col, K = row): 64-column blocks LBO = rows*128 B apart, 8-row K groups SBO = 1024 B
//                       apart; a K=16 step is two K groups (+2048 B)
// (descriptor bit layout as in CUTLASS cute/arch/mma_sm100_desc.hpp; version = 1, layout = SWIZZLE_128B)
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace dlas {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte offset of the 16-byte chunk holding columns [c, c+8) (c % 8 == 0) of row r
__device__ __forceinline__ uint32_t tile_chunk_off(int rows, int r, int c) {
  return (uint32_t)((c >> 6) * rows * 128 + (r >> 3) * 1024 + (r & 7) * 128 + ((((c & 63) >> 3) ^ (r & 7)) << 4));
}

// ---- descriptors ------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;   // SWIZZLE_128B
  return d;
}
// K-major view starting at K offset k0 (multiple of 16) of a tile with `rows` rows, first row row0 (multiple of 8)
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int rows, int row0, int k0) {
  uint32_t a = base + (uint32_t)((k0 >> 6) * rows * 128 + (row0 >> 3) * 1024 + (k0 & 63) * 2);
  return make_desc(a, 16, 1024);
}
// MN-major view: MN starts at column mn0 (multiple of 64), K starts at row k0 (multiple of 8)
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int rows, int mn0, int k0) {
  uint32_t a = base + (uint32_t)((mn0 >> 6) * rows * 128 + (k0 >> 3) * 1024);
  return make_desc(a, (uint32_t)rows * 128, 1024);
}
// Descriptor offsets (added to a base descriptor; the address field counts 16-byte units and never
// carries out of its 14 bits for addresses inside one CTA's shared memory):
//   K step i (16 elements) of a K-major view of a tile with `rows` rows
__host__ __device__ constexpr uint64_t koff_kmajor(int rows, int i) { return (uint64_t)(((i >> 2) * rows * 128 + (i & 3) * 32) >> 4); }
//   K step i (16 rows) of an MN-major view
__host__ __device__ constexpr uint64_t koff_mnmajor(int i) { return (uint64_t)((i * 2048) >> 4); }
//   MN offset mn0 (multiple of 64 columns) of an MN-major view; row offset row0 (multiple of 8) of a K-major view
__host__ __device__ constexpr uint64_t mnoff(int rows, int mn0) { return (uint64_t)(((mn0 >> 6) * rows * 128) >> 4); }
__host__ __device__ constexpr uint64_t rowoff(int row0) { return (uint64_t)(((row0 >> 3) * 1024) >> 4); }
// kind::f16, A and B fp16, D fp32
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn, bool b_mn) {
  return (1u << 4) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// ---- tcgen05 ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy smem writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// all MMAs issued so far by this thread arrive on the mbarrier when they complete
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 32 columns of 32-bit: thread i of the warp gets lane (base lane + i), columns col..col+31
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// one lane of the (converged) warp
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

// ---- mbarrier -----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Bounded wait: returns false if the phase did not complete within ~2^26 polls (a hang would otherwise
// take the GPU box down); the kernel then raises its error flag and bails out.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t a = smem_u32(bar);
  for (uint32_t it = 0; it < (1u << 26); ++it) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}

// ---- explicit shared-space accesses (32-bit shared addresses; keeps LDS/STS instead of generic LD/ST) ----
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds128f(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v)); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// ---- cluster / DSMEM ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// execution barrier only (no release of this thread's global stores: the kernel boundary publishes those)
__device__ __forceinline__ void cluster_sync_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// asynchronous 4-byte store into a peer CTA's shared memory; completes 4 transaction bytes on the peer's mbarrier
__device__ __forceinline__ void st_async_f32(uint32_t caddr, float v, uint32_t cbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(caddr), "r"(__float_as_uint(v)), "r"(cbar)
               : "memory");
}
__device__ __forceinline__ void st_cluster_f32(uint32_t caddr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(caddr), "f"(v) : "memory");
}

// bulk copy global -> this CTA's shared memory (TMA engine, no tensor map: UBLKCP); completes `bytes` transaction bytes on
// the mbarrier.  Addresses and size are multiples of 16.
__device__ __forceinline__ void bulk_g2s(uint32_t sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sdst), "l"(gsrc),
               "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- cross-CTA hand-over of operand tiles (LSTM recurrence: each CTA of a pair writes its half of h_t into the peer's
// A tile, which the peer's tcgen05.mma then reads) -------------------------------------------------------------------
// 16-byte store into a peer CTA's shared memory (address from map_to_cta)
__device__ __forceinline__ void st_cluster_v4(uint32_t caddr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(caddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// 16 bytes into a peer CTA's shared memory as an asynchronous store that completes 16 transaction bytes on the PEER's
// mbarrier when the data has landed: no release fence on the sender's side (a releasing arrive would first wait for every
// global store the thread still has in flight)
__device__ __forceinline__ void st_async_v4(uint32_t caddr, uint4 v, uint32_t cbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(caddr), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w), "r"(cbar)
               : "memory");
}
// arrive on a peer's mbarrier without ordering this thread's earlier writes (a buffer has been READ and may be overwritten;
// the reads' results have been consumed by earlier instructions of the arriving warp)
__device__ __forceinline__ void mbar_arrive_remote_relaxed(uint32_t cbar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cbar) : "memory");
}
// generic-proxy writes (own or peer shared memory) -> visible to async-proxy readers (tcgen05.mma operand fetch)
__device__ __forceinline__ void fence_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
// arrive on a mbarrier of a peer CTA (address from map_to_cta), releasing this thread's earlier writes at cluster scope
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cbar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cbar) : "memory");
}
// bounded wait with cluster-scope acquire: pairs with mbar_arrive_remote
__device__ __forceinline__ bool mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t a = smem_u32(bar);
  for (uint32_t it = 0; it < (1u << 26); ++it) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}

// bring the 128-byte line at p into L2 (no register, no wait): the recurrences touch each stash line once, a round ahead
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// bulk copy this CTA's shared memory -> global (tile images); completion through bulk async-groups of the issuing thread
__device__ __forceinline__ void bulk_s2g(void* gdst, uint32_t ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }   // all but the newest group
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }   // sources read
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }             // writes done

// packed fp16 helpers for the epilogues (one instruction per two elements)
__device__ __forceinline__ uint32_t relu_h2(uint32_t a) {
  uint32_t r;
  asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(0u));
  return r;
}
// 1.0 where a > 0 else 0.0, per half
__device__ __forceinline__ uint32_t gt0_h2(uint32_t a) {
  uint32_t r;
  asm("set.gt.f16x2.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(0u));
  return r;
}
__device__ __forceinline__ uint32_t mul_h2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

}  // namespace tc
}  // namespace dlas
