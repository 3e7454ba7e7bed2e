// Tensor-core training kernel for the 32-256-256-1 static ANN (DLAS_MODE_TC): forward, loss and
// backward of cnn.py:127-128's batch body on tcgen05.mma with fp32 accumulators in TMEM.
//
// Operands are fp16 (11-bit significand, the same as TF32; fp32 accumulate), because fp32/TF32
// operands do not fit on chip: W2 alone is 256 KB of the 227 KB of shared memory, and dW2's fp32
// accumulator (256x256) is all 512 TMEM columns of one SM.  The network is therefore split over a
// CTA PAIR (cluster of 2) along layer 2's OUTPUT units; CTA `rank` owns out-columns
// [rank*128, rank*128+128):
//     Z1 = X W1, H1 = relu(Z1+b1)                 both CTAs (K = 32, cheap, redundant)
//     Z2h = H1 W2[:, half]                        N = 128
//     p = relu(Z2h+b2h) . w3h                     per-row partial dot; the ONLY cross-CTA exchange
//                                                 (128 floats through distributed shared memory)
//     dZ2h = dy w3h relu'(Z2h)
//     dW2[:, half] += H1^T dZ2h, db2h += dZ2h^T 1 256 + 16 TMEM columns, accumulated over all tiles
//     dH1p = dZ2h W2[:, half]^T                   PARTIAL dH1; dZ1p = relu'(H1) o dH1p is linear in
//     dW1p^T += dZ1p^T [X | 1]                    the partials, so each CTA just adds its share to
//                                                 dW1 / db1 (column 32 of [X | 1] gives db1)
// MMA floor per 128-row tile per CTA: 3584 cycles (256 + 1024 + 1024 + 1024 + 256), 93 % of the
// 6680 SM-cycles the unsplit network needs per tile.
//
// Warp roles (384 threads): warps 0-7 epilogue (warp w: TMEM lane quarter w%4 = rows, column half w/4 of every
// 128-column chunk, 64 columns per thread), warp 8 = MMA issuer, warps 9-11 idle: they complete the issuer's
// warpgroup so that setmaxnreg can move its registers to the epilogue warps (104 / 200 per thread), which keep Z2h,
// the dw3 accumulators and a prefetched block of constants in registers.  Z1 and dH1p are produced as two N=128
// halves through ONE 128-column TMEM work buffer (mbarrier waits cost ~140 cycles each whatever their state --
// tools/mbar_lat.cu -- so fewer, larger hand-overs win over double-buffered N=64 chunks); the issuer interleaves
//   Z1 half 1 | Z2h (first K half as soon as H1 columns 0..127 are stored) | dH1p half 0 | dW2^T half 0 | dH1p half 1 |
//   dW2^T half 1 | dW1 half 0 | next tile's Z1 half 0 | dW1 half 1      (db2 = column sums of dZ2: two sibling warps of the issuer)
// so that the epilogue of one piece runs under the MMAs of the next.  dW2^T / db2 / dW1^T accumulate in their own TMEM
// columns for the whole launch (dW2 transposed: lane = layer-2 unit, so the drain stores 128 contiguous bytes per
// instruction).  X tiles are double-buffered; their loads are issued where the epilogue warps are about to wait.
// mbarriers: B_DONE (tcgen05.commit -> epilogue), B_FREE (work buffer read -> issuer), B_X*, B_H1A/B_H1, B_DZ2,
// B_DW2A/B (H1 half may be overwritten), B_DZ1A/B_DZ1, B_DW1, B_XC* (pair exchange).
// The step is finished by tc_reduce_kernel (below): per-CTA partials -> sums [-> NVLink push/poll exchange] -> update.
//
// Shared memory (fp16 tiles in the column-blocked SWIZZLE_128B layout of tc_common.cuh):
//   sW1 [48][256] 24 KB ([W1; b1; 0]) | sW2 [256][128] 64 KB | sX[2] [128][64] 2x16 KB ([x | 1 | 0]) |
//   sH1 [128][256] 64 KB (dZ1 overwrites it in place) | sDZ2 [128][128] 32 KB | ones [16][64] 2 KB | fp32 vectors
// TMEM columns: [0,128) work | [128,384) dW2^T half | [384,432) [448,496) dW1^T halves
#include "common.cuh"
#include "optim_math.cuh"
#include "tc_common.cuh"
#include <math.h>
#include <stdlib.h>

namespace dlas {
size_t simt_workspace_bytes(const int* dims, int n_layers);

namespace {
using namespace tc;

constexpr int D0 = 32, D1 = 256, D2 = 256;
constexpr int TILE = 128;
constexpr int CSPLIT = 2;             // epilogue warps per TMEM lane quarter (4 measured no faster: the epilogues are latency- and smem-bound): each takes 64/CSPLIT columns of a chunk
constexpr int CW = 128 / CSPLIT;      // work-buffer columns per epilogue thread
constexpr int N_EPI = 128 * CSPLIT;   // epilogue threads (warps 0 .. 4*CSPLIT-1)
constexpr int ISSUER_WARP = N_EPI / 32;
constexpr int NT = N_EPI + 128;       // + the issuer warpgroup (warp ISSUER_WARP issues; its three siblings only give their registers away)
constexpr int EPI_REGS = 200, ISSUER_REGS = 104;  // setmaxnreg: 2 x 128 x 200 + 128 x 104 = 64,512 <= 65,536
constexpr int W1_OFF = 0, B1_OFF = W1_OFF + D0 * D1, W2_OFF = B1_OFF + D1, B2_OFF = W2_OFF + D1 * D2,
              W3_OFF = B2_OFF + D2, B3_OFF = W3_OFF + D2, N_PARAMS = B3_OFF + 1;
constexpr size_t PSTRIDE = ((size_t)N_PARAMS + 1 + 3) & ~(size_t)3;

constexpr int K1 = 48;                // layer-1 contraction: 32 features + the ones column (bias row of W1) + 15 zero rows
constexpr uint32_t S_W1 = 0, S_W2 = S_W1 + K1 * 4 * 128, S_X = S_W2 + 65536, S_H1 = S_X + 2 * 16384, S_DZ2 = S_H1 + 65536,
                   S_ONES = S_DZ2 + 32768, S_F32 = S_ONES + 2048;
// fp32 area (float index)
constexpr int F_B2 = 0, F_W3 = 128, F_Y = 256 /*[2][128]*/, F_PART = 512 /*[2][128]*/, F_RP = 768 /*[CSPLIT][128]*/,
              F_RED = F_RP + CSPLIT * 128 /*[4][128]*/, F_END = F_RED + 512;
constexpr uint32_t S_BAR = S_F32 + F_END * 4;
enum { B_DONE = 0, B_FREE, B_X0, B_X1, B_H1A, B_H1, B_DZ2, B_DW2A, B_DW2B, B_DZ1A, B_DZ1, B_DW1, B_ALL, B_XC0, B_XC1, B_WIMG, B_H1L, B_H1M, B_DW1A, B_DB2, N_BARS };
// fp16 image of the weights in the kernel's own tile layout, kept in the workspace by the finishing kernel of a step loop
// (tc_reduce_kernel writes every updated parameter into it), so that the next step's set-up is three bulk copies instead of
// a 160 KB fp32 read + convert per CTA:  [sW1 tile 24 KB: W1 rows, b1 row, zero rows][sW2 tile of rank 0 64 KB][rank 1 64 KB]
constexpr uint32_t IMG_W1_BYTES = K1 * 4 * 128, IMG_W2_BYTES = 65536, IMG_BYTES = IMG_W1_BYTES + 2 * IMG_W2_BYTES;
enum { K_TRAIN = 0, K_EVAL = 1, K_PREDICT = 2, K_TRAIN_EXT = 3, K_EVAL_EXT = 4, K_PREDICT_EXT = 5 };
// K_TRAIN_EXT: layers 2 and 3 of the wide-input network d0-256-256-1 (mlp_tc_l1.cu holds layer 1): H1 arrives as a 64 KB
// tile image per 128 rows (one bulk copy instead of Z1 / E1) and each CTA's partial dZ1 tile leaves as an image (one bulk
// copy instead of the dW1 MMAs); `params` points (d0 - 32) * 256 floats into the real vector, so the W2.. offsets hold.
constexpr uint32_t IMG_TILE_BYTES = 65536;
struct ExtIo {
  const uint8_t* h1_img;   // [tile][64 KB], written by l1_fwd_tc_kernel
  uint8_t* dz1_img;        // [2 CTA ranks][img_stride bytes], read by l1_dw_tc_kernel
  size_t img_stride;
};
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8;
constexpr uint32_t SMEM_BYTES = S_TMEM + 16;

constexpr uint32_t TM_W0 = 0, TM_DW2 = 128, TM_DW1 = 384, TM_COLS = 512;

// v[0..W) per lane -> lane l (< W) returns the sum over the warp's 32 lanes of v[l]  (W = 32 or 16)
template <int W>
__device__ __forceinline__ float warp_colsum(float (&v)[W], int lane) {
#pragma unroll
  for (int s = W / 2; s >= 1; s >>= 1) {
    const bool hi = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      float send = hi ? v[i] : v[i + s];
      float keep = hi ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  float r = v[0];
  if (W == 16) r += __shfl_xor_sync(0xffffffffu, r, 16);   // lanes l and l+16 hold the two halves of column l % 16
  return r;
}

template <int W> __device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&v)[W]);
template <> __device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, uint32_t (&v)[32]) { tmem_ld32(taddr, v); }
template <> __device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, uint32_t (&v)[16]) { tmem_ld16(taddr, v); }

// every lane has done its own fences; one arrival for the warp
__device__ __forceinline__ void warp_arrive(int lane, uint64_t* bar) {
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}

__device__ __forceinline__ uint4 pack8(float4 a, float4 b) {
  return make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(b.x, b.y), pack_half2(b.z, b.w));
}

// SM clock as the kernel itself sees it (dlas_prof_sm_clock): CTA 0 adds its clock64 and globaltimer deltas of every
// launch; the ratio of the sums is the average SM frequency while this kernel runs.  For the benchmark's clocks line at
// N > 1, where NVML queries during the timed region were measured to stall the ranks' steps.
__device__ unsigned long long g_clk_sum[2] = {0ull, 0ull};
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Debug trace (library built with `build.py --trace`, run with DLAS_TC_TRACE=<file>): SM-clock stamps of CTA 0,
// epilogue warp 0 (slots 0..), the issuer (slots 512..) and the last epilogue warp (slots 1024..)
#ifdef DLAS_TC_TRACE
#define TRACE_E(k) do { if (trace && blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == N_EPI / 32 - 1) && te < 500) trace[(warp ? 1024 : 0) + te++] = ((long long)(k) << 48) | (clock64() & 0xFFFFFFFFFFFFll); } while (0)
#define TRACE_M(k) do { if (trace && blockIdx.x == 0 && lane == 0 && tmm < 500) trace[512 + tmm++] = ((long long)(k) << 48) | (clock64() & 0xFFFFFFFFFFFFll); } while (0)
#else
#define TRACE_E(k) do { } while (0)
#define TRACE_M(k) do { } while (0)
#endif

// KIND: K_TRAIN forward + loss + backward (partials out); K_EVAL forward + loss (loss partial out: validation pass,
// cnn.py:121); K_PREDICT forward only (yhat out: model.predict / test_on_batch, cnn.py:132).  The forward-only kinds run
// Z1 | E1 | Z2h | partial dot + exchange per tile and leave the accumulator columns unused.
template <int KIND>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1)
mlp_tc_kernel(const int loss_kind, const float* __restrict__ params, const float* __restrict__ X_base,
              const float* __restrict__ y_base, const size_t n_rows, float* __restrict__ partials,
              unsigned* __restrict__ err_flag, const unsigned call_id, long long* __restrict__ trace,
              const long long* __restrict__ cursor, const int rows_ready, const size_t ldx,
              const uint8_t* __restrict__ w_image, float* __restrict__ yhat_out, const ExtIo ext) {
  constexpr bool EXT = (KIND >= K_TRAIN_EXT);           // layer 1 lives in mlp_tc_l1.cu (H1 tile images in)
  constexpr int BK = EXT ? KIND - K_TRAIN_EXT : KIND;   // what the kernel computes: K_TRAIN / K_EVAL / K_PREDICT
  constexpr bool TRAIN = (BK == K_TRAIN);
  // ldx: floats between consecutive rows of X (32 for a dense matrix; smaller for lag windows laid over a raw series)
  // device-resident batch cursor (replayable steps): this launch works on rows [cursor*n_rows, (cursor+1)*n_rows)
  extern __shared__ __align__(1024) uint8_t smem[];
  float* f32 = reinterpret_cast<float*>(smem + S_F32);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sW1 = sbase + S_W1, sW2 = sbase + S_W2, sX = sbase + S_X, sH1 = sbase + S_H1, sDZ2 = sbase + S_DZ2,
                 sONES = sbase + S_ONES, sF = sbase + S_F32;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  bool ok = (sbase & 1023u) == 0;   // the swizzled tile layout needs a 1024-byte aligned base
  [[maybe_unused]] int te = 0, tmm = 0;   // trace slots (DLAS_TC_TRACE builds)
  TRACE_E(1);
  const bool clk_owner = (blockIdx.x == 0 && tid == 0);
  const long long clk_c0 = clk_owner ? clock64() : 0;
  const unsigned long long clk_g0 = clk_owner ? globaltimer_ns() : 0ull;

  // ---- one-time set-up: barriers, TMEM, weights -> fp16 tiles -------------------------------------------
  if (tid == 0) {
    mbar_init(&bars[B_DONE], 1);
    // one arrival per epilogue WARP (lane 0, after __syncwarp): 256 per-thread arrivals on one word would serialise
    mbar_init(&bars[B_FREE], N_EPI / 32);
    mbar_init(&bars[B_X0], N_EPI / 32); mbar_init(&bars[B_X1], N_EPI / 32);
    mbar_init(&bars[B_H1A], N_EPI / 32); mbar_init(&bars[B_H1], N_EPI / 32); mbar_init(&bars[B_DZ2], N_EPI / 32); mbar_init(&bars[B_DZ1A], N_EPI / 32);
    mbar_init(&bars[B_DZ1], N_EPI / 32);
    mbar_init(&bars[B_DB2], 2);                                // the two db2 warps have read sDZ2
    mbar_init(&bars[B_ALL], 1); mbar_init(&bars[B_DW1], 1); mbar_init(&bars[B_DW1A], 1); mbar_init(&bars[B_DW2A], 1); mbar_init(&bars[B_DW2B], 1);
    mbar_init(&bars[B_XC0], 1); mbar_init(&bars[B_XC1], 1);   // exchange of the partial output dots: 1 arrival + 512 tx bytes
    mbar_init(&bars[B_WIMG], 1);                               // weight image: 1 arrival + 88 KB of bulk-copy bytes
    mbar_init(&bars[B_H1L], 1); mbar_init(&bars[B_H1M], 1);   // layer 1 outside: H1 tile image, columns 0..127 / 128..255: 1 arrival + 32 KB each
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  // X tile loader (epilogue threads): thread -> (row tid / XT, XF features), XT threads per row
  constexpr int XT = N_EPI / TILE, XF = D0 / XT;      // 2 threads per row, 16 features each; a warp covers 32 rows of one
  const int xr = tid % TILE, xh = tid / TILE;         // half, so its 16-byte stores into the swizzled tile do not conflict
  float4 xv[XF / 4];
  float yv = 0.f;
  auto x_issue = [&](size_t tile, const float* __restrict__ X, const float* __restrict__ y) {
    const size_t row0 = tile * TILE;
    const bool in = row0 + xr < n_rows;
    const float* src = X + (row0 + xr) * ldx + xh * XF;
#pragma unroll
    for (int i = 0; i < XF / 4; ++i) xv[i] = (in && !EXT) ? __ldcs(reinterpret_cast<const float4*>(src + i * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (BK != K_PREDICT && tid < TILE) yv = (row0 + tid < n_rows) ? __ldcs(y + row0 + tid) : 0.f;
  };
  // rows_ready: the host states that X / y were NOT written by the kernel that precedes this one in the stream (the
  // library's own step loops: the predecessor is the previous step's finishing kernel).  Only then, and with a host-side
  // batch offset (no device cursor), the first tile's loads are issued before the dependency wait and complete under the
  // weight set-up.  Single calls (dlas_mlp_grad, the first step of a loop) load after the wait: their predecessor may be
  // the producer of the rows and may have released its dependents early.
  const int my_tiles = (n_tiles > (size_t)pair) ? (int)((n_tiles - pair + n_pairs - 1) / n_pairs) : 0;
  const bool early_x = (cursor == nullptr) && rows_ready != 0;
  if (early_x && tid < N_EPI && my_tiles > 0) x_issue(pair, X_base, y_base);
  // Programmatic dependent launch: everything above overlaps the tail of the previous kernel in the stream
  // (the optimiser update that produced `params`); from here on its writes are visible.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const size_t batch0 = cursor ? (size_t)__ldcg(cursor) * n_rows : 0;   // the cursor is written by the previous step's kernels
  // forward-only with layer 1 outside: a timed-out wait in l1_fwd_tc_kernel (same call id) poisons the predictions too
  if (EXT && BK == K_PREDICT && err_flag != nullptr && __ldcg(err_flag) == call_id) ok = false;
  const float* __restrict__ X = X_base + batch0 * ldx;
  const float* __restrict__ y = (BK == K_PREDICT) ? nullptr : y_base + batch0;
  if (w_image != nullptr) {
    // the finishing kernel of the previous step left the updated weights as fp16 tiles: fetch them with the copy engine
    if (tid == 0) {
      mbar_arrive_expect_tx(&bars[B_WIMG], (EXT ? 0u : IMG_W1_BYTES) + IMG_W2_BYTES);
      if (!EXT) bulk_g2s(sW1, w_image, IMG_W1_BYTES, &bars[B_WIMG]);
#pragma unroll
      for (int c = 0; c < 4; ++c)
        bulk_g2s(sW2 + c * 16384, w_image + IMG_W1_BYTES + rank * IMG_W2_BYTES + c * 16384, 16384, &bars[B_WIMG]);
    }
  } else {
    // sW1 = [W1 ; b1 ; 0]: rows 0..31 weights, row 32 the bias (it multiplies the ones column of [x | 1 | 0]),
    // rows 33..47 zero -> Z1 comes out of the MMA with the bias already added
    for (int idx = tid; idx < (EXT ? 0 : K1 * (D1 / 8)); idx += NT) {
      const int r = idx / (D1 / 8), c = (idx % (D1 / 8)) * 8;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (r <= D0) {
        const float* src = (r < D0) ? params + W1_OFF + r * D1 + c : params + B1_OFF + c;
        a = __ldg(reinterpret_cast<const float4*>(src));
        b = __ldg(reinterpret_cast<const float4*>(src + 4));
      }
      sts128(sW1 + tile_chunk_off(K1, r, c), pack8(a, b));
    }
    const float* W2 = params + W2_OFF + rank * 128;           // [256][256], my 128 out-columns
    for (int i0 = 0; i0 < D1 * 16; i0 += NT * 4) {             // 4 chunks (8 loads) in flight per thread
      float4 a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * NT + tid;
        if (idx < D1 * 16) {
          const int r = idx >> 4, c = (idx & 15) * 8;
          a[u] = __ldg(reinterpret_cast<const float4*>(W2 + (size_t)r * D2 + c));
          b[u] = __ldg(reinterpret_cast<const float4*>(W2 + (size_t)r * D2 + c + 4));
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * NT + tid;
        if (idx < D1 * 16) sts128(sW2 + tile_chunk_off(D1, idx >> 4, (idx & 15) * 8), pack8(a[u], b[u]));
      }
    }
  }
  {
    if (tid < 128) {
      f32[F_B2 + tid] = __ldg(params + B2_OFF + rank * 128 + tid);
      f32[F_W3 + tid] = __ldg(params + W3_OFF + rank * 128 + tid);
      // constant part of [x | 1 | 0] in both X buffers: column 32 = 1, columns 33..63 = 0
      for (int j = 0; j < 2; ++j) {
        sts128(sX + j * 16384 + tile_chunk_off(TILE, tid, 32), make_uint4(0x00003C00u, 0u, 0u, 0u));
        for (int c = 40; c < 64; c += 8) sts128(sX + j * 16384 + tile_chunk_off(TILE, tid, c), make_uint4(0u, 0u, 0u, 0u));
      }
      sts128(sONES + tid * 16, make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u));  // 2 KB of fp16 1.0
    }
  }
  const float b3 = __ldg(params + B3_OFF);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  cluster_sync_all();  // the peer CTA is running: its shared memory may be written from here on
  const uint32_t tm = *tmem_slot;
  TRACE_E(2);

  // dependents (the reduction kernel) may be scheduled from now on; they block in griddepcontrol.wait until this grid is done
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp >= ISSUER_WARP) {
    // the issuer warpgroup hands most of its registers to the two epilogue warpgroups
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(ISSUER_REGS));
  }
  if (warp > ISSUER_WARP) {
    // Siblings of the issuer warp.  Two of them compute db2h = dZ2h^T 1 on the CUDA cores: thread j sums columns 2j, 2j+1 of
    // every dZ2 tile over its 128 rows (one conflict-free 4-byte read per row).  As eight N = 16 MMAs the same sums cost the
    // tensor pipe ~410 cycles per tile in the half of the tile where it is the bottleneck.
    if (TRAIN && warp <= ISSUER_WARP + 2) {
      const int j = (warp - ISSUER_WARP - 1) * 32 + lane;     // 0..63
      float s0 = 0.f, s1 = 0.f;
      for (int it = 0; it < my_tiles; ++it) {
        ok = mbar_wait(&bars[B_DZ2], it & 1) && ok;
#pragma unroll 8
        for (int r = 0; r < TILE; ++r) {
          const uint32_t w = lds32(sDZ2 + tile_chunk_off(TILE, r, (2 * j) & ~7) + (uint32_t)((2 * j) & 7) * 2);
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
          s0 += f.x; s1 += f.y;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[B_DB2]);
      }
      if (my_tiles > 0) {
        float* my = partials + (size_t)blockIdx.x * PSTRIDE;
        my[B2_OFF + rank * 128 + 2 * j] = s0;
        my[B2_OFF + rank * 128 + 2 * j + 1] = s1;
      }
    }
  } else if (warp == ISSUER_WARP) {
    // =========================== MMA issuer ===========================
    // The whole warp runs the control flow (so descriptor arithmetic stays warp-uniform); one elected
    // lane issues.  Descriptors are a constant base plus a compile-time offset per K step.
    constexpr uint32_t ID_KM_MN_128 = make_idesc(128, 128, false, true);
    constexpr uint32_t ID_KM_KM_128 = make_idesc(128, 128, false, false);
    constexpr uint32_t ID_MN_MN_128 = make_idesc(128, 128, true, true);
    constexpr uint32_t ID_MN_MN_48 = make_idesc(128, 48, true, true);
    const bool leader = elect_one();
    // the second X buffer is 16 KB further: +1024 in the descriptor's (address >> 4) field
    uint64_t dX_km0 = desc_kmajor(sX, TILE, 0, 0), dX_mn0 = desc_mnmajor(sX, TILE, 0, 0);
    uint64_t dW1_mn = desc_mnmajor(sW1, K1, 0, 0);
    uint64_t dW2_mn = desc_mnmajor(sW2, D1, 0, 0), dW2_km = desc_kmajor(sW2, D1, 0, 0);
    uint64_t dH1_km = desc_kmajor(sH1, TILE, 0, 0), dH1_mn = desc_mnmajor(sH1, TILE, 0, 0);
    uint64_t dDZ2_km = desc_kmajor(sDZ2, TILE, 0, 0), dDZ2_mn = desc_mnmajor(sDZ2, TILE, 0, 0);
    uint64_t dONES = desc_kmajor(sONES, 16, 0, 0);
    // Re-derive "base + compile-time offset" inside every tile: hoisted out of the loop, the ~100 descriptor sums
    // do not fit the issuer's reduced register budget and come back as local-memory loads in front of every MMA.
    auto pin_descs = [&]() {
      asm volatile("" : "+l"(dX_km0), "+l"(dX_mn0), "+l"(dW1_mn), "+l"(dW2_mn), "+l"(dW2_km));
      asm volatile("" : "+l"(dH1_km), "+l"(dH1_mn), "+l"(dDZ2_km), "+l"(dDZ2_mn), "+l"(dONES));
    };
    uint32_t n = 0;  // running index of work-buffer uses (five per tile): use n completes phase n of B_DONE / B_FREE
    auto wait_free = [&]() {   // the epilogue has read use n-1 out of the work buffer
      ok = mbar_wait(&bars[B_FREE], (n & 1) ^ 1) && ok;
      fence_after();
    };
    // Z1 + b1 = [X | 1 | 0] [W1 ; b1 ; 0], K = 48, as two N=128 halves.  Half 0 of tile t+1 is issued near the END of
    // tile t, so that it is already in TMEM when the epilogue warps come back for it.
    auto issue_z1 = [&](int xb, int h) {
      wait_free();
      if (leader) {
#pragma unroll
        for (int i = 0; i < K1 / 16; ++i)
          mma_f16(tm + TM_W0, dX_km0 + (uint64_t)(xb * 1024) + koff_kmajor(TILE, i), dW1_mn + mnoff(K1, 128) * h + koff_mnmajor(i),
                  ID_KM_MN_128, i > 0);
        mma_commit(&bars[B_DONE]);
      }
      __syncwarp();
      ++n;
    };
    if (w_image != nullptr) ok = mbar_wait(&bars[B_WIMG], 0) && ok;   // bulk copies of the weight tiles have landed
    if (EXT) {
      if (my_tiles > 0 && leader) {   // H1 of the first tile: the image l1_fwd_tc_kernel left, as two bulk copies (K halves of Z2h)
        const uint8_t* src = ext.h1_img + (size_t)pair * IMG_TILE_BYTES;
        mbar_arrive_expect_tx(&bars[B_H1L], IMG_TILE_BYTES / 2);
        bulk_g2s(sH1, src, IMG_TILE_BYTES / 2, &bars[B_H1L]);
        mbar_arrive_expect_tx(&bars[B_H1M], IMG_TILE_BYTES / 2);
        bulk_g2s(sH1 + IMG_TILE_BYTES / 2, src + IMG_TILE_BYTES / 2, IMG_TILE_BYTES / 2, &bars[B_H1M]);
      }
      __syncwarp();
    } else if (my_tiles > 0) {
      ok = mbar_wait(&bars[B_X0], 0) && ok;
      issue_z1(0, 0);
    }
    for (int it = 0; it < my_tiles; ++it) {
      const int xb = it & 1;
      const bool first = (it == 0);
      pin_descs();
      TRACE_M(10);
      if (!EXT) issue_z1(xb, 1);
      // Z2h = H1 W2h: one N=128 chain, K = 256; its first half only needs H1 columns 0..127, which are in place
      // while the epilogue is still converting columns 128..255
      TRACE_M(12);
      wait_free();
      ok = mbar_wait(&bars[EXT ? B_H1L : B_H1A], it & 1) && ok;
      fence_after();
      if (leader) {
#pragma unroll
        for (int i = 0; i < D1 / 32; ++i)
          mma_f16(tm + TM_W0, dH1_km + koff_kmajor(TILE, i), dW2_mn + koff_mnmajor(i), ID_KM_MN_128, i > 0);
      }
      __syncwarp();
      ok = mbar_wait(&bars[EXT ? B_H1M : B_H1], it & 1) && ok;
      TRACE_M(13);
      fence_after();
      if (leader) {
#pragma unroll
        for (int i = D1 / 32; i < D1 / 16; ++i)
          mma_f16(tm + TM_W0, dH1_km + koff_kmajor(TILE, i), dW2_mn + koff_mnmajor(i), ID_KM_MN_128, 1);
        mma_commit(&bars[B_DONE]);
      }
      __syncwarp();
      ++n;
      TRACE_M(14);
      if (!TRAIN && EXT) {   // forward only, layer 1 outside: the next H1 image may land once Z2h has read this one
        if (it + 1 < my_tiles) {
          ok = mbar_wait(&bars[B_DONE], (n - 1) & 1) && ok;
          if (leader) {
            const uint8_t* src = ext.h1_img + ((size_t)pair + (size_t)(it + 1) * n_pairs) * IMG_TILE_BYTES;
            mbar_arrive_expect_tx(&bars[B_H1L], IMG_TILE_BYTES / 2);
            bulk_g2s(sH1, src, IMG_TILE_BYTES / 2, &bars[B_H1L]);
            mbar_arrive_expect_tx(&bars[B_H1M], IMG_TILE_BYTES / 2);
            bulk_g2s(sH1 + IMG_TILE_BYTES / 2, src + IMG_TILE_BYTES / 2, IMG_TILE_BYTES / 2, &bars[B_H1M]);
          }
          __syncwarp();
        }
        continue;
      }
      if (!TRAIN) {   // forward only: the next tile's first Z1 half follows at once (work buffer free after the partial-dot pass)
        if (it + 1 < my_tiles) {
          ok = mbar_wait(&bars[B_X0 + (xb ^ 1)], ((it + 1) >> 1) & 1) && ok;
          issue_z1(xb ^ 1, 0);
        }
        continue;
      }
      ok = mbar_wait(&bars[B_DZ2], it & 1) && ok;
      TRACE_M(16);
      fence_after();
      // Order: dH1p half 0 | dW2 rows 0..127 | dH1p half 1 | dW2 rows 128..255 | dW1 rows 0..127 | next tile's
      // Z1 half 0 | dW1 rows 128..255.  The epilogue overwrites H1 with dZ1 in place, half by half: half h (H1 columns
      // 128h..128h+127) may be written once the dW2 half that reads those columns as its M rows has completed
      // (B_DW2A / B_DW2B), so the epilogue fetches dH1p half h and the ReLU mask while that dW2 half still runs.
      auto issue_dw2 = [&](int mh) {
        if (leader) {
#pragma unroll
          for (int i = 0; i < TILE / 16; ++i)
            // dW2h^T[c][k] (lane = layer-2 unit c, column = H1 unit k): the drain then stores 128 contiguous bytes per instruction
            mma_f16(tm + TM_DW2 + mh * 128, dDZ2_mn + koff_mnmajor(i), dH1_mn + mnoff(TILE, mh * 128) + koff_mnmajor(i), ID_MN_MN_128,
                    !(first && i == 0));
        }
        __syncwarp();
      };
      auto issue_dh1 = [&](int h) {   // dH1p half h = dZ2h W2h^T[:, 128h..128h+127]
        wait_free();
        TRACE_M(61);
        if (leader) {
#pragma unroll
          for (int i = 0; i < 128 / 16; ++i)
            mma_f16(tm + TM_W0, dDZ2_km + koff_kmajor(TILE, i), dW2_km + rowoff(h * 128) + koff_kmajor(D1, i), ID_KM_KM_128, i > 0);
          mma_commit(&bars[B_DONE]);
        }
        __syncwarp();
        ++n;
      };
      auto issue_dw1 = [&](int mh) {  // dW1p^T rows mh*128.. += dZ1p^T [X | 1]
        if (leader) {
#pragma unroll
          for (int i = 0; i < TILE / 16; ++i)
            mma_f16(tm + TM_DW1 + mh * 64, dH1_mn + mnoff(TILE, mh * 128) + koff_mnmajor(i), dX_mn0 + (uint64_t)(xb * 1024) + koff_mnmajor(i),
                    ID_MN_MN_48, !(first && i == 0));
        }
        __syncwarp();
      };
      issue_dh1(0);
      TRACE_M(60);
      issue_dw2(0);
      if (leader) mma_commit(&bars[B_DW2A]);   // H1 columns 0..127 may be overwritten once this arrives
      __syncwarp();
      issue_dh1(1);
      issue_dw2(1);
      if (leader) mma_commit(&bars[B_DW2B]);   // ... and columns 128..255 after this one
      __syncwarp();
      TRACE_M(17);
      if (EXT) {
        // layer 1 lives in mlp_tc_l1.cu: the partial dZ1 tile leaves as an image; then the next tile's H1 image comes in
        if (leader) mma_commit(&bars[B_DW1]);   // sDZ2 is free once this arrives (dW2, db2 were its last readers)
        __syncwarp();
        // half by half (columns 0..127, then 128..255): the partial dZ1 leaves as soon as the epilogue has written it, and the
        // next tile's H1 half comes in as soon as the copy engine has read that half (the writers fenced for the async proxy)
        const size_t tile = (size_t)pair + (size_t)it * n_pairs;
        uint8_t* dzo = ext.dz1_img + (size_t)rank * ext.img_stride + tile * IMG_TILE_BYTES;
        const uint8_t* nxt = ext.h1_img + (tile + n_pairs) * IMG_TILE_BYTES;
        const bool more = it + 1 < my_tiles;
        ok = mbar_wait(&bars[B_DZ1A], it & 1) && ok;
        if (leader) {
          bulk_s2g(dzo, sH1, IMG_TILE_BYTES / 2);
          bulk_commit();
          bulk_wait_read0();                    // the copy engine has read the half: it may be overwritten
          if (more) {                           // ... by the next tile's H1 columns 0..127, under the epilogue's second dZ1 half
            mbar_arrive_expect_tx(&bars[B_H1L], IMG_TILE_BYTES / 2);
            bulk_g2s(sH1, nxt, IMG_TILE_BYTES / 2, &bars[B_H1L]);
          }
        }
        __syncwarp();
        ok = mbar_wait(&bars[B_DZ1], it & 1) && ok;
        if (leader) {
          bulk_s2g(dzo + IMG_TILE_BYTES / 2, sH1 + IMG_TILE_BYTES / 2, IMG_TILE_BYTES / 2);
          bulk_commit();
          bulk_wait_read0();
          if (more) {
            mbar_arrive_expect_tx(&bars[B_H1M], IMG_TILE_BYTES / 2);
            bulk_g2s(sH1 + IMG_TILE_BYTES / 2, nxt + IMG_TILE_BYTES / 2, IMG_TILE_BYTES / 2, &bars[B_H1M]);
          }
        }
        __syncwarp();
        continue;
      }
      ok = mbar_wait(&bars[B_DZ1A], it & 1) && ok;    // dZ1 columns 0..127 are in place
      fence_after();
      issue_dw1(0);
      if (leader) mma_commit(&bars[B_DW1A]);   // H1 / dZ1 columns 0..127 (and sDZ2) are free once this arrives: the next tile's
      __syncwarp();                            // first H1 half does not have to wait for dW1 rows 128..255
      // next tile's Z1 half 0 (the work buffer is free as soon as the second dH1p half has been read)
      if (it + 1 < my_tiles) {
        ok = mbar_wait(&bars[B_X0 + (xb ^ 1)], ((it + 1) >> 1) & 1) && ok;
        issue_z1(xb ^ 1, 0);
      }
      ok = mbar_wait(&bars[B_DZ1], it & 1) && ok;
      TRACE_M(18);
      fence_after();
      issue_dw1(1);
      if (leader) mma_commit(&bars[B_DW1]);   // sH1 (dZ1), sDZ2 and sX[xb] are free once this arrives
      __syncwarp();
    }
    if (leader) {
      mma_commit(&bars[B_ALL]);
      if (EXT) bulk_wait0();   // every dZ1 image has reached memory before the kernel ends
    }
    __syncwarp();
  } else {
    // =========================== epilogue warps ===========================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(EPI_REGS));
    // warp w: TMEM lane quarter q = w % 4 (rows 32q..32q+31), column slice hh = w / 4 (CW columns) of every 128-column chunk
    const int q = warp & 3, hh = warp >> 2;
    const int row = q * 32 + lane;                      // TMEM lane = tile row
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t t_work = tm + TM_W0 + lane_base + hh * CW;
    const uint32_t peer_part = map_to_cta(sF + F_PART * 4, rank ^ 1u);
    const uint32_t peer_xc = map_to_cta(smem_u32(&bars[B_XC0]), rank ^ 1u);
    float acc_w3[CW / 32][32];   // dw3 partials of this thread's row and CW columns, summed over rows once at the end
#pragma unroll
    for (int j = 0; j < CW; ++j) acc_w3[j >> 5][j & 31] = 0.f;
    float acc_loss = 0.f, acc_b3 = 0.f;
    uint32_t n = 0;
    // this thread's CW columns of the work buffer (use n) -> registers; the buffer is handed back at once
    auto take = [&](uint32_t (&v)[CW]) {
      ok = mbar_wait(&bars[B_DONE], n & 1) && ok;
      fence_after();
#pragma unroll
      for (int c = 0; c < CW / 32; ++c) tmem_ld32(t_work + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&v[c * 32]));
      tmem_ld_wait();
      fence_before();
      warp_arrive(lane, &bars[B_FREE]);
      ++n;
    };
    auto x_store = [&](int j) {
      // pin the fp32 -> fp16 conversion HERE: hoisted to x_issue it would make the loads blocking (a DRAM latency per tile)
#pragma unroll
      for (int i = 0; i < XF / 4; ++i) asm volatile("" : "+f"(xv[i].x), "+f"(xv[i].y), "+f"(xv[i].z), "+f"(xv[i].w));
#pragma unroll
      for (int i = 0; i < (EXT ? 0 : XF / 8); ++i)
        sts128(sX + j * 16384 + tile_chunk_off(TILE, xr, xh * XF + i * 8), pack8(xv[2 * i], xv[2 * i + 1]));
      if (tid < TILE) f32[F_Y + j * 128 + tid] = yv;
      fence_async_smem();
      warp_arrive(lane, &bars[B_X0 + j]);
    };
    if (my_tiles > 0) {
      if (!early_x) x_issue(pair, X, y);
      x_store(0);
    }
    for (int it = 0; it < my_tiles; ++it) {
      const size_t tile = (size_t)pair + (size_t)it * n_pairs;
      const size_t row0 = tile * TILE;
      const bool row_valid = row0 + row < n_rows;
      TRACE_E(20);
      // ---- (b) H1 = relu(Z1 + b1) -> sH1, two halves of 128 columns (the bias is already in the accumulator) ----
      if (EXT && TRAIN && it > 0) { ok = mbar_wait(&bars[B_DW1], (it - 1) & 1) && ok; ok = mbar_wait(&bars[B_DB2], (it - 1) & 1) && ok; }   // dW2 / db2 of the previous tile still read sDZ2
#pragma unroll 1
      for (int h = 0; h < (EXT ? 0 : 2); ++h) {
        uint32_t v[CW];
        TRACE_E(40);
        take(v);
        TRACE_E(43);
        // dW1 of the previous tile still reads sH1 (as dZ1): its rows 0..127 read columns 0..127, its rows 128..255 the rest
        if (TRAIN && it > 0) ok = mbar_wait(&bars[h == 0 ? B_DW1A : B_DW1], (it - 1) & 1) && ok;
        if (TRAIN && it > 0 && h == 1) ok = mbar_wait(&bars[B_DB2], (it - 1) & 1) && ok;   // ... and the db2 warps the previous dZ2 tile
        // (forward only: sH1 was last read by the previous tile's Z2h, whose completion this thread has already seen)
        const int c0 = h * 128 + hh * CW;
#pragma unroll
        for (int g8 = 0; g8 < CW / 8; ++g8) {   // relu on packed halves: one F2FP + one HMNMX2 per two elements
          uint4 o;
          o.x = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 0]), __uint_as_float(v[g8 * 8 + 1])));
          o.y = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 2]), __uint_as_float(v[g8 * 8 + 3])));
          o.z = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 4]), __uint_as_float(v[g8 * 8 + 5])));
          o.w = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 6]), __uint_as_float(v[g8 * 8 + 7])));
          sts128(sH1 + tile_chunk_off(TILE, row, c0 + g8 * 8), o);
        }
        fence_async_smem();
        warp_arrive(lane, &bars[h == 0 ? B_H1A : B_H1]);
      }
      TRACE_E(21);
      // next tile's rows: loads issued here (the warps are about to wait for Z2h), converted and stored at the start of
      // phase (d), where they wait for dH1p.  The other X buffer was last read by tile it-1's dW1 (waited for in (b)).
      if (it + 1 < my_tiles) x_issue(tile + n_pairs, X, y);
      // ---- (c) Z2h: partial output dot, exchange, dZ2h -> sDZ2 ---------------------------------------
      TRACE_E(22);
      const int buf = it & 1;
      const int cz = hh * CW;                              // my columns of this CTA's 128 layer-2 units
      uint32_t z[CW];                                      // Z2h stays in registers across the exchange
      take(z);
      TRACE_E(44);
      float p4[4] = {0.f, 0.f, 0.f, 0.f};   // four independent chains (a single one is CW dependent FMAs per thread)
      {
        // z <- H2 = relu(Z2 + b2) in place; partial dot with w3.  The constants of the next 8 columns are loaded
        // before the current 8 are used: with two warps per scheduler a load-to-use distance of one instruction
        // leaves the shared-memory latency exposed 16 times per pass.
        float4 cb[2][2], cw[2][2];
        auto fetch = [&](int blk, int s) {
          cb[s][0] = lds128f(sF + (F_B2 + cz + blk * 8) * 4); cb[s][1] = lds128f(sF + (F_B2 + cz + blk * 8 + 4) * 4);
          cw[s][0] = lds128f(sF + (F_W3 + cz + blk * 8) * 4); cw[s][1] = lds128f(sF + (F_W3 + cz + blk * 8 + 4) * 4);
        };
        fetch(0, 0);
#pragma unroll
        for (int blk = 0; blk < CW / 8; ++blk) {
          const int s = blk & 1;
          if (blk + 1 < CW / 8) fetch(blk + 1, s ^ 1);
          const float b8[8] = {cb[s][0].x, cb[s][0].y, cb[s][0].z, cb[s][0].w, cb[s][1].x, cb[s][1].y, cb[s][1].z, cb[s][1].w};
          const float w8[8] = {cw[s][0].x, cw[s][0].y, cw[s][0].z, cw[s][0].w, cw[s][1].x, cw[s][1].y, cw[s][1].z, cw[s][1].w};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float r = fmaxf(__uint_as_float(z[blk * 8 + e]) + b8[e], 0.f);
            p4[e & 3] = fmaf(r, w8[e], p4[e & 3]);
            z[blk * 8 + e] = __float_as_uint(r);
          }
        }
      }
      const float part = (p4[0] + p4[1]) + (p4[2] + p4[3]);
      TRACE_E(23);
      f32[F_RP + hh * 128 + row] = part;
      named_bar_sync(1, N_EPI);
      float mine = 0.f;
#pragma unroll
      for (int h = 0; h < CSPLIT; ++h) mine += f32[F_RP + h * 128 + row];     // fixed order: same value in every warp
      // exchange with the peer CTA: each of the 128 rows is stored straight into the peer's shared memory with an
      // asynchronous store that completes 4 transaction bytes on the PEER's mbarrier; nobody else has to synchronise
      if (tid == 0) mbar_arrive_expect_tx(&bars[B_XC0 + buf], 512);
      if (hh == 0) st_async_f32(peer_part + (uint32_t)(buf * 128 + row) * 4, mine, peer_xc + (uint32_t)buf * 8);
      TRACE_E(24);
      // the next tile's rows (loaded during the Z2h wait) are converted and stored while the partial dots cross the pair
      if (it + 1 < my_tiles) x_store((it + 1) & 1);
      ok = mbar_wait(&bars[B_XC0 + buf], (it >> 1) & 1) && ok;
      TRACE_E(25);
      const float other = f32[F_PART + buf * 128 + row];
      const float yhat = (rank == 0 ? mine + other : other + mine) + b3;   // same value on both CTAs
      if (BK == K_PREDICT) {
        // a timed-out wait in this thread poisons its rows (the predict entry point has no status words)
        if (rank == 0 && hh == 0 && row_valid) yhat_out[row0 + row] = ok ? yhat : __int_as_float(0x7fc00000);
        continue;
      }
      float g = 0.f;
      if (row_valid) {
        const float d = yhat - f32[F_Y + buf * 128 + row];
        float lo;
        if (loss_kind == DLAS_LOSS_MSE) {
          lo = d * d;
          g = 2.0f * d;
        } else {  // cnn.py:29-32, CLIP_VALUE = 6
          const float a = fabsf(d);
          lo = fminf(a, 6.0f);
          g = (a <= 6.0f) ? (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) : 0.f;
        }
        if (hh == 0) { acc_loss += lo; acc_b3 += g; }
      }
      if (BK == K_EVAL) continue;
      {
        float4 cw[2][2];
        auto fetch = [&](int blk, int s) {
          cw[s][0] = lds128f(sF + (F_W3 + cz + blk * 8) * 4); cw[s][1] = lds128f(sF + (F_W3 + cz + blk * 8 + 4) * 4);
        };
        fetch(0, 0);
#pragma unroll
        for (int g8 = 0; g8 < CW / 8; ++g8) {
          const int s = g8 & 1;
          if (g8 + 1 < CW / 8) fetch(g8 + 1, s ^ 1);
          const float w8[8] = {cw[s][0].x, cw[s][0].y, cw[s][0].z, cw[s][0].w, cw[s][1].x, cw[s][1].y, cw[s][1].z, cw[s][1].w};
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int j = g8 * 8 + 2 * e;
            const float r0 = __uint_as_float(z[j]), r1 = __uint_as_float(z[j + 1]);       // H2, from the first pass
            acc_w3[j >> 5][j & 31] = fmaf(r0, g, acc_w3[j >> 5][j & 31]);                      // -> dw3
            acc_w3[(j + 1) >> 5][(j + 1) & 31] = fmaf(r1, g, acc_w3[(j + 1) >> 5][(j + 1) & 31]);
            // dZ2 = (H2 > 0) * (g * w3): mask and product on packed halves
            o[e] = mul_h2(gt0_h2(pack_half2(r0, r1)), pack_half2(g * w8[2 * e], g * w8[2 * e + 1]));
          }
          sts128(sDZ2 + tile_chunk_off(TILE, row, cz + g8 * 8), make_uint4(o[0], o[1], o[2], o[3]));
        }
      }
      fence_async_smem();
      warp_arrive(lane, &bars[B_DZ2]);
      TRACE_E(26);
      // ---- (d) dZ1p = relu'(H1) o dH1p in place over H1, two halves of 128 columns ----------------------
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        uint32_t v[CW];
        TRACE_E(50);
        take(v);
        TRACE_E(52);
        const int c0 = h * 128 + hh * CW;
#pragma unroll
        for (int g32 = 0; g32 < CW / 32; ++g32) {
          uint4 m[4];   // loads first: volatile shared-memory accesses keep program order
#pragma unroll
          for (int k = 0; k < 4; ++k) m[k] = lds128(sH1 + tile_chunk_off(TILE, row, c0 + g32 * 32 + k * 8));
          if (g32 == 0) ok = mbar_wait(&bars[B_DW2A + h], it & 1) && ok;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int j = g32 * 32 + k * 8;
            uint32_t o[4];   // dZ1 = (H1 > 0) * dH1p on packed halves
            o[0] = mul_h2(gt0_h2(m[k].x), pack_half2(__uint_as_float(v[j + 0]), __uint_as_float(v[j + 1])));
            o[1] = mul_h2(gt0_h2(m[k].y), pack_half2(__uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
            o[2] = mul_h2(gt0_h2(m[k].z), pack_half2(__uint_as_float(v[j + 4]), __uint_as_float(v[j + 5])));
            o[3] = mul_h2(gt0_h2(m[k].w), pack_half2(__uint_as_float(v[j + 6]), __uint_as_float(v[j + 7])));
            sts128(sH1 + tile_chunk_off(TILE, row, c0 + j), make_uint4(o[0], o[1], o[2], o[3]));
          }
        }
        fence_async_smem();
        warp_arrive(lane, &bars[h == 0 ? B_DZ1A : B_DZ1]);   // half 0 in place: the issuer may start dW1 rows 0..127
      }
      TRACE_E(27);
    }

    // ---- drain the accumulators into this CTA's partial vector -----------------------------------------
    float* my = partials + (size_t)blockIdx.x * PSTRIDE;
    if (BK == K_EVAL && my_tiles > 0) {   // loss sum of this pair's rows: fixed-order sum over the 128 rows
      if (hh == 0) f32[F_RED + row] = acc_loss;
      named_bar_sync(1, N_EPI);
      if (rank == 0 && tid == 0) {
        float s = 0.f;
        for (int i = 0; i < 128; ++i) s += f32[F_RED + i];
        my[N_PARAMS] = s;
      }
    }
    if (TRAIN && my_tiles > 0) {
      TRACE_E(30);
      ok = mbar_wait(&bars[B_ALL], 0) && ok;
      fence_after();
      TRACE_E(31);
      constexpr int DC = 128 / CSPLIT;                    // H1 units (accumulator columns) per thread and half
#pragma unroll 1
      for (int mh = 0; mh < 2; ++mh) {
        // accumulator: lane = layer-2 unit (row), column = H1 unit k  ->  W2[k][rank*128 + row]
        float* dst = my + W2_OFF + (size_t)(mh * 128 + hh * DC) * D2 + rank * 128 + row;
#pragma unroll 1
        for (int c0 = 0; c0 < DC; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tm + TM_DW2 + mh * 128 + lane_base + hh * DC + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) dst[(size_t)(c0 + i) * D2] = __uint_as_float(v[i]);
        }
        // dW1p^T: lane = h1 unit, column = input feature; column 32 = db1
        const int h1 = mh * 128 + row;
        if (EXT) {
          // dW1 / db1 come from l1_dw_tc_kernel
        } else if (hh == 0) {
          uint32_t v[32];
          tmem_ld32(tm + TM_DW1 + mh * 64 + lane_base, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) my[W1_OFF + i * D1 + h1] = __uint_as_float(v[i]);
        } else if (hh == 1) {
          uint32_t w[16];
          tmem_ld16(tm + TM_DW1 + mh * 64 + lane_base + 32, w);
          tmem_ld_wait();
          my[B1_OFF + h1] = __uint_as_float(w[0]);
        }
      }
      // (db2h is written by the db2 warps)
      // dw3h: per-warp column sums (lanes 0..CW-1 hold them) -> add the four row quarters in order
#pragma unroll
      for (int c = 0; c < CW / 32; ++c) {
        const float cs = warp_colsum<32>(acc_w3[c], lane);
        f32[F_RED + q * 128 + hh * CW + c * 32 + lane] = cs;
      }
      named_bar_sync(1, N_EPI);
      if (tid < 128)
        my[W3_OFF + rank * 128 + tid] = (f32[F_RED + tid] + f32[F_RED + 128 + tid]) + (f32[F_RED + 256 + tid] + f32[F_RED + 384 + tid]);
      named_bar_sync(1, N_EPI);
      // loss and db3: fixed-order sum over the 128 rows (only rank 0 contributes them)
      if (hh == 0) { f32[F_RED + row] = acc_loss; f32[F_RED + 128 + row] = acc_b3; }
      named_bar_sync(1, N_EPI);
      if (rank == 0 && tid < 2) {
        float s = 0.f;
        for (int i = 0; i < 128; ++i) s += f32[F_RED + tid * 128 + i];
        my[tid == 0 ? N_PARAMS : B3_OFF] = s;
      }
    }
  }
  TRACE_E(32);
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  if (clk_owner) {   // one writer, launches are stream-ordered
    g_clk_sum[0] += (unsigned long long)(clock64() - clk_c0);
    g_clk_sum[1] += globaltimer_ns() - clk_g0;
  }

  // ---- teardown ----------------------------------------------------------------------------------------
  fence_before();
  __syncthreads();
  // do not exit while the peer may still write our exchange buffer.  Relaxed: a releasing arrive would first wait for the
  // acknowledgement of the 160 KB of partials this CTA has just stored (12 % of all stall samples in ncu's source view)
  cluster_sync_relaxed();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// grad_sum[i] = sum of the partials of the CTAs that own entry i, in CTA order:
//   W1, b1: every active CTA (each holds a partial);  W2[:, c], b2[c], w3[c]: CTAs of rank c/128;
//   b3 and the loss sum: rank-0 CTAs.
// FIN_GRAD:     the sums go to `out` (dlas_mlp_grad; a collective of the caller follows).
// FIN_UPDATE:   the optimiser update of entry i is applied in the same pass (single-GPU step: no all-reduce sits
//               between the gradient and the update), and the batch-mean loss goes to *loss_out.
// FIN_EXCHANGE: data-parallel step over NVLink peer memory.  Every block PUSHES its 64 sums into every peer's receive
//               buffer as plain 4-byte floats, then polls ITS OWN receive buffer -- local memory -- until the peers' words
//               are there, adds the `world` values in rank order (same bits on every rank), applies the update, and the
//               last block to finish advances the device-resident cursor and iteration.  A word is its own flag: the
//               buffer starts as zeros, +0.0 travels as -0.0 (and is read back as +0.0), so "non-zero bits" means
//               "arrived"; the reader clears every word it has consumed.  100 % of the NVLink bytes are payload (the
//               {value, epoch} pairs of the first version spent half of them on the epoch, and at 8 ranks the poll was
//               bound by the 4.2 MB each rank receives per step), and no access wider than a 4-byte word has to arrive
//               whole.  One NVLink one-way latency per step, no flags, no separate all-reduce / update / advance
//               kernels.  Receive buffer of a rank: [2 parities][world sources][recv_stride words]; parity =
//               iteration & 1: a rank one step ahead writes the other parity, and it cannot get two steps ahead -- it
//               needs this rank's next words first, which leave after this step's kernel (and its clearing) has ended.
enum { FIN_GRAD = 0, FIN_UPDATE = 1, FIN_EXCHANGE = 2 };
struct FinDp {
  const unsigned long long* recv_bases;   // device array: address of every rank's receive buffer as mapped HERE
  int rank, world;
  size_t recv_stride;                     // 8-byte words per (parity, source)
  dlas_opt_cfg cfg;
  dlas_step_state* state;                 // device-resident cursor / iteration (graph replay), or NULL: host-driven step
  long long it_host;                      //   ... whose iteration is this and whose scalars are the kernel's `opt` argument
  long long n_batches;
  float* loss_ring;
  long long ring_len;
};
#ifdef DLAS_TC_TRACE
// trace build: per step (ring of 64) globaltimer stamps of block 0 and the last block: after the dependency wait, after
// the reduction + push, after the poll; read back by tools/fin_trace.py through dlas_debug_fin_trace
__device__ unsigned long long g_fin_trace[64][2][4];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define FIN_STAMP(k) do { if (threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) g_fin_trace[it & 63][blockIdx.x ? 1 : 0][k] = gtime(); } while (0)
#else
#define FIN_STAMP(k) do { } while (0)
#endif
__device__ unsigned g_fin_done = 0;       // blocks of the current FIN_EXCHANGE launch that have finished (reset by the last)

// One 4-byte word per value; all-zero bits = "not arrived".  +0.0 is sent as -0.0 and every zero is read as +0.0.
__device__ __forceinline__ void st_push(unsigned long long addr, float v) {
  unsigned w = __float_as_uint(v);
  if (w == 0u) w = 0x80000000u;
  asm volatile("st.relaxed.sys.global.b32 [%0], %1;" ::"l"(addr), "r"(w) : "memory");
}
__device__ __forceinline__ unsigned ld_poll(unsigned long long addr) {
  unsigned w;
  asm volatile("ld.relaxed.sys.global.b32 %0, [%1];" : "=r"(w) : "l"(addr) : "memory");
  return w;
}
__device__ __forceinline__ void st_clear(unsigned long long addr) {
  asm volatile("st.relaxed.sys.global.b32 [%0], %1;" ::"l"(addr), "r"(0u) : "memory");
}

// the updated parameter i as fp16 at its place in the weight image (see IMG_*); b2, w3, b3 are not part of the operand tiles
__device__ __forceinline__ void img_store(uint8_t* img, int i, float p) {
  uint32_t off;
  if (i < B1_OFF) {
    const int r = i / D1, c = i - r * D1;
    off = tile_chunk_off(K1, r, c & ~7) + (uint32_t)(c & 7) * 2;
  } else if (i < W2_OFF) {
    const int c = i - B1_OFF;
    off = tile_chunk_off(K1, D0, c & ~7) + (uint32_t)(c & 7) * 2;        // the bias row of [W1 ; b1 ; 0]
  } else if (i < B2_OFF) {
    const int k = (i - W2_OFF) / D2, c2 = (i - W2_OFF) - k * D2, cc = c2 & 127;
    off = IMG_W1_BYTES + (uint32_t)(c2 >> 7) * IMG_W2_BYTES + tile_chunk_off(D1, k, cc & ~7) + (uint32_t)(cc & 7) * 2;
  } else {
    return;
  }
  *reinterpret_cast<__half*>(img + off) = __float2half_rn(p);
}

constexpr int RG = 4;   // thread groups per block of 64 entries (8 measured no faster)
constexpr int FIN_MAXW = 8;
template <int MODE>
__global__ void __launch_bounds__(64 * RG, 8)   // 8 blocks per SM: the whole grid is resident
tc_reduce_kernel(const float* __restrict__ partials, int n_active_pairs, unsigned* __restrict__ err_flag,
                 unsigned call_id, float* __restrict__ out, OptScalars opt, float grad_scale, float* __restrict__ params,
                 float* __restrict__ m, float* __restrict__ v, float* __restrict__ loss_out,
                 const long long* __restrict__ state, size_t parity_stride, const FinDp dp, uint8_t* __restrict__ img_out,
                 const int pshift) {
  // pshift != 0: the wide-input network d0-256-256-1 (K_TRAIN_EXT partials).  Entries below W2_OFF belong to layer 1, which
  // l1_reduce_kernel finishes; entry i >= W2_OFF is parameter i + pshift of the real vector (params / m / v / out are its base)
  // 64 consecutive entries per block; the list of contributing CTAs is cut into RG contiguous parts summed by RG
  // thread groups (more loads in flight: the pass is latency-bound), then added in quarter order -> still deterministic
  __shared__ float red[RG - 1][64];
  __shared__ OptScalars o_dev;
  __shared__ int failed;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // next gradient kernel: set-up may overlap this pass
  asm volatile("griddepcontrol.wait;" ::: "memory");   // the gradient kernel has completed; its partials are visible
  const int i = blockIdx.x * 64 + (threadIdx.x & 63), grp = threadIdx.x >> 6;
  if (pshift != 0 && i < W2_OFF) return;     // whole blocks: W2_OFF is a multiple of 64
  const int pi = i + pshift;
  const bool live = i <= N_PARAMS;
  long long it = 0, cur = 0;
  if (MODE == FIN_EXCHANGE && dp.state == nullptr) {
    it = dp.it_host;
    if (threadIdx.x == 0) { o_dev = opt; failed = 0; }
  } else if (MODE == FIN_EXCHANGE) {
    it = __ldcg(&dp.state->iteration);
    cur = __ldcg(&dp.state->cursor);
    if (threadIdx.x == 0) {   // step-dependent scalars from the device-resident iteration (as opt_at_kernel, optim.cu)
      double lr_t = (double)dp.cfg.lr / (1.0 + (double)dp.cfg.decay * (double)it);
      if (dp.cfg.kind == DLAS_OPT_ADAM) {
        const double t = (double)it + 1.0;
        lr_t = lr_t * sqrt(1.0 - pow((double)dp.cfg.beta2, t)) / (1.0 - pow((double)dp.cfg.beta1, t));
      }
      o_dev.kind = dp.cfg.kind; o_dev.lr_t = (float)lr_t; o_dev.momentum = dp.cfg.momentum; o_dev.nesterov = dp.cfg.nesterov;
      o_dev.beta1 = dp.cfg.beta1; o_dev.beta2 = dp.cfg.beta2; o_dev.eps = dp.cfg.eps;
      failed = 0;
    }
  }
  if (MODE == FIN_EXCHANGE) FIN_STAMP(0);
  int first = 0, step = 1;
  if (i >= W2_OFF && i < B2_OFF) { first = ((i - W2_OFF) % D2) / 128; step = 2; }
  else if (i >= B2_OFF && i < W3_OFF) { first = (i - B2_OFF) / 128; step = 2; }
  else if (i >= W3_OFF && i < B3_OFF) { first = (i - W3_OFF) / 128; step = 2; }
  else if (i >= B3_OFF) { first = 0; step = 2; }
  const int n_src = (2 * n_active_pairs - first + step - 1) / step;        // contributing CTAs
  const int per = (n_src + RG - 1) / RG, k0 = grp * per, k1 = min(n_src, k0 + per);
  float s = 0.f;
  if (live) {
#pragma unroll 8
    for (int k = k0; k < k1; ++k) s += partials[(size_t)(first + k * step) * PSTRIDE + i];
  }
  if (grp > 0) red[grp - 1][threadIdx.x & 63] = s;
  __syncthreads();
  if (grp == 0 && live) {
#pragma unroll
    for (int g2 = 0; g2 < RG - 1; ++g2) s += red[g2][threadIdx.x];
    // an mbarrier wait timed out inside this call's tcgen05 kernel: fail loudly, never return numbers
    if (*err_flag == call_id) s = __int_as_float(0x7fc00000);
    // fp16 operands: a row value, activation or back-propagated term beyond +-65504 became inf in an operand tile, and
    // inf (or inf * 0 = NaN) always reaches a gradient sum.  A non-finite sum is reported through the status words
    // (dlas_mlp_status -> DLAS_ERR_RANGE); a diverged training raises the same flag.
    if (!(fabsf(s) <= 3.0e38f)) err_flag[1] = call_id;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) err_flag[2] = call_id;   // the launch the status words describe
  if (MODE != FIN_EXCHANGE) {
    if (grp > 0 || !live) return;
    if (MODE == FIN_GRAD) {
      out[(state && (state[1] & 1)) ? parity_stride + pi : pi] = s;   // state[1] = iteration: double-buffered bucket
    } else if (i == N_PARAMS) {
      if (loss_out) *loss_out = s * grad_scale;
    } else {
      float p = params[pi], mi = m[pi], vi = (opt.kind == 0) ? 0.f : v[pi];
      opt_update(opt, s * grad_scale, p, mi, vi);
      params[pi] = p;
      m[pi] = mi;
      if (opt.kind != 0) v[pi] = vi;
      if (img_out) {
        img_store(img_out, i, p);
        if (i < (K1 - D0 - 1) * D1)      // rows 33..47 of the [W1 ; b1 ; 0] tile are zero (the workspace is not initialised)
          *reinterpret_cast<__half*>(img_out + tile_chunk_off(K1, D0 + 1 + i / D1, (i % D1) & ~7) + (uint32_t)(i & 7) * 2) = __float2half_rn(0.f);
      }
    }
    return;
  }
  // ---- FIN_EXCHANGE -------------------------------------------------------------------------------------------
  if (grp == 0 && live) {
    const size_t slot = (size_t)(it & 1) * dp.world * dp.recv_stride + (size_t)i;   // + source * recv_stride
    // push my sum into every peer's buffer, at source index = my rank (64 threads x 4 bytes: one 256-byte row per peer)
    for (int r = 0; r < dp.world; ++r)
      if (r != dp.rank) st_push(dp.recv_bases[r] + (slot + (size_t)dp.rank * dp.recv_stride) * 4, s);
    FIN_STAMP(1);
    // collect the peers' sums from my own buffer; every load first, spinning only on words that have not arrived
    float q[FIN_MAXW];
    const unsigned long long mine = dp.recv_bases[dp.rank];
    bool ok = true;
#pragma unroll
    for (int r = 0; r < FIN_MAXW; ++r) {
      q[r] = 0.f;
      if (r < dp.world && r != dp.rank) {
        const unsigned long long a = mine + (slot + (size_t)r * dp.recv_stride) * 4;
        unsigned w = ld_poll(a);
        for (unsigned spin = 0; w == 0u && spin < (1u << 21); ++spin) {
          __nanosleep(20);
          w = ld_poll(a);
        }
        if (w == 0u) ok = false;
        else st_clear(a);                                   // consumed: the slot is free for iteration it + 2
        q[r] = (w == 0x80000000u) ? 0.f : __uint_as_float(w);
      } else if (r == dp.rank) {
        q[r] = (s == 0.f) ? 0.f : s;                        // the same canonical zero the peers read
      }
    }
    FIN_STAMP(2);
    if (!ok) failed = 1;
    float sum = 0.f;
#pragma unroll
    for (int r = 0; r < FIN_MAXW; ++r) sum += q[r];          // rank order: bit-identical on every rank (absent ranks add +0)
    if (!ok) sum = __int_as_float(0x7fc00000);
    if (!(fabsf(sum) <= 3.0e38f)) err_flag[1] = call_id;   // a peer's sum was non-finite, or the exchange timed out
    if (i == N_PARAMS) {
      if (dp.loss_ring) dp.loss_ring[it % dp.ring_len] = sum * grad_scale;
    } else {
      float p = params[i], mi = m[i], vi = (o_dev.kind == 0) ? 0.f : v[i];
      opt_update(o_dev, sum * grad_scale, p, mi, vi);
      params[i] = p;
      m[i] = mi;
      if (o_dev.kind != 0) v[i] = vi;
      if (img_out) {
        img_store(img_out, i, p);
        if (i < (K1 - D0 - 1) * D1)
          *reinterpret_cast<__half*>(img_out + tile_chunk_off(K1, D0 + 1 + i / D1, (i % D1) & ~7) + (uint32_t)(i & 7) * 2) = __float2half_rn(0.f);
      }
    }
  }
  FIN_STAMP(3);
  // the last block to finish moves the cursor: every block has read the state by then
  if (dp.state == nullptr) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&g_fin_done, 1u) == gridDim.x - 1) {
      g_fin_done = 0;
      dp.state->cursor = (cur + 1) % dp.n_batches;
      dp.state->iteration = it + 1;
    }
  }
}


// validation / test loss (K_EVAL launches): sum of the rank-0 CTAs' loss partials in CTA order
__global__ void __launch_bounds__(32)
tc_loss_reduce_kernel(const float* __restrict__ partials, int n_active_pairs, unsigned* __restrict__ err_flag, unsigned call_id,
                      float* __restrict__ loss_sum, int accumulate) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int p = 0; p < n_active_pairs; ++p) s += partials[(size_t)(2 * p) * PSTRIDE + N_PARAMS];
    if (*err_flag == call_id) s = __int_as_float(0x7fc00000);
    if (!(fabsf(s) <= 3.0e38f)) err_flag[1] = call_id;
    err_flag[2] = call_id;
    *loss_sum = accumulate ? *loss_sum + s : s;   // accumulate: a later 65,536-row part of the same call
  }
}

int n_ctas() { return (sm_count() / 2) * 2; }
unsigned g_call_id = 0x5eed0000u;

size_t ws_partials_bytes() { return (size_t)n_ctas() * PSTRIDE * sizeof(float); }
size_t ws_need() { return ws_partials_bytes() + 256 + IMG_BYTES; }
uint8_t* ws_image(void* workspace) { return (uint8_t*)workspace + ws_partials_bytes() + 256; }

template <int KIND = K_TRAIN>
int launch_main(int loss, const float* params, const float* X, const float* y, size_t n_rows, void* workspace,
                size_t workspace_bytes, cudaStream_t st, float** partials_out, unsigned** err_out, int* n_active_pairs,
                const long long* cursor = nullptr, bool rows_ready = false, size_t ldx = 0, const uint8_t* w_image = nullptr,
                float* yhat = nullptr, const ExtIo& ext = ExtIo{}) {
  if (ldx == 0) ldx = D0;
  DLAS_REQUIRE(ldx % 4 == 0, "mlp (tc): the row stride of X must be a multiple of 4 floats (16-byte row loads), got %zu", ldx);
  size_t need = ws_need();
  constexpr int BK = KIND >= K_TRAIN_EXT ? KIND - K_TRAIN_EXT : KIND;
  if (BK != K_PREDICT && (workspace == nullptr || workspace_bytes < need)) {
    set_error("mlp (tc): workspace of %zu bytes given, %zu needed", workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  DLAS_REQUIRE((((uintptr_t)X) & 15) == 0, "mlp (tc): X must be 16-byte aligned");
  float* partials = (BK == K_PREDICT) ? nullptr : (float*)workspace;
  unsigned* err_flag = (BK == K_PREDICT && workspace == nullptr) ? nullptr : (unsigned*)((char*)workspace + ws_partials_bytes());
  static bool attr_set[64] = {};   // the attribute is per device
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(mlp_tc_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int grid = n_ctas();
  ++g_call_id;   // the error word is never cleared: a kernel that times out writes THIS call's id into it
  int pi = prof_begin(DLAS_PROF_MLP_GRAD, st);
  long long* trace = nullptr;
#ifdef DLAS_TC_TRACE
  const char* trace_path = getenv("DLAS_TC_TRACE");   // trace library only: this branch allocates and synchronises
  if (trace_path) {
    DLAS_CUDA(cudaMalloc(&trace, 1536 * sizeof(long long)));
    DLAS_CUDA(cudaMemsetAsync(trace, 0, 1536 * sizeof(long long), st));
  }
#endif
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DLAS_CUDA(cudaLaunchKernelEx(&cfg, mlp_tc_kernel<KIND>, loss, params, X, y, n_rows, partials, err_flag, g_call_id, trace, cursor,
                                 rows_ready ? 1 : 0, ldx, w_image, yhat, ext));
  }
  prof_end(pi, st);
  DLAS_LAUNCH_CHECK();
#ifdef DLAS_TC_TRACE
  if (trace_path) {
    long long h[1536];
    DLAS_CUDA(cudaMemcpyAsync(h, trace, sizeof(h), cudaMemcpyDeviceToHost, st));
    DLAS_CUDA(cudaStreamSynchronize(st));
    cudaFree(trace);
    FILE* f = fopen(trace_path, "w");
    if (f) {
      for (int i = 0; i < 1536; ++i)
        if (h[i]) fprintf(f, "%s %lld %lld\n", i < 512 ? "E" : (i < 1024 ? "M" : "F"), h[i] >> 48, h[i] & 0xFFFFFFFFFFFFll);
      fclose(f);
    }
  }
#endif
  size_t n_tiles = (n_rows + TILE - 1) / TILE;
  if (n_active_pairs) *n_active_pairs = (int)(n_tiles < (size_t)(grid / 2) ? n_tiles : (size_t)(grid / 2));
  if (partials_out) *partials_out = partials;
  if (err_out) *err_out = err_flag;
  return DLAS_OK;
}

template <int MODE>
int launch_reduce(cudaStream_t st, const float* partials, int nap, unsigned* err, float* out, OptScalars o,
                  float grad_scale, float* params, float* m, float* v, float* loss_out, const long long* state = nullptr,
                  size_t parity_stride = 0, const FinDp& dp = FinDp{}, uint8_t* img_out = nullptr, int pshift = 0) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((N_PARAMS + 1 + 63) / 64);
  cfg.blockDim = dim3(64 * RG);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, tc_reduce_kernel<MODE>, partials, nap, err, g_call_id, out, o, grad_scale, params, m, v,
                               loss_out, state, parity_stride, dp, img_out, pshift));
  return DLAS_OK;
}

}  // namespace

int make_opt_scalars(const dlas_opt_cfg* cfg, uint64_t iteration, OptScalars& o);   // optim.cu

bool tc_supported(const int* dims, int n_layers) {
  return n_layers == 3 && dims[0] == D0 && dims[1] == D1 && dims[2] == D2 && dims[3] == 1;
}

size_t tc_workspace_bytes(const int* dims, int n_layers) {
  size_t tc = ws_need();   // partials + status words + fp16 weight image
  size_t simt = simt_workspace_bytes(dims, n_layers);             // eval / predict share the workspace
  return tc > simt ? tc : simt;
}

// Status words at the end of the workspace (written by the kernels, never cleared): [0] id of the last launch in which an
// mbarrier / exchange wait timed out, [1] id of the last launch that produced a non-finite gradient sum, [2] id of the
// last launch.  A word describes the most recent launch iff it equals word [2]; NaN parameters keep every later launch
// non-finite, so looking at the latest launch is enough.
int tc_status(const void* workspace, size_t workspace_bytes, cudaStream_t st) {
  const size_t off = (size_t)n_ctas() * PSTRIDE * sizeof(float);
  if (workspace == nullptr || workspace_bytes < off + 256) {
    set_error("dlas_mlp_status: workspace of %zu bytes given, %zu needed", workspace_bytes, off + 256);
    return DLAS_ERR_WORKSPACE;
  }
  unsigned w[4] = {0u, 0u, 0u, 0u};
  DLAS_CUDA(cudaMemcpyAsync(w, (const char*)workspace + off, sizeof(w), cudaMemcpyDeviceToHost, st));
  DLAS_CUDA(cudaStreamSynchronize(st));
  if (w[2] <= 0x5eed0000u || w[2] > g_call_id) return DLAS_OK;   // uninitialised memory: no launch of this process has used the workspace
  if (w[0] == w[2]) {
    set_error("tensor-core MLP step: a wait inside the kernels timed out (a peer rank did not deliver its sums, or the "
              "GPU was pre-empted); the step's results were replaced by NaN");
    return DLAS_ERR_TIMEOUT;
  }
  if (w[1] == w[2]) {
    set_error("tensor-core MLP step: non-finite gradient sum.  DLAS_MODE_TC keeps rows, activations and back-propagated "
              "terms as fp16 operands (|value| <= 65504, 11-bit significand); a value beyond that range became inf, or "
              "the training diverged.  Rescale the feature / target columns or use DLAS_MODE_FP32");
    return DLAS_ERR_RANGE;
  }
  return DLAS_OK;
}

// forward + loss + backward: grad_sum[0..n) gradient sums, grad_sum[n] loss sum
int tc_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y,
            size_t n_rows, float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st,
            const long long* cursor, size_t parity_stride, size_t ldx) {
  if (!tc_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC supports the 32-256-256-1 network only");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)N_PARAMS + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  float* partials; unsigned* err; int nap;
  int rc = launch_main(loss, params, X, y, n_rows, workspace, workspace_bytes, st, &partials, &err, &nap, cursor, false, ldx);
  if (rc != DLAS_OK) return rc;
  rc = launch_reduce<FIN_GRAD>(st, partials, nap, err, grad_sum, OptScalars{}, 0.f, nullptr, nullptr, nullptr, nullptr,
                            parity_stride ? cursor : nullptr, parity_stride);
  if (rc != DLAS_OK) return rc;
  return DLAS_OK;
}

// model.predict / test_on_batch on the tensor cores (forward-only kind of the same kernel; no scratch, no second launch)
int tc_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows, float* yhat, cudaStream_t st,
               size_t ldx) {
  if (!tc_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC supports the 32-256-256-1 network only");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) return DLAS_OK;
  return launch_main<K_PREDICT>(DLAS_LOSS_MSE, params, X, nullptr, n_rows, nullptr, 0, st, nullptr, nullptr, nullptr, nullptr, false, ldx,
                                nullptr, yhat);
}

// validation / test loss sum on the tensor cores: forward + loss kind, then the sum of the pairs' loss partials
int tc_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
                 float* loss_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx) {
  if (!tc_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC supports the 32-256-256-1 network only");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
    return DLAS_OK;
  }
  float* partials; unsigned* err; int nap;
  int rc = launch_main<K_EVAL>(loss, params, X, y, n_rows, workspace, workspace_bytes, st, &partials, &err, &nap, nullptr, false, ldx);
  if (rc != DLAS_OK) return rc;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(32);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, tc_loss_reduce_kernel, (const float*)partials, nap, err, g_call_id, loss_sum, 0));
  return DLAS_OK;
}

// one whole single-GPU step: gradient kernel, then reduction fused with the optimiser update
int tc_fit_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                const float* X, const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace,
                size_t workspace_bytes, cudaStream_t st, bool chained, size_t ldx, bool more_follow) {
  // chained: the kernel before this step in the stream is the previous step's finishing kernel of the same call (which left
  // the fp16 weight image in the workspace); more_follow: another step of the same call follows (leave the image for it)
  if (!tc_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC supports the 32-256-256-1 network only");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  OptScalars o;
  int rc = make_opt_scalars(cfg, iteration, o);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(o.kind == DLAS_OPT_SGD || v != nullptr, "Adam needs the second-moment buffer v");
  float* partials; unsigned* err; int nap;
  rc = launch_main(loss, params, X, y, n_rows, workspace, workspace_bytes, st, &partials, &err, &nap, nullptr, chained, ldx,
                   chained ? ws_image(workspace) : nullptr);
  if (rc != DLAS_OK) return rc;
  rc = launch_reduce<FIN_UPDATE>(st, partials, nap, err, nullptr, o, 1.0f / (float)n_rows, params, m, v, loss_out, nullptr, 0, FinDp{},
                                 more_follow ? ws_image(workspace) : nullptr);
  if (rc != DLAS_OK) return rc;
  return DLAS_OK;
}


// ---- d0-256-256-1 (64 <= d0 <= 240, d0 % 8 == 0: the dynamic ANN's 24 x 8 lag window is d0 = 192) --------------------
// Five launches per step, chained by programmatic dependent launch:
//   l1_fwd (H1 tile images) -> mlp_tc_kernel<K_TRAIN_EXT> (layers 2, 3; partial dZ1 images) -> tc_reduce (W2.. [update])
//   -> l1_dw (dW1, db1 partials) -> l1_reduce ([update of] W1, b1)
// Workspace: [pair partials | status words | weight image] as for 32-256-256-1, then the H1 images, the two partial dZ1
// image sets and layer 1's partials.  The images bound the rows of one call (TCX_MAX_ROWS).
int l1_fwd(int d0, const float* params, const float* X, size_t ldx, size_t n_rows, uint8_t* h1_img, unsigned* err_flag, unsigned call_id,
           cudaStream_t st);
int l1_dw(int d0, const float* X, size_t ldx, size_t n_rows, const uint8_t* dz1_img, size_t img_stride, float* partials,
          unsigned* err_flag, unsigned call_id, int mode, float* grad_sum, const OptScalars& opt, float grad_scale, float* params,
          float* m, float* v, cudaStream_t st);
size_t l1_partial_bytes(int d0);

constexpr size_t TCX_MAX_ROWS = 65536;
namespace {
size_t tcx_base() { return (ws_need() + 1023) & ~(size_t)1023; }
size_t tcx_img_bytes() { return (TCX_MAX_ROWS / TILE) * (size_t)IMG_TILE_BYTES; }
}  // namespace

bool tcx_supported(const int* dims, int n_layers) {
  return n_layers == 3 && dims[1] == D1 && dims[2] == D2 && dims[3] == 1 && dims[0] >= 64 && dims[0] <= 240 && dims[0] % 8 == 0;
}

size_t tcx_workspace_bytes(const int* dims, int n_layers) {
  size_t tc = tcx_base() + 3 * tcx_img_bytes() + l1_partial_bytes(dims[0]);
  size_t simt = simt_workspace_bytes(dims, n_layers);             // eval / predict run on the fp32 kernels
  return tc > simt ? tc : simt;
}

// mode 0: gradient sums -> grad_sum (dlas_mlp_grad);  mode 1: one whole step (update fused into the two reductions)
static int tcx_run(const int* dims, int loss, int mode, const float* params_c, float* params, float* m, float* v, const OptScalars& o,
                   float grad_scale, const float* X, size_t ldx, const float* y, size_t n_rows, float* grad_sum, float* loss_out,
                   void* workspace, size_t workspace_bytes, cudaStream_t st, bool chained, bool more_follow) {
  const int d0 = dims[0];
  const size_t need = tcx_base() + 3 * tcx_img_bytes() + l1_partial_bytes(d0);
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("mlp (tc, %d-256-256-1): workspace of %zu bytes given, %zu needed", d0, workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  if (n_rows > TCX_MAX_ROWS) {
    set_error("mlp (tc, %d-256-256-1): %zu rows in one call; the tensor-core path of this shape takes at most %zu (its "
              "workspace holds the layer-1 tiles of one call) -- split the batch or use DLAS_MODE_FP32", d0, n_rows, TCX_MAX_ROWS);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  DLAS_REQUIRE(ldx % 4 == 0, "mlp (tc): the row stride of X must be a multiple of 4 floats (16-byte row loads), got %zu", ldx);
  DLAS_REQUIRE((((uintptr_t)X) & 15) == 0, "mlp (tc): X must be 16-byte aligned");
  uint8_t* base = (uint8_t*)workspace + tcx_base();
  ExtIo ext;
  ext.h1_img = base;
  ext.dz1_img = base + tcx_img_bytes();
  ext.img_stride = tcx_img_bytes();
  float* l1_part = (float*)(base + 3 * tcx_img_bytes());
  unsigned* err = (unsigned*)((char*)workspace + ws_partials_bytes());
  const int pshift = (d0 - D0) * D1;
  int rc = l1_fwd(d0, params_c, X, ldx, n_rows, base, err, g_call_id + 1, st);   // launch_main gives the step its id
  if (rc != DLAS_OK) return rc;
  float* partials; int nap;
  rc = launch_main<K_TRAIN_EXT>(loss, params_c + pshift, X, y, n_rows, workspace, workspace_bytes, st, &partials, &err, &nap, nullptr, false,
                                ldx, chained ? ws_image(workspace) : nullptr, nullptr, ext);
  if (rc != DLAS_OK) return rc;
  if (mode == 0)
    rc = launch_reduce<FIN_GRAD>(st, partials, nap, err, grad_sum, OptScalars{}, 0.f, nullptr, nullptr, nullptr, nullptr, nullptr, 0,
                                 FinDp{}, nullptr, pshift);
  else
    rc = launch_reduce<FIN_UPDATE>(st, partials, nap, err, nullptr, o, grad_scale, params, m, v, loss_out, nullptr, 0, FinDp{},
                                   more_follow ? ws_image(workspace) : nullptr, pshift);
  if (rc != DLAS_OK) return rc;
  return l1_dw(d0, X, ldx, n_rows, ext.dz1_img, ext.img_stride, l1_part, err, g_call_id, mode, grad_sum, o, grad_scale, params, m, v, st);
}

int tcx_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
             float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx) {
  if (!tcx_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC: not a d0-256-256-1 network");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)N_PARAMS + (size_t)(dims[0] - D0) * D1 + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  return tcx_run(dims, loss, 0, params, nullptr, nullptr, nullptr, OptScalars{}, 0.f, X, ldx, y, n_rows, grad_sum, nullptr, workspace,
                 workspace_bytes, st, false, false);
}

int tcx_fit_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* X,
                 const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace, size_t workspace_bytes,
                 cudaStream_t st, bool chained, size_t ldx, bool more_follow) {
  if (!tcx_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC: not a d0-256-256-1 network");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  OptScalars o;
  int rc = make_opt_scalars(cfg, iteration, o);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(o.kind == DLAS_OPT_SGD || v != nullptr, "Adam needs the second-moment buffer v");
  return tcx_run(dims, loss, 1, params, params, m, v, o, 1.0f / (float)n_rows, X, ldx, y, n_rows, nullptr, loss_out, workspace,
                 workspace_bytes, st, chained, more_follow);
}

// model.predict / validation loss of d0-256-256-1 on the tensor cores: per 65,536-row part l1_fwd (H1 images) and the
// forward-only kinds of the pair kernel; the workspace holds the images, so these take one (dlas_mlp_predict_ws)
int tcx_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows, float* yhat, void* workspace,
                size_t workspace_bytes, cudaStream_t st, size_t ldx) {
  if (!tcx_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC: not a d0-256-256-1 network");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  const int d0 = dims[0];
  const size_t need = tcx_base() + tcx_img_bytes();
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("mlp predict (tc, %d-256-256-1): workspace of %zu bytes given, %zu needed", d0, workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  DLAS_REQUIRE(ldx % 4 == 0, "mlp (tc): the row stride of X must be a multiple of 4 floats (16-byte row loads), got %zu", ldx);
  ExtIo ext;
  ext.h1_img = (uint8_t*)workspace + tcx_base();
  ext.dz1_img = nullptr;
  ext.img_stride = 0;
  unsigned* err = (unsigned*)((char*)workspace + ws_partials_bytes());
  for (size_t r0 = 0; r0 < n_rows; r0 += TCX_MAX_ROWS) {
    const size_t n = n_rows - r0 < TCX_MAX_ROWS ? n_rows - r0 : TCX_MAX_ROWS;
    int rc = l1_fwd(d0, params, X + r0 * ldx, ldx, n, (uint8_t*)workspace + tcx_base(), err, g_call_id + 1, st);
    if (rc != DLAS_OK) return rc;
    rc = launch_main<K_PREDICT_EXT>(DLAS_LOSS_MSE, params + (size_t)(d0 - D0) * D1, X + r0 * ldx, nullptr, n, workspace, workspace_bytes, st,
                                    nullptr, nullptr, nullptr, nullptr, false, ldx, nullptr, yhat + r0, ext);
    if (rc != DLAS_OK) return rc;
  }
  return DLAS_OK;
}

int tcx_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
                  float* loss_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx) {
  if (!tcx_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC: not a d0-256-256-1 network");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
    return DLAS_OK;
  }
  const int d0 = dims[0];
  const size_t need = tcx_base() + tcx_img_bytes();
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("mlp eval_loss (tc, %d-256-256-1): workspace of %zu bytes given, %zu needed", d0, workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  DLAS_REQUIRE(ldx % 4 == 0, "mlp (tc): the row stride of X must be a multiple of 4 floats (16-byte row loads), got %zu", ldx);
  ExtIo ext;
  ext.h1_img = (uint8_t*)workspace + tcx_base();
  ext.dz1_img = nullptr;
  ext.img_stride = 0;
  for (size_t r0 = 0; r0 < n_rows; r0 += TCX_MAX_ROWS) {
    const size_t n = n_rows - r0 < TCX_MAX_ROWS ? n_rows - r0 : TCX_MAX_ROWS;
    unsigned* err = (unsigned*)((char*)workspace + ws_partials_bytes());
    int rc = l1_fwd(d0, params, X + r0 * ldx, ldx, n, (uint8_t*)workspace + tcx_base(), err, g_call_id + 1, st);
    if (rc != DLAS_OK) return rc;
    float* partials; int nap;
    rc = launch_main<K_EVAL_EXT>(loss, params + (size_t)(d0 - D0) * D1, X + r0 * ldx, y + r0, n, workspace, workspace_bytes, st, &partials, &err,
                                 &nap, nullptr, false, ldx, nullptr, nullptr, ext);
    if (rc != DLAS_OK) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(32);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DLAS_CUDA(cudaLaunchKernelEx(&cfg, tc_loss_reduce_kernel, (const float*)partials, nap, err, g_call_id, loss_sum, r0 > 0 ? 1 : 0));
  }
  return DLAS_OK;
}

// one whole data-parallel step in two launches (forward/backward, then reduce + exchange + update + advance)
size_t tc_dp_recv_words(int world) { return 2 * (size_t)world * (((size_t)N_PARAMS + 1 + 63) & ~(size_t)63); }
int tc_dp_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v,
               const float* X_base, const float* y_base, size_t batch_rows, dlas_step_state* state,
               int64_t cursor0, int64_t iteration0, const int64_t* batch_order_host, const uint64_t* recv_bases_dev, int rank,
               int world, float grad_scale, int64_t n_batches, float* loss_ring, int64_t ring_len, int64_t n_steps,
               void* workspace, size_t workspace_bytes, cudaStream_t st) {
  if (!tc_supported(dims, n_layers)) {
    set_error("DLAS_MODE_TC supports the 32-256-256-1 network only");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  DLAS_REQUIRE(cfg && params && m && recv_bases_dev && batch_rows > 0 && n_batches >= 1, "dlas_mlp_dp_step_at: bad arguments");
  DLAS_REQUIRE(state || (cursor0 >= 0 && cursor0 < n_batches && iteration0 >= 0), "dlas_mlp_dp_step_at: host-driven call needs cursor0 in [0, n_batches) and iteration0 >= 0");
  DLAS_REQUIRE(cfg->kind == DLAS_OPT_SGD || (cfg->kind == DLAS_OPT_ADAM && v), "dlas_mlp_dp_step_at: bad optimiser / missing v");
  DLAS_REQUIRE(world >= 1 && world <= FIN_MAXW && rank >= 0 && rank < world, "dlas_mlp_dp_step_at: rank %d / world %d", rank, world);
  DLAS_REQUIRE(loss_ring == nullptr || ring_len >= 1, "dlas_mlp_dp_step_at: ring_len < 1");
  DLAS_REQUIRE(n_steps >= 0, "dlas_mlp_dp_step_at: n_steps < 0");
  DLAS_REQUIRE(!(state && batch_order_host), "dlas_mlp_dp_step_at: batch_order_host needs the host-driven flavour (state_dev = NULL)");
  FinDp dp;
  dp.recv_bases = reinterpret_cast<const unsigned long long*>(recv_bases_dev);
  dp.rank = rank; dp.world = world;
  dp.recv_stride = ((size_t)N_PARAMS + 1 + 63) & ~(size_t)63;
  dp.cfg = *cfg;
  dp.state = state;
  dp.it_host = 0;
  dp.n_batches = n_batches;
  dp.loss_ring = loss_ring;
  dp.ring_len = ring_len;
  // the state lives on the device: every step is the same two launches, chained by programmatic dependent launch
  for (int64_t s = 0; s < n_steps; ++s) {
    float* partials; unsigned* err; int nap;
    OptScalars o{};
    const float* Xs = X_base; const float* ys = y_base;
    if (!state) {   // host-driven: batch offset and step scalars computed here, exactly as dlas_mlp_fit_steps does
      const int64_t bo = batch_order_host ? batch_order_host[s] : (cursor0 + s) % n_batches;   // model.fit's shuffled order
      DLAS_REQUIRE(bo >= 0 && bo < n_batches, "dlas_mlp_dp_step_at: batch index %lld out of range (%lld batches)", (long long)bo,
                   (long long)n_batches);
      const size_t b = (size_t)bo;
      Xs += b * batch_rows * D0;
      ys += b * batch_rows;
      dp.it_host = iteration0 + s;
      int rc = make_opt_scalars(cfg, (uint64_t)dp.it_host, o);
      if (rc != DLAS_OK) return rc;
    }
    int rc = launch_main(loss, params, Xs, ys, batch_rows, workspace, workspace_bytes, st, &partials, &err, &nap,
                         reinterpret_cast<const long long*>(state), /*rows_ready=*/s > 0, 0, s > 0 ? ws_image(workspace) : nullptr);
    if (rc != DLAS_OK) return rc;
    rc = launch_reduce<FIN_EXCHANGE>(st, partials, nap, err, nullptr, o, grad_scale, params, m, v, nullptr, nullptr, 0, dp,
                                     s + 1 < n_steps ? ws_image(workspace) : nullptr);
    if (rc != DLAS_OK) return rc;
  }
  return DLAS_OK;
}

}  // namespace dlas

extern "C" int dlas_prof_sm_clock(double* mhz_host, int reset) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  unsigned long long h[2] = {0ull, 0ull};
  DLAS_CUDA(cudaDeviceSynchronize());
  DLAS_CUDA(cudaMemcpyFromSymbol(h, g_clk_sum, sizeof(h)));
  if (mhz_host) *mhz_host = h[1] ? (double)h[0] / (double)h[1] * 1e3 : 0.0;
  if (reset) {
    const unsigned long long z[2] = {0ull, 0ull};
    DLAS_CUDA(cudaMemcpyToSymbol(g_clk_sum, z, sizeof(z)));
  }
  return DLAS_OK;
}

#ifdef DLAS_TC_TRACE
extern "C" int dlas_debug_fin_trace(unsigned long long* host64x2x4) {   // trace library only
  return cudaMemcpyFromSymbol(host64x2x4, dlas::g_fin_trace, sizeof(unsigned long long) * 64 * 2 * 4) == cudaSuccess ? 0 : -2;
}
#endif
