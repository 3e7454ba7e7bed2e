// Tensor-core kernel for the small static ANN d0-64-64-1 (BASELINE.json configs[0]: 8-64-64-1; d0 <= 15), DLAS_MODE_TC:
// forward, loss and backward of cnn.py:127-128's batch body on tcgen05.mma, fp16 operands, fp32 accumulators in TMEM.
//
// The whole network fits one CTA (W2 is 8 KB as fp16), so there is no CTA pair and no exchange: one CTA walks 128-row
// tiles, and TWO CTAs share an SM (106 KB of shared memory, 256 TMEM columns, 160 threads each) -- while one waits
// for its MMAs the other runs its epilogue.  Per tile (thread = row, TMEM lane = row):
//     Z1 = [x | 1 | 0] [W1 ; b1 ; 0]      M128 N64 K16 (the bias rides in the GEMM)        1 MMA
//     H1 = relu(Z1) -> sH1 (fp16)
//     Z2 = H1 W2                           N64 K64                                          4 MMAs
//     yhat = relu(Z2 + b2) . w3 + b3       whole dot product inside the row's thread
//     dZ2 = dy w3 relu'(Z2) -> sDZ2
//     dH1 = dZ2 W2^T                       N64 K64                                          4 MMAs
//     dW2^T += dZ2^T H1, db2 += dZ2^T 1    M = layer-2 unit (64 of 128 lanes used), K = rows   8 + 8 MMAs
//     dZ1 = relu'(H1) o dH1 -> over H1 in place
//     dW1^T += dZ1^T [x | 1]               column d0 of the result is db1                   8 MMAs
// The tensor pipe is nearly idle by construction (K = 8 and N = 64 tiles): the kernel is bound by
// its epilogues; it is still ~15x the CUDA-core kernel because the 27,008 flop per row leave the FMA pipe.
// Shared memory (fp16 tiles, column-blocked SWIZZLE_128B of tc_common.cuh):
//   sX[2] [128][64] 2 x 16 KB | sH1 [128][128] 32 KB (columns 64.. zero) | sDZ2 [128][128] 32 KB (columns 64.. zero: the M
//   padding of dW2^T) | sW2 [64][64] 8 KB | sW1 [16][64] 2 KB | ones 2 KB | fp32 vectors
// TMEM columns: [0,64) Z1, then dH1 | [64,128) Z2 | [128,192) dW2^T | [192,208) db2 | [208,224) dW1^T
#include "common.cuh"
#include "optim_math.cuh"
#include "tc_common.cuh"
#include <math.h>

namespace dlas {
int make_opt_scalars(const dlas_opt_cfg* cfg, uint64_t iteration, OptScalars& o);   // optim.cu

namespace {
using namespace tc;

constexpr int TILE = 128, H = 64, K1 = 16;
constexpr int N_EPI = 128, ISSUER_WARP = 4, NT = 160;
constexpr uint32_t S_X = 0, S_H1 = 32768, S_DZ2 = 65536, S_W2 = 98304, S_W1 = 106496, S_ONES = 108544, S_F32 = 110592;
constexpr int F_B2 = 0, F_W3 = 64, F_Y = 128 /*[2][128]*/, F_RED = 384 /*[4][64]*/, F_END = F_RED + 256;
constexpr uint32_t S_BAR = S_F32 + F_END * 4;
enum { B_X0 = 0, B_X1, B_Z1, B_H1, B_Z2, B_DZ2, B_DH1, B_DZ1, B_DW1, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8;
constexpr uint32_t SMEM_BYTES = S_TMEM + 16;
constexpr uint32_t TM_Z1 = 0, TM_Z2 = 64, TM_DW2 = 128, TM_DB2 = 192, TM_DW1 = 208, TM_COLS = 256;
enum { K_TRAIN = 0, K_EVAL = 1, K_PREDICT = 2 };

struct Net {   // flat parameter offsets of d0-64-64-1
  int d0, w1, b1, w2, b2, w3, b3, n;
};
__host__ __device__ inline Net make_net(int d0) {
  Net t;
  t.d0 = d0; t.w1 = 0; t.b1 = d0 * H; t.w2 = t.b1 + H; t.b2 = t.w2 + H * H; t.w3 = t.b2 + H; t.b3 = t.w3 + H; t.n = t.b3 + 1;
  return t;
}
__host__ __device__ inline size_t pstride(int n) { return ((size_t)n + 1 + 3) & ~(size_t)3; }

__device__ __forceinline__ void warp_arrive(int lane, uint64_t* bar) {
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}

// v[0..32) per lane -> lane l returns the sum over the warp's 32 lanes of v[l]
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool hi = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      float send = hi ? v[i] : v[i + s];
      float keep = hi ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// KIND: K_TRAIN forward + loss + backward (per-CTA partials out), K_EVAL forward + loss sum, K_PREDICT forward (yhat out)
template <int KIND>
__global__ void __launch_bounds__(NT, 2)
mlp_tc64_kernel(const Net net, const int loss_kind, const float* __restrict__ params, const float* __restrict__ X_base,
                const float* __restrict__ y_base, const size_t n_rows, const size_t ldx, float* __restrict__ partials,
                unsigned* __restrict__ err_flag, const unsigned call_id, const long long* __restrict__ cursor,
                float* __restrict__ yhat_out) {
  constexpr bool TRAIN = (KIND == K_TRAIN);
  extern __shared__ __align__(1024) uint8_t smem[];
  float* f32 = reinterpret_cast<float*>(smem + S_F32);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sX = sbase + S_X, sH1 = sbase + S_H1, sDZ2 = sbase + S_DZ2, sW2 = sbase + S_W2, sW1 = sbase + S_W1,
                 sONES = sbase + S_ONES, sF = sbase + S_F32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int d0 = net.d0;
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  const int my_tiles = (n_tiles > (size_t)blockIdx.x) ? (int)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  bool ok = (sbase & 1023u) == 0;   // the swizzled tile layout needs a 1024-byte aligned base

  if (tid == 0) {
    mbar_init(&bars[B_X0], N_EPI / 32); mbar_init(&bars[B_X1], N_EPI / 32);
    mbar_init(&bars[B_H1], N_EPI / 32); mbar_init(&bars[B_DZ2], N_EPI / 32); mbar_init(&bars[B_DZ1], N_EPI / 32);
    mbar_init(&bars[B_Z1], 1); mbar_init(&bars[B_Z2], 1); mbar_init(&bars[B_DH1], 1); mbar_init(&bars[B_DW1], 1);
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  // everything above overlaps the tail of the previous kernel in the stream (programmatic dependent launch)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const size_t batch0 = cursor ? (size_t)__ldcg(cursor) * n_rows : 0;
  const float* __restrict__ X = X_base + batch0 * ldx;
  const float* __restrict__ y = (KIND == K_PREDICT) ? nullptr : y_base + batch0;

  // ---- weights -> fp16 tiles ---------------------------------------------------------------------------
  if (tid < 128) {
    {  // sW1 = [W1 ; b1 ; 0], 16 rows x 64 columns: one 16-byte chunk per thread
      const int r = tid >> 3, c = (tid & 7) * 8;
      float w[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) w[e] = (r < d0) ? __ldg(params + net.w1 + r * H + c + e) : (r == d0 ? __ldg(params + net.b1 + c + e) : 0.f);
      sts128(sW1 + tile_chunk_off(K1, r, c), make_uint4(pack_half2(w[0], w[1]), pack_half2(w[2], w[3]), pack_half2(w[4], w[5]), pack_half2(w[6], w[7])));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {   // sW2 [64][64]: 512 chunks
      const int idx = tid + u * 128, r = idx >> 3, c = (idx & 7) * 8;
      const float4 a = __ldg(reinterpret_cast<const float4*>(params + net.w2 + r * H + c));
      const float4 b = __ldg(reinterpret_cast<const float4*>(params + net.w2 + r * H + c + 4));
      sts128(sW2 + tile_chunk_off(H, r, c), make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(b.x, b.y), pack_half2(b.z, b.w)));
    }
    if (tid < H) {
      f32[F_B2 + tid] = __ldg(params + net.b2 + tid);
      f32[F_W3 + tid] = __ldg(params + net.w3 + tid);
    }
    // columns 64..127 of sH1 / sDZ2 pad the M = 128 views of dZ1^T and dZ2^T (their accumulator lanes 64..127 are never read)
    for (int c = 64; c < 128; c += 8) {
      sts128(sH1 + tile_chunk_off(TILE, tid, c), make_uint4(0u, 0u, 0u, 0u));
      sts128(sDZ2 + tile_chunk_off(TILE, tid, c), make_uint4(0u, 0u, 0u, 0u));
    }
    sts128(sONES + tid * 16, make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u));   // 2 KB of fp16 1.0
  }
  const float b3 = __ldg(params + net.b3);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tm = *tmem_slot;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp == ISSUER_WARP) {
    // =========================== MMA issuer (one elected lane) ===========================
    constexpr uint32_t ID_KM_MN_64 = make_idesc(128, 64, false, true);
    constexpr uint32_t ID_KM_KM_64 = make_idesc(128, 64, false, false);
    constexpr uint32_t ID_MN_MN_64 = make_idesc(128, 64, true, true);
    constexpr uint32_t ID_MN_MN_16 = make_idesc(128, 16, true, true);
    constexpr uint32_t ID_MN_KM_16 = make_idesc(128, 16, true, false);
    const bool leader = elect_one();
    const uint64_t dX_km = desc_kmajor(sX, TILE, 0, 0), dX_mn = desc_mnmajor(sX, TILE, 0, 0);
    const uint64_t dW1_mn = desc_mnmajor(sW1, K1, 0, 0);
    const uint64_t dW2_mn = desc_mnmajor(sW2, H, 0, 0), dW2_km = desc_kmajor(sW2, H, 0, 0);
    const uint64_t dH1_km = desc_kmajor(sH1, TILE, 0, 0), dH1_mn = desc_mnmajor(sH1, TILE, 0, 0);
    const uint64_t dDZ2_km = desc_kmajor(sDZ2, TILE, 0, 0), dDZ2_mn = desc_mnmajor(sDZ2, TILE, 0, 0);
    const uint64_t dONES = desc_kmajor(sONES, 16, 0, 0);
    for (int it = 0; it < my_tiles; ++it) {
      const int xb = it & 1, ph = it & 1;
      const bool first = (it == 0);
      const uint64_t xoff = (uint64_t)(xb * 1024);      // the second X buffer is 16 KB further: +1024 sixteen-byte units
      ok = mbar_wait(&bars[B_X0 + xb], (it >> 1) & 1) && ok;
      fence_after();
      if (leader) {   // Z1 + b1 = [x | 1 | 0] [W1 ; b1 ; 0]
        mma_f16(tm + TM_Z1, dX_km + xoff, dW1_mn, ID_KM_MN_64, 0);
        mma_commit(&bars[B_Z1]);
      }
      __syncwarp();
      ok = mbar_wait(&bars[B_H1], ph) && ok;
      fence_after();
      if (leader) {   // Z2 = H1 W2
#pragma unroll
        for (int i = 0; i < H / 16; ++i)
          mma_f16(tm + TM_Z2, dH1_km + koff_kmajor(TILE, i), dW2_mn + koff_mnmajor(i), ID_KM_MN_64, i > 0);
        mma_commit(&bars[B_Z2]);
      }
      __syncwarp();
      ok = mbar_wait(&bars[B_DZ2], ph) && ok;    // the epilogue has read Z2 (and, when training, stored dZ2)
      fence_after();
      if (!TRAIN) continue;
      if (leader) {
#pragma unroll
        for (int i = 0; i < H / 16; ++i)          // dH1 = dZ2 W2^T (into the Z1 columns: the epilogue has read Z1)
          mma_f16(tm + TM_Z1, dDZ2_km + koff_kmajor(TILE, i), dW2_km + koff_kmajor(H, i), ID_KM_KM_64, i > 0);
#pragma unroll
        for (int i = 0; i < TILE / 16; ++i)       // dW2^T[c][k] += dZ2^T H1 (lane = layer-2 unit c)
          mma_f16(tm + TM_DW2, dDZ2_mn + koff_mnmajor(i), dH1_mn + koff_mnmajor(i), ID_MN_MN_64, !(first && i == 0));
#pragma unroll
        for (int i = 0; i < TILE / 16; ++i)       // db2 += dZ2^T 1
          mma_f16(tm + TM_DB2, dDZ2_mn + koff_mnmajor(i), dONES, ID_MN_KM_16, !(first && i == 0));
        mma_commit(&bars[B_DH1]);                 // all three: sH1 may be overwritten by dZ1, sDZ2 is free
      }
      __syncwarp();
      ok = mbar_wait(&bars[B_DZ1], ph) && ok;
      fence_after();
      if (leader) {
#pragma unroll
        for (int i = 0; i < TILE / 16; ++i)       // dW1^T[k][j] += dZ1^T [x | 1]; column d0 = db1
          mma_f16(tm + TM_DW1, dH1_mn + koff_mnmajor(i), dX_mn + xoff + koff_mnmajor(i), ID_MN_MN_16, !(first && i == 0));
        mma_commit(&bars[B_DW1]);                 // sH1 and sX[xb] are free once this arrives
      }
      __syncwarp();
    }
  } else {
    // =========================== epilogue warps: thread = tile row = TMEM lane ===========================
    const int row = tid;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    const bool vec4 = (d0 % 4 == 0) && (ldx % 4 == 0) && ((((uintptr_t)X) & 15) == 0);
    float xr[K1];
    float yv = 0.f;
    auto x_issue = [&](size_t tile) {
      const size_t r = tile * TILE + row;
      const bool in = r < n_rows;
      const float* src = X + r * ldx;
      if (vec4) {
#pragma unroll
        for (int q = 0; q < K1 / 4; ++q) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (in && q * 4 < d0) v = __ldg(reinterpret_cast<const float4*>(src + q * 4));
          xr[q * 4 + 0] = v.x; xr[q * 4 + 1] = v.y; xr[q * 4 + 2] = v.z; xr[q * 4 + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int k = 0; k < K1; ++k) xr[k] = (in && k < d0) ? __ldg(src + k) : 0.f;
      }
      if (KIND != K_PREDICT) yv = in ? __ldcs(y + r) : 0.f;
    };
    auto x_store = [&](int j) {   // [x | 1 | 0] as fp16: the ones column multiplies the bias row of [W1 ; b1 ; 0]
      float h[K1];
#pragma unroll
      for (int k = 0; k < K1; ++k) h[k] = (k == d0) ? 1.f : xr[k];
      sts128(sX + j * 16384 + tile_chunk_off(TILE, row, 0), make_uint4(pack_half2(h[0], h[1]), pack_half2(h[2], h[3]), pack_half2(h[4], h[5]), pack_half2(h[6], h[7])));
      sts128(sX + j * 16384 + tile_chunk_off(TILE, row, 8), make_uint4(pack_half2(h[8], h[9]), pack_half2(h[10], h[11]), pack_half2(h[12], h[13]), pack_half2(h[14], h[15])));
      if (KIND != K_PREDICT) f32[F_Y + j * 128 + row] = yv;
      fence_async_smem();
      warp_arrive(lane, &bars[B_X0 + j]);
    };
    uint32_t h1p[H / 2];          // H1 of this row as packed halves: the ReLU mask of the backward pass
    float acc_w3[2][32];          // dw3 partials of this row, summed over rows once at the end
#pragma unroll
    for (int j = 0; j < H; ++j) acc_w3[j >> 5][j & 31] = 0.f;
    float acc_loss = 0.f, acc_b3 = 0.f;
    if (my_tiles > 0) {
      x_issue(blockIdx.x);
      x_store(0);
    }
    for (int it = 0; it < my_tiles; ++it) {
      const size_t tile = (size_t)blockIdx.x + (size_t)it * gridDim.x;
      const size_t row0 = tile * TILE;
      const bool row_valid = row0 + row < n_rows;
      const int ph = it & 1;
      if (it + 1 < my_tiles) x_issue(tile + gridDim.x);     // next tile's row: in flight under this tile's first phases
      // ---- H1 = relu(Z1 + b1) -> sH1 (the bias is already in the accumulator) ----
      {
        uint32_t v[H];
        ok = mbar_wait(&bars[B_Z1], ph) && ok;
        fence_after();
        tmem_ld32(tm + TM_Z1 + lane_base, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        tmem_ld32(tm + TM_Z1 + lane_base + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
        tmem_ld_wait();
        fence_before();
        if (TRAIN && it > 0) ok = mbar_wait(&bars[B_DW1], (it - 1) & 1) && ok;   // dW1 of the previous tile still reads sH1 / sX
#pragma unroll
        for (int g8 = 0; g8 < H / 8; ++g8) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            o[e] = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 2 * e]), __uint_as_float(v[g8 * 8 + 2 * e + 1])));
            h1p[g8 * 4 + e] = o[e];
          }
          sts128(sH1 + tile_chunk_off(TILE, row, g8 * 8), make_uint4(o[0], o[1], o[2], o[3]));
        }
        fence_async_smem();
        warp_arrive(lane, &bars[B_H1]);
      }
      if (it + 1 < my_tiles) x_store((it + 1) & 1);         // under the Z2 MMAs
      // ---- Z2: output dot product, loss, dZ2 -> sDZ2 ----
      uint32_t z[H];
      ok = mbar_wait(&bars[B_Z2], ph) && ok;
      fence_after();
      tmem_ld32(tm + TM_Z2 + lane_base, *reinterpret_cast<uint32_t(*)[32]>(&z[0]));
      tmem_ld32(tm + TM_Z2 + lane_base + 32, *reinterpret_cast<uint32_t(*)[32]>(&z[32]));
      tmem_ld_wait();
      fence_before();
      float p4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int blk = 0; blk < H / 4; ++blk) {   // z <- H2 = relu(Z2 + b2) in place; dot with w3 (four independent chains)
        const float4 cb = lds128f(sF + (F_B2 + blk * 4) * 4), cw = lds128f(sF + (F_W3 + blk * 4) * 4);
        const float b4[4] = {cb.x, cb.y, cb.z, cb.w}, w4[4] = {cw.x, cw.y, cw.z, cw.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float r = fmaxf(__uint_as_float(z[blk * 4 + e]) + b4[e], 0.f);
          p4[e] = fmaf(r, w4[e], p4[e]);
          z[blk * 4 + e] = __float_as_uint(r);
        }
      }
      const float yhat = ((p4[0] + p4[1]) + (p4[2] + p4[3])) + b3;
      if (KIND == K_PREDICT) {
        if (row_valid) yhat_out[row0 + row] = ok ? yhat : __int_as_float(0x7fc00000);
        warp_arrive(lane, &bars[B_DZ2]);        // Z2 has been read: the next tile may overwrite it
        continue;
      }
      float g = 0.f;
      if (row_valid) {
        const float d = yhat - f32[F_Y + (it & 1) * 128 + row];
        float lo;
        if (loss_kind == DLAS_LOSS_MSE) {
          lo = d * d;
          g = 2.0f * d;
        } else {  // cnn.py:29-32, CLIP_VALUE = 6
          const float a = fabsf(d);
          lo = fminf(a, 6.0f);
          g = (a <= 6.0f) ? (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) : 0.f;
        }
        acc_loss += lo;
        acc_b3 += g;
      }
      if (KIND == K_EVAL) {
        warp_arrive(lane, &bars[B_DZ2]);
        continue;
      }
#pragma unroll
      for (int g8 = 0; g8 < H / 8; ++g8) {
        const float4 c0 = lds128f(sF + (F_W3 + g8 * 8) * 4), c1 = lds128f(sF + (F_W3 + g8 * 8 + 4) * 4);
        const float w8[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = g8 * 8 + 2 * e;
          const float r0 = __uint_as_float(z[j]), r1 = __uint_as_float(z[j + 1]);        // H2
          acc_w3[j >> 5][j & 31] = fmaf(r0, g, acc_w3[j >> 5][j & 31]);                     // -> dw3
          acc_w3[(j + 1) >> 5][(j + 1) & 31] = fmaf(r1, g, acc_w3[(j + 1) >> 5][(j + 1) & 31]);
          o[e] = mul_h2(gt0_h2(pack_half2(r0, r1)), pack_half2(g * w8[2 * e], g * w8[2 * e + 1]));   // dZ2 = (H2 > 0) g w3
        }
        sts128(sDZ2 + tile_chunk_off(TILE, row, g8 * 8), make_uint4(o[0], o[1], o[2], o[3]));
      }
      fence_async_smem();
      warp_arrive(lane, &bars[B_DZ2]);
      // ---- dZ1 = relu'(H1) o dH1, in place over H1 ----
      {
        uint32_t v[H];
        ok = mbar_wait(&bars[B_DH1], ph) && ok;   // dH1 is in TMEM; dW2^T and db2 have finished reading sH1 / sDZ2
        fence_after();
        tmem_ld32(tm + TM_Z1 + lane_base, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        tmem_ld32(tm + TM_Z1 + lane_base + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
        tmem_ld_wait();
        fence_before();
#pragma unroll
        for (int g8 = 0; g8 < H / 8; ++g8) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e)
            o[e] = mul_h2(gt0_h2(h1p[g8 * 4 + e]), pack_half2(__uint_as_float(v[g8 * 8 + 2 * e]), __uint_as_float(v[g8 * 8 + 2 * e + 1])));
          sts128(sH1 + tile_chunk_off(TILE, row, g8 * 8), make_uint4(o[0], o[1], o[2], o[3]));
        }
        fence_async_smem();
        warp_arrive(lane, &bars[B_DZ1]);
      }
    }

    // ---- drain the accumulators into this CTA's partial vector -----------------------------------------
    float* my = partials + (size_t)blockIdx.x * pstride(net.n);
    if (TRAIN && my_tiles > 0) {
      ok = mbar_wait(&bars[B_DW1], (my_tiles - 1) & 1) && ok;   // the last tile's dW1: every accumulator is complete
      fence_after();
      if (warp < 2) {   // accumulator lanes 0..63 hold the 64 units; lanes 64..127 are padding
        // dW2^T: lane = layer-2 unit c, column = H1 unit k -> W2[k][c]: 32 lanes store 128 contiguous bytes
#pragma unroll 1
        for (int c0 = 0; c0 < H; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tm + TM_DW2 + lane_base + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) my[net.w2 + (c0 + i) * H + row] = __uint_as_float(v[i]);
        }
        uint32_t w[16];
        tmem_ld16(tm + TM_DB2 + lane_base, w);       // every column of the db2 accumulator holds the same sum
        tmem_ld_wait();
        my[net.b2 + row] = __uint_as_float(w[0]);
        tmem_ld16(tm + TM_DW1 + lane_base, w);       // dW1^T: lane = H1 unit, column = input feature; column d0 = db1
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < K1; ++j) {
          if (j < d0) my[net.w1 + j * H + row] = __uint_as_float(w[j]);
          else if (j == d0) my[net.b1 + row] = __uint_as_float(w[j]);
        }
      }
      // dw3: per-warp column sums over the warp's 32 rows, then the four warps in order
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const float cs = warp_colsum32(acc_w3[c], lane);
        f32[F_RED + warp * 64 + c * 32 + lane] = cs;
      }
      named_bar_sync(1, N_EPI);
      if (tid < H) my[net.w3 + tid] = (f32[F_RED + tid] + f32[F_RED + 64 + tid]) + (f32[F_RED + 128 + tid] + f32[F_RED + 192 + tid]);
      named_bar_sync(1, N_EPI);
    }
    if (KIND != K_PREDICT && my_tiles > 0) {   // loss (and db3): fixed-order sums over the 128 rows
      f32[F_RED + row] = acc_loss;
      f32[F_RED + 128 + row] = acc_b3;
      named_bar_sync(1, N_EPI);
      if (tid == 0 || (TRAIN && tid == 1)) {
        float s = 0.f;
        for (int i = 0; i < 128; ++i) s += f32[F_RED + tid * 128 + i];
        my[tid == 0 ? net.n : net.b3] = s;
      }
    }
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  // ---- teardown ----
  fence_before();
  __syncthreads();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// grad_sum[i] = sum of the active CTAs' partials in CTA order (every CTA holds a partial of every entry).
// FIN_GRAD: the sums go to `out`; FIN_UPDATE: the optimiser update is applied in the same pass and the batch-mean loss goes
// to *loss_out; FIN_LOSS: only the loss sum (validation pass).  Status words as in mlp_tc.cu (dlas_mlp_status).
enum { FIN_GRAD = 0, FIN_UPDATE = 1, FIN_LOSS = 2 };
constexpr int RG = 8;   // thread groups per block of 32 entries: the pass is latency-bound, more loads in flight
template <int MODE>
__global__ void __launch_bounds__(32 * RG)
tc64_reduce_kernel(const float* __restrict__ partials, const int n_active, const int n_params, unsigned* __restrict__ err_flag,
                   const unsigned call_id, float* __restrict__ out, const OptScalars opt, const float grad_scale,
                   float* __restrict__ params, float* __restrict__ m, float* __restrict__ v, float* __restrict__ loss_out) {
  __shared__ float red[RG - 1][32];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int i = (MODE == FIN_LOSS) ? n_params : blockIdx.x * 32 + (threadIdx.x & 31);
  const int grp = threadIdx.x >> 5;
  const bool live = (MODE == FIN_LOSS) ? (threadIdx.x & 31) == 0 : i <= n_params;
  const size_t ps = pstride(n_params);
  const int per = (n_active + RG - 1) / RG, k0 = grp * per, k1 = min(n_active, k0 + per);
  float s = 0.f;
  if (live) {
#pragma unroll 8
    for (int k = k0; k < k1; ++k) s += partials[(size_t)k * ps + i];
  }
  if (grp > 0) red[grp - 1][threadIdx.x & 31] = s;
  __syncthreads();
  if (grp > 0 || !live) return;
#pragma unroll
  for (int g2 = 0; g2 < RG - 1; ++g2) s += red[g2][threadIdx.x];
  if (*err_flag == call_id) s = __int_as_float(0x7fc00000);     // a wait timed out in this call's kernel: never return numbers
  if (!(fabsf(s) <= 3.0e38f)) err_flag[1] = call_id;            // fp16 operand range exceeded / diverged (DLAS_ERR_RANGE)
  if (blockIdx.x == 0 && threadIdx.x == 0) err_flag[2] = call_id;
  if (MODE == FIN_GRAD) {
    out[i] = s;
  } else if (MODE == FIN_LOSS) {
    *out = s;
  } else if (i == n_params) {
    if (loss_out) *loss_out = s * grad_scale;
  } else {
    float p = params[i], mi = m[i], vi = (opt.kind == 0) ? 0.f : v[i];
    opt_update(opt, s * grad_scale, p, mi, vi);
    params[i] = p;
    m[i] = mi;
    if (opt.kind != 0) v[i] = vi;
  }
}

int n_ctas64() { return 2 * sm_count(); }
unsigned g_call_id64 = 0x5eed0000u;
size_t ws_partials64(int n) { return (size_t)n_ctas64() * pstride(n) * sizeof(float); }

template <int KIND>
int launch_main64(const Net& net, int loss, const float* params, const float* X, const float* y, size_t n_rows, size_t ldx,
                  void* workspace, size_t workspace_bytes, cudaStream_t st, const long long* cursor, float* yhat, int* n_active) {
  if (KIND != K_PREDICT && (workspace == nullptr || workspace_bytes < ws_partials64(net.n) + 256)) {
    set_error("mlp (tc, %d-64-64-1): workspace of %zu bytes given, %zu needed", net.d0, workspace_bytes, ws_partials64(net.n) + 256);
    return DLAS_ERR_WORKSPACE;
  }
  float* partials = (KIND == K_PREDICT) ? nullptr : (float*)workspace;
  unsigned* err = (KIND == K_PREDICT) ? nullptr : (unsigned*)((char*)workspace + ws_partials64(net.n));
  static bool attr_set[64] = {};   // per instantiation and per device
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(mlp_tc64_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  ++g_call_id64;
  const int grid = n_ctas64();
  int pi = (KIND == K_TRAIN) ? prof_begin(DLAS_PROF_MLP_GRAD, st) : -1;
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
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, mlp_tc64_kernel<KIND>, net, loss, params, X, y, n_rows, ldx, partials, err, g_call_id64, cursor, yhat));
  if (KIND == K_TRAIN) prof_end(pi, st);
  DLAS_LAUNCH_CHECK();
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  if (n_active) *n_active = (int)(n_tiles < (size_t)grid ? n_tiles : (size_t)grid);
  return DLAS_OK;
}

template <int MODE>
int launch_reduce64(const Net& net, cudaStream_t st, void* workspace, int n_active, float* out, OptScalars o, float grad_scale,
                    float* params, float* m, float* v, float* loss_out) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(MODE == FIN_LOSS ? 1 : (net.n + 1 + 31) / 32);
  cfg.blockDim = dim3(32 * RG);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  unsigned* err = (unsigned*)((char*)workspace + ws_partials64(net.n));
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, tc64_reduce_kernel<MODE>, (const float*)workspace, n_active, net.n, err, g_call_id64, out, o,
                               grad_scale, params, m, v, loss_out));
  return DLAS_OK;
}

}  // namespace

// ---- entry points used by mlp_api.cu -----------------------------------------------------------------
bool tc64_supported(const int* dims, int n_layers) {
  return n_layers == 3 && dims[0] >= 1 && dims[0] < K1 && dims[1] == H && dims[2] == H && dims[3] == 1;
}

size_t tc64_workspace_bytes(const int* dims) { return ws_partials64(make_net(dims[0]).n) + 256; }

int tc64_status(const int* dims, const void* workspace, size_t workspace_bytes, cudaStream_t st) {
  const size_t off = ws_partials64(make_net(dims[0]).n);
  if (workspace == nullptr || workspace_bytes < off + 256) {
    set_error("dlas_mlp_status: workspace of %zu bytes given, %zu needed", workspace_bytes, off + 256);
    return DLAS_ERR_WORKSPACE;
  }
  unsigned w[4] = {0u, 0u, 0u, 0u};
  DLAS_CUDA(cudaMemcpyAsync(w, (const char*)workspace + off, sizeof(w), cudaMemcpyDeviceToHost, st));
  DLAS_CUDA(cudaStreamSynchronize(st));
  if (w[2] <= 0x5eed0000u || w[2] > g_call_id64) return DLAS_OK;   // no launch of this process has used the workspace
  if (w[0] == w[2]) {
    set_error("tensor-core MLP step: a wait inside the kernel timed out; the step's results were replaced by NaN");
    return DLAS_ERR_TIMEOUT;
  }
  if (w[1] == w[2]) {
    set_error("tensor-core MLP step: non-finite gradient or loss sum.  DLAS_MODE_TC keeps rows, activations and back-propagated "
              "terms as fp16 operands (|value| <= 65504, 11-bit significand); a value beyond that range became inf, or "
              "the training diverged.  Rescale the feature / target columns or use DLAS_MODE_FP32");
    return DLAS_ERR_RANGE;
  }
  return DLAS_OK;
}

int tc64_grad(const int* dims, int loss, const float* params, const float* X, const float* y, size_t n_rows, float* grad_sum,
              void* workspace, size_t workspace_bytes, cudaStream_t st, const long long* cursor, size_t parity_stride, size_t ldx) {
  const Net net = make_net(dims[0]);
  if (parity_stride != 0) {
    set_error("DLAS_MODE_TC, %d-64-64-1: the double-buffered bucket of dlas_dp_allreduce_update is an fp32-mode path", net.d0);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)net.n + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  int na = 0;
  int rc = launch_main64<K_TRAIN>(net, loss, params, X, y, n_rows, ldx ? ldx : (size_t)net.d0, workspace, workspace_bytes, st, cursor, nullptr, &na);
  if (rc != DLAS_OK) return rc;
  return launch_reduce64<FIN_GRAD>(net, st, workspace, na, grad_sum, OptScalars{}, 0.f, nullptr, nullptr, nullptr, nullptr);
}

// one whole single-GPU step: gradient kernel, then the reduction fused with the optimiser update
int tc64_fit_step(const int* dims, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* X,
                  const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace, size_t workspace_bytes,
                  cudaStream_t st, size_t ldx) {
  const Net net = make_net(dims[0]);
  OptScalars o;
  int rc = make_opt_scalars(cfg, iteration, o);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(o.kind == DLAS_OPT_SGD || v != nullptr, "Adam needs the second-moment buffer v");
  int na = 0;
  rc = launch_main64<K_TRAIN>(net, loss, params, X, y, n_rows, ldx ? ldx : (size_t)net.d0, workspace, workspace_bytes, st, nullptr, nullptr, &na);
  if (rc != DLAS_OK) return rc;
  return launch_reduce64<FIN_UPDATE>(net, st, workspace, na, nullptr, o, 1.0f / (float)n_rows, params, m, v, loss_out);
}

int tc64_predict(const int* dims, const float* params, const float* X, size_t n_rows, float* yhat, cudaStream_t st, size_t ldx) {
  const Net net = make_net(dims[0]);
  if (n_rows == 0) return DLAS_OK;
  return launch_main64<K_PREDICT>(net, DLAS_LOSS_MSE, params, X, nullptr, n_rows, ldx ? ldx : (size_t)net.d0, nullptr, 0, st, nullptr, yhat, nullptr);
}

int tc64_eval_loss(const int* dims, int loss, const float* params, const float* X, const float* y, size_t n_rows, float* loss_sum,
                   void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx) {
  const Net net = make_net(dims[0]);
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
    return DLAS_OK;
  }
  int na = 0;
  int rc = launch_main64<K_EVAL>(net, loss, params, X, y, n_rows, ldx ? ldx : (size_t)net.d0, workspace, workspace_bytes, st, nullptr, nullptr, &na);
  if (rc != DLAS_OK) return rc;
  return launch_reduce64<FIN_LOSS>(net, st, workspace, na, loss_sum, OptScalars{}, 0.f, nullptr, nullptr, nullptr, nullptr);
}

}  // namespace dlas
