// LSTM regressor on the tensor cores (DLAS_MODE_TC; Readme.md:21, BASELINE.json configs[4]: 2 layers, hidden 128, T = 48):
// tcgen05.mma with fp16 operands and fp32 accumulators in TMEM for the gate GEMMs and the weight gradient.  Hidden width
// 128, d_in <= 16.  Same stash and parameter layout as the fp32 kernels of lstm.cu (lstm_common.cuh), so that every stage
// here can be exchanged for its fp32 counterpart (the tests do exactly that).
//
//   lstm_dw_tc_kernel   dK_l = [x_t, h_{t-1}]^T dz and db_l = dz^T 1 as one long-K GEMM over all (sequence, t) rows.
//                       CTA b owns gate chunk b % 4 (128 of the 512 gate columns = one gate: the accumulator's 128 lanes)
//                       and every (148/4)-th 128-row chunk; per chunk the loader warps convert the fp32 rows
//                       to fp16 tiles (double-buffered) and one thread issues  acc[gate col][k] += dz^T [x | h_prev]
//                       (M = 128, N = 144 / 256, K = 128 rows) plus the ones-vector MMA for db.  Per-CTA partials are
//                       added in slice order by lstm_dw_tc_reduce_kernel (deterministic).
#include "common.cuh"
#include "lstm_common.cuh"
#include "tc_common.cuh"

namespace dlas {
namespace {
using namespace tc;
using namespace lstm;

constexpr int HT = 128;                 // hidden width of the tensor-core path
constexpr int CH = 128;                 // rows per chunk (the K extent of one accumulation step)
constexpr int XW0 = 16;                 // layer 0: the d_in <= 16 input features are padded to one K = 16 block
__host__ __device__ inline int xw_of(int l) { return l == 0 ? XW0 : HT; }       // width of the input part of [x | h_prev]
__host__ __device__ inline int nk_of(int l) { return xw_of(l) + HT; }           // 144 or 256 accumulator columns

// ---- weight gradient -----------------------------------------------------------------------------------------------
namespace dw {
constexpr int N_LOAD = 256, ISSUER_WARP = 8, NT = 288;
constexpr uint32_t S_DZ = 0 /*2 x 32 KB*/, S_A = 65536 /*2 x 64 KB*/, S_ONES = 196608, S_BAR = S_ONES + 2048;
enum { B_FULL0 = 0, B_FULL1, B_FREE0, B_FREE1, B_ACC, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8, SMEM_BYTES = S_TMEM + 16;
constexpr uint32_t TM_ACC = 0, TM_DB = 256, TM_COLS = 512;
constexpr int PART = 256 * 128 + 128;   // floats per (layer, CTA) partial: [k 0..255][128 gate cols] then db[128]
}  // namespace dw

template <bool FAST>
__global__ void __launch_bounds__(dw::NT, 1)
lstm_dw_tc_kernel(const LstmDesc d, const float* __restrict__ X, const float* __restrict__ stash, const size_t n_seq,
                  float* __restrict__ partials, unsigned* __restrict__ err_flag, const unsigned call_id,
                  const uint16_t* __restrict__ h16, const uint16_t* __restrict__ dz16) {
  // FAST: both operands come as fp16 from the side arrays the recurrences wrote (h16, dz16: lstm_common.cuh), 16-byte
  // chunks copied into the swizzled tiles as they are.
  using namespace dw;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sDZ = sbase + S_DZ, sA = sbase + S_A, sONES = sbase + S_ONES;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int jc = blockIdx.x & 3, slice = blockIdx.x >> 2, n_slices = gridDim.x >> 2;
  const int T = d.T;
  const size_t n_rows = n_seq * (size_t)T;
  // chunk = 128 rows of the long K dimension.  fp32-compatible stash: consecutive (sequence, t) rows; fast layout: 128
  // consecutive sequences at one t (the order of the rows does not matter to the sum, only to the memory pattern)
  const size_t n_blk = (n_seq + CH - 1) / CH;
  const size_t n_chunks = FAST ? n_blk * (size_t)T : (n_rows + CH - 1) / CH;
  const int my_chunks = (n_chunks > (size_t)slice) ? (int)((n_chunks - slice + n_slices - 1) / n_slices) : 0;
  bool ok = (sbase & 1023u) == 0;
  if (tid == 0) {
    mbar_init(&bars[B_FULL0], N_LOAD / 32); mbar_init(&bars[B_FULL1], N_LOAD / 32);
    mbar_init(&bars[B_FREE0], 1); mbar_init(&bars[B_FREE1], 1); mbar_init(&bars[B_ACC], 1);
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  if (tid < 128) sts128(sONES + tid * 16, make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u));
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tm = *tmem_slot;
  uint32_t use = 0;       // running count of buffer uses over both layers: use u fills buffer u & 1, phase (u >> 1) & 1

  for (int l = 0; l < d.L; ++l) {
    const int dl = d.d_l[l], xw = xw_of(l), nk = nk_of(l);
    float* my = partials + ((size_t)l * gridDim.x + blockIdx.x) * PART;
    if (warp == ISSUER_WARP) {
      const bool leader = elect_one();
      const uint32_t idesc = make_idesc(128, nk, true, true);
      constexpr uint32_t ID_DB = make_idesc(128, 16, true, false);
      const uint64_t dONES = desc_kmajor(sONES, 16, 0, 0);
      uint32_t u = use;
      for (int c = 0; c < my_chunks; ++c, ++u) {
        const int b = u & 1;
        ok = mbar_wait(&bars[B_FULL0 + b], (u >> 1) & 1) && ok;
        fence_after();
        if (leader) {
          const uint64_t dDZ = desc_mnmajor(sDZ + b * 32768, CH, 0, 0), dAm = desc_mnmajor(sA + b * 65536, CH, 0, 0);
#pragma unroll
          for (int i = 0; i < CH / 16; ++i) {
            mma_f16(tm + TM_ACC, dDZ + koff_mnmajor(i), dAm + koff_mnmajor(i), idesc, !(c == 0 && i == 0));
            mma_f16(tm + TM_DB, dDZ + koff_mnmajor(i), dONES, ID_DB, !(c == 0 && i == 0));
          }
          mma_commit(&bars[B_FREE0 + b]);
          if (c == my_chunks - 1) mma_commit(&bars[B_ACC]);
        }
        __syncwarp();
      }
    } else {
      // ---- loaders: two threads per row; half 0 = the input part of [x | h_prev] and dz columns 0..63, half 1 = h_prev
      // and dz columns 64..127.  fp32 from the stash / X -> fp16 chunks of the swizzled tiles.
      const int r = tid & 127, half = tid >> 7;     // a warp = 32 consecutive rows of one half: coalesced in the fast layout
      uint32_t u = use;
      for (int c = 0; c < my_chunks; ++c, ++u) {
        const int b = u & 1;
        const size_t ci = (size_t)slice + (size_t)c * n_slices;
        size_t seq; int t; bool in;
        if (FAST) {
          t = (int)(ci % T);
          seq = (ci / T) * CH + r;
          in = seq < n_seq;
        } else {
          const size_t row_ = ci * CH + r;
          in = row_ < n_rows;
          seq = in ? row_ / T : 0;
          t = in ? (int)(row_ - seq * T) : 0;
        }
        if (!in) { seq = 0; t = 0; }
        const size_t row = seq * T + t;               // row of X
        const StashPos sp = stash_pos(d, seq);
        if (u >= 2) ok = mbar_wait(&bars[B_FREE0 + b], ((u >> 1) - 1) & 1) && ok;   // the MMAs of use u-2 have read this buffer
        const uint32_t tDZ = sDZ + b * 32768, tA = sA + b * 65536;
        if (FAST) {
          // every 16-byte chunk of the row -- 8 of dz (gate jc, this thread's unit half), 16 of [x | h_prev] -- is requested before
          // the first one is stored: one trip to L2 per chunk of rows instead of three
          uint4 vz[8], va[16];
#pragma unroll
          for (int q = 0; q < 8; ++q)
            vz[q] = in ? __ldcs(reinterpret_cast<const uint4*>(dz16 + idx16(sp, ((size_t)l * T + t) * 4 + jc, half * 64 + q * 8))) : make_uint4(0u, 0u, 0u, 0u);
          const bool hx = (half == 0 && l == 0);            // x_t comes from X (fp32), below
          const bool have = in && !hx && (half == 0 || t > 0);
          const int sl = (half == 0) ? l - 1 : l, stt = (half == 0) ? t : (t > 0 ? t - 1 : 0);
#pragma unroll
          for (int q = 0; q < 16; ++q)
            va[q] = have ? __ldg(reinterpret_cast<const uint4*>(h16 + idx16(sp, (size_t)(hx ? 0 : sl) * T + stt, q * 8))) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int q = 0; q < 8; ++q) sts128(tDZ + tile_chunk_off(CH, r, half * 64 + q * 8), vz[q]);
          if (!hx) {
            const int c0 = (half == 0) ? 0 : xw;
#pragma unroll
            for (int q = 0; q < 16; ++q) sts128(tA + tile_chunk_off(CH, r, c0 + q * 8), va[q]);
          }
        } else {      // dz: 64 floats of gate jc
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), e = a;
            if (in) {
              a = __ldcs(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, l, t, jc, half * 64 + q * 8)));
              e = __ldcs(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, l, t, jc, half * 64 + q * 8 + 4)));
            }
            sts128(tDZ + tile_chunk_off(CH, r, half * 64 + q * 8),
                   make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(e.x, e.y), pack_half2(e.z, e.w)));
          }
        }
        if (half == 0 && l == 0) {   // x_t: d_in <= 16 features, zero-padded to 16
          float xv[XW0];
#pragma unroll
          for (int k = 0; k < XW0; ++k) xv[k] = (in && k < dl) ? __ldg(X + row * (size_t)dl + k) : 0.f;
          sts128(tA + tile_chunk_off(CH, r, 0), make_uint4(pack_half2(xv[0], xv[1]), pack_half2(xv[2], xv[3]), pack_half2(xv[4], xv[5]), pack_half2(xv[6], xv[7])));
          sts128(tA + tile_chunk_off(CH, r, 8), make_uint4(pack_half2(xv[8], xv[9]), pack_half2(xv[10], xv[11]), pack_half2(xv[12], xv[13]), pack_half2(xv[14], xv[15])));
        } else {
          // half 0 (l > 0): h of the layer below at t -> columns 0..127; half 1: h_{t-1} of this layer -> columns xw..xw+127
          const bool have = in && (half == 0 || t > 0);
          const int sl = (half == 0) ? l - 1 : l, stt = (half == 0) ? t : (t > 0 ? t - 1 : 0);
          const int c0 = (half == 0) ? 0 : xw;
          if (FAST) {
            // (copied above, together with dz)
          } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) {
              float4 a = make_float4(0.f, 0.f, 0.f, 0.f), e = a;
              if (have) {
                a = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, sl, stt, 5, q * 8)));
                e = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, sl, stt, 5, q * 8 + 4)));
              }
              sts128(tA + tile_chunk_off(CH, r, c0 + q * 8),
                     make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(e.x, e.y), pack_half2(e.z, e.w)));
            }
          }
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[B_FULL0 + b]);
      }
      // ---- drain this layer's accumulators: lane = gate column jc*128 + lane, column n = k index of [x | h_prev] ----
      if (warp < 4) {
        const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
        const int j = warp * 32 + lane;
        if (my_chunks > 0) {
          ok = mbar_wait(&bars[B_ACC], l & 1) && ok;
          fence_after();
#pragma unroll 1
          for (int n0 = 0; n0 < nk; n0 += 16) {
            uint32_t v[16];
            tmem_ld16(tm + TM_ACC + lane_base + n0, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int n = n0 + i;
              const int k = (n < xw) ? (n < dl ? n : -1) : dl + (n - xw);       // accumulator column -> row of K_l
              if (k >= 0) my[(size_t)k * 128 + j] = __uint_as_float(v[i]);
            }
          }
          uint32_t w[16];
          tmem_ld16(tm + TM_DB + lane_base, w);
          tmem_ld_wait();
          my[256 * 128 + j] = __uint_as_float(w[0]);
          fence_before();
        } else {   // no chunk for this slice: its partial is zero
          for (int k = 0; k < dl + HT; ++k) my[(size_t)k * 128 + j] = 0.f;
          my[256 * 128 + j] = 0.f;
        }
      }
    }
    use += (uint32_t)my_chunks;
    // the next layer's first MMA overwrites the accumulators: everybody is done with this layer's
    fence_before();
    __syncthreads();
    fence_after();
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  fence_before();
  __syncthreads();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// grad_sum[K_l[k][jc*128 + j]] = sum over the slices of gate chunk jc, in slice order; same for b_l
__global__ void __launch_bounds__(256)
lstm_dw_tc_reduce_kernel(const LstmDesc d, const float* __restrict__ partials, const int n_ctas, const unsigned* __restrict__ err_flag,
                         const unsigned call_id, float* __restrict__ grad_sum) {
  using namespace dw;
  const int l = blockIdx.y;
  const int rows = d.d_l[l] + HT + 1;                   // K_l rows, then the bias row
  const int idx = blockIdx.x * 256 + threadIdx.x;       // (row, gate column)
  if (idx >= rows * 4 * HT) return;
  const int k = idx / (4 * HT), jg = idx - k * (4 * HT), jc = jg >> 7, j = jg & 127;
  const size_t off = (k < rows - 1) ? (size_t)k * 128 + j : (size_t)256 * 128 + j;
  float s = 0.f;
  for (int b = jc; b < n_ctas; b += 4) s += partials[((size_t)l * n_ctas + b) * PART + off];
  if (*err_flag == call_id) s = __int_as_float(0x7fc00000);
  grad_sum[(k < rows - 1) ? d.k_off[l] + (size_t)k * 4 * HT + jg : d.b_off[l] + jg] = s;
}

}  // namespace

// ---- entry points used by lstm.cu ----------------------------------------------------------------------------------
bool lstm_tc_supported(const lstm::LstmDesc& d) { return d.H == HT && d.d_in >= 1 && d.d_in <= XW0 && d.L >= 1; }

int lstm_tc_ctas() { return (sm_count() / 4) * 4; }
size_t lstm_tc_dw_partial_floats(const lstm::LstmDesc& d) { return (size_t)d.L * lstm_tc_ctas() * dw::PART; }
unsigned g_lstm_call_id = 0x5eed0000u;

// dK_l, db_l of every layer from X and the stash (h in comp 5, dz in comps 0..3) into grad_sum[0 .. hw_off)
int lstm_tc_dw(const lstm::LstmDesc& d, const float* X, const float* stash, size_t n_seq, float* grad_sum, float* tc_partials,
               unsigned* err_flag, const uint16_t* h16, const uint16_t* dz16, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(lstm_dw_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dw::SMEM_BYTES));
    DLAS_CUDA(cudaFuncSetAttribute(lstm_dw_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dw::SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  ++g_lstm_call_id;
  const int grid = lstm_tc_ctas();
  if (d.fast) lstm_dw_tc_kernel<true><<<grid, dw::NT, dw::SMEM_BYTES, st>>>(d, X, stash, n_seq, tc_partials, err_flag, g_lstm_call_id, h16, dz16);
  else lstm_dw_tc_kernel<false><<<grid, dw::NT, dw::SMEM_BYTES, st>>>(d, X, stash, n_seq, tc_partials, err_flag, g_lstm_call_id, h16, dz16);
  DLAS_LAUNCH_CHECK();
  const int max_rows = 2 * HT + 1;
  dim3 rg((unsigned)((max_rows * 4 * HT + 255) / 256), (unsigned)d.L);
  lstm_dw_tc_reduce_kernel<<<rg, 256, 0, st>>>(d, tc_partials, grid, err_flag, g_lstm_call_id, grad_sum);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

}  // namespace dlas
