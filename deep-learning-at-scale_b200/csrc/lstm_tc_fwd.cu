// LSTM forward recurrence on the tensor cores (DLAS_MODE_TC stage "forward"; Readme.md:21, BASELINE.json configs[4]).
// A persistent kernel: a CTA PAIR walks a tile of 128 sequences through all T steps of every layer with the state on chip.
//   * CTA `rank` of the pair owns hidden units [64 rank, 64 rank + 64): its 4 x 64 = 256 gate columns of K_l stay in shared
//     memory as one fp16 tile (128 KB for a 256-row layer), the cell state c of its units stays in registers (fp32), and
//     h_{t-1} of ALL 128 units sits in the A tile [128 sequences][x_t | h_{t-1}] (fp16) that tcgen05.mma reads.
//   * per step: z = [x_t | h_{t-1}] K_l[:, own columns]  (M = 128 sequences, N = 256, K = 144 / 256; fp32 in TMEM), then the
//     epilogue warps (thread = sequence row x 16 units) apply the gate non-linearities, update c, write i,f,g,o,c,h to the
//     stash (fp32, the layout of lstm.cu, so the fp32 BPTT can consume it) and store h_t as fp16 into BOTH CTAs' A tiles
//     (own half locally, the other through distributed shared memory) -- the only exchange of the recurrence.
//   * hand-over by mbarriers only: B_A (own rows of the next step are in place), B_PH (the peer's half of h has arrived:
//     st.async stores that complete their bytes on the receiver's barrier -- a releasing arrive would first wait for the
//     sender's stash stores), B_Z (tcgen05.commit), B_PZ (the peer's MMA of this step has completed, so its A tile may be
//     overwritten).
#include "common.cuh"
#include "lstm_common.cuh"
#include "tc_common.cuh"

namespace dlas {
namespace {
using namespace tc;
using namespace lstm;

constexpr int HT = 128, TS = 128, XW0 = 16, UH = 64;      // hidden width, sequences per tile, padded d_in, units per CTA
__host__ __device__ inline int xw_of(int l) { return l == 0 ? XW0 : HT; }
constexpr int N_EPI = 512, ISSUER_WARP = 16, NT = 544;   // 16 epilogue warps (4 per TMEM lane quarter, 16 units each): the epilogue
                                                           // is bound by memory and MUFU latency, so more warps in flight pay
constexpr int UT = 16;                                     // units per epilogue thread
constexpr uint32_t S_W = 0 /*[256 rows][256 gate cols] 128 KB*/, S_A = 131072 /*[128][256] 64 KB*/, S_F32 = 196608 /*bias[256]*/,
                   S_BAR = S_F32 + 1024;
enum { B_A = 0, B_PH, B_Z, B_PZ, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8, SMEM_BYTES = S_TMEM + 16, TM_COLS = 256;

__device__ __forceinline__ float sigm(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_(float x) { return 2.0f * sigm(2.0f * x) - 1.0f; }

template <bool FAST>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1)
lstm_fwd_tc_kernel(const LstmDesc d, const float* __restrict__ params, const float* __restrict__ X, const size_t n_seq,
                   float* __restrict__ stash, unsigned* __restrict__ err_flag, const unsigned call_id, unsigned* __restrict__ prog,
                   uint16_t* __restrict__ h16, uint16_t* __restrict__ g16) {
  // FAST (every stage of the step on the tensor cores): h goes out as fp16 (h16, lstm_common.cuh) for the layer above and the
  // weight-gradient GEMM; the fp32 h of the stash is written for the head only (top layer, last step).
  // Layer pipeline: the grid is n_lanes x L pairs; pair p works on layer p / n_lanes of the tiles lane, lane + n_lanes, ...
  // Layer l reads h of layer l-1 from the stash as soon as that pair has published the step (prog[tile * L + l - 1]), so the
  // layers of one tile run one step apart on different SM pairs instead of one after the other on the same pair.
  extern __shared__ __align__(1024) uint8_t smem[];
  float* bias = reinterpret_cast<float*>(smem + S_F32);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sW = sbase + S_W, sA = sbase + S_A;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int T = d.T;
  const size_t n_tiles = (n_seq + TS - 1) / TS;
  bool ok = (sbase & 1023u) == 0;
  if (tid == 0) {
    mbar_init(&bars[B_A], N_EPI / 32); mbar_init(&bars[B_PH], 1 + N_EPI / 32);   // B_PH: 1 local arrival (expect_tx) + the peer's 16 warps + its 16 KB of st.async bytes
    mbar_init(&bars[B_Z], 1); mbar_init(&bars[B_PZ], 1);
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  fence_before();
  __syncthreads();
  fence_after();
  cluster_sync_all();       // the peer is running and its barriers are initialised: remote arrivals / stores may start
  const uint32_t tm = *tmem_slot;
  const uint32_t peerA = map_to_cta(sA, rank ^ 1u);
  const uint32_t peer_PH = map_to_cta(smem_u32(&bars[B_PH]), rank ^ 1u), peer_PZ = map_to_cta(smem_u32(&bars[B_PZ]), rank ^ 1u);
  uint32_t q = 0;           // round = (tile, layer, step): every barrier completes exactly one phase per round

  // epilogue thread -> (row of the tile, 32 of this CTA's 64 units)
  const int q4 = warp & 3, hh = (warp >> 2) & 3;
  const int row = q4 * 32 + lane;
  const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;

  const int n_lanes = n_pairs / d.L, my_layer = pair / n_lanes;
  for (size_t tile = pair % n_lanes; tile < n_tiles; tile += n_lanes) {
    const size_t seq = tile * TS + row;
    const bool valid = (warp < ISSUER_WARP) && seq < n_seq;
    const StashPos sp = stash_pos(d, valid ? seq : 0);
    unsigned* const prog_mine = prog + tile * d.L + my_layer;          // published by this pair (layers below the top)
    const unsigned* const prog_below = prog_mine - (my_layer > 0 ? 1 : 0);                   // polled by layers above the first
    for (int l = my_layer; l == my_layer; ++l) {
      const int dl = d.d_l[l], xw = xw_of(l), nk = xw + HT;
      // ---- this layer's weights -> fp16 tile: row n of [x | h] (n < xw: input row n, zero beyond d_l; else recurrent row),
      // column g*64 + u = gate g of this CTA's unit u.  Everybody is past the previous layer here (barrier below).
      {
        const float* K = params + d.k_off[l];
        for (int idx = tid; idx < nk * 32; idx += NT) {       // 32 chunks of 8 columns per row
          const int n = idx >> 5, c = (idx & 31) * 8, g = c >> 6, u = c & 63;
          const int k = (n < xw) ? (n < dl ? n : -1) : dl + (n - xw);
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f), e = a;
          if (k >= 0) {
            const float* src = K + (size_t)k * 4 * HT + g * HT + UH * rank + u;
            a = __ldg(reinterpret_cast<const float4*>(src));
            e = __ldg(reinterpret_cast<const float4*>(src + 4));
          }
          sts128(sW + tile_chunk_off(256, n, c), make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(e.x, e.y), pack_half2(e.z, e.w)));
        }
        if (tid < 256) bias[tid] = __ldg(params + d.b_off[l] + (tid >> 6) * HT + UH * rank + (tid & 63));
      }
      // x part of step t for this thread's row(s): layer 0 -> 16 columns from X (threads of unit half 0), layers above ->
      // 64 of the 128 columns of h_t of the layer below (stash, written earlier by this pair; read through L2)
      auto x_store = [&](int t) {
        if (l == 0) {
          if (hh == 0) {
            float xv[XW0];
#pragma unroll
            for (int k = 0; k < XW0; ++k) xv[k] = (valid && k < dl) ? __ldg(X + (seq * T + t) * (size_t)dl + k) : 0.f;
            sts128(sA + tile_chunk_off(TS, row, 0), make_uint4(pack_half2(xv[0], xv[1]), pack_half2(xv[2], xv[3]), pack_half2(xv[4], xv[5]), pack_half2(xv[6], xv[7])));
            sts128(sA + tile_chunk_off(TS, row, 8), make_uint4(pack_half2(xv[8], xv[9]), pack_half2(xv[10], xv[11]), pack_half2(xv[12], xv[13]), pack_half2(xv[14], xv[15])));
          }
        } else if (FAST) {
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {
            uint4 o = make_uint4(0u, 0u, 0u, 0u);
            if (valid) o = __ldcg(reinterpret_cast<const uint4*>(h16 + idx16(sp, (size_t)(l - 1) * T + t, hh * 32 + c8 * 8)));
            sts128(sA + tile_chunk_off(TS, row, hh * 32 + c8 * 8), o);
          }
        } else {
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), e = a;
            if (valid) {
              a = __ldcg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, l - 1, t, 5, hh * 32 + c8 * 8)));
              e = __ldcg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, seq, l - 1, t, 5, hh * 32 + c8 * 8 + 4)));
            }
            sts128(sA + tile_chunk_off(TS, row, hh * 32 + c8 * 8), make_uint4(pack_half2(a.x, a.y), pack_half2(a.z, a.w), pack_half2(e.x, e.y), pack_half2(e.z, e.w)));
          }
        }
      };
      // ---- layer prologue: h_{-1} = 0 in the whole recurrent part of the A tile, x_0 in place -------------
      float c_state[UT];
      if (warp < ISSUER_WARP) {
#pragma unroll
        for (int j = 0; j < UT; ++j) c_state[j] = 0.f;
#pragma unroll
        for (int c8 = 0; c8 < 4; ++c8) sts128(sA + tile_chunk_off(TS, row, xw + hh * 32 + c8 * 8), make_uint4(0u, 0u, 0u, 0u));
        if (l > 0) {                       // step 0 of the layer below has been published
          if (lane == 0) ok = prog_wait(prog_below, PROG_WARPS) && ok;
          __syncwarp();
        }
        x_store(0);
      }
      fence_async_smem();
      __syncthreads();          // weights, bias, A tile of step 0 complete in this CTA
      if (warp < ISSUER_WARP) {
        // (the remote arrival carries no data: it keeps the pair in step, so that no barrier ever gets two phases ahead of a waiter)
        if (lane == 0) { mbar_arrive(&bars[B_A]); mbar_arrive_remote_relaxed(peer_PH); }
      }
      if (warp == ISSUER_WARP) {
        // =========================== MMA issuer ===========================
        const bool leader = elect_one();
        constexpr uint32_t ID_Z = make_idesc(128, 256, false, true);
        const uint64_t dA = desc_kmajor(sA, TS, 0, 0), dW = desc_mnmajor(sW, 256, 0, 0);
        for (int t = 0; t < T; ++t, ++q) {
          ok = mbar_wait(&bars[B_A], q & 1) && ok;
          // the peer's half of h_{t-1}: 128 rows x 64 units as asynchronous stores that complete their bytes on this barrier
          // (nothing at step 0: the phase still completes, so every barrier stays in step)
          if (leader) mbar_arrive_expect_tx(&bars[B_PH], t == 0 ? 0u : (uint32_t)(TS * UH * 2));
          ok = mbar_wait_cluster(&bars[B_PH], q & 1) && ok;
          fence_async_all();   // the peer's st.async bytes (observed through the barrier) before this thread's async-proxy reads
          fence_after();
          if (leader) {
            for (int i = 0; i < nk / 16; ++i)
              mma_f16(tm, dA + koff_kmajor(TS, i), dW + koff_mnmajor(i), ID_Z, i > 0);
            mma_commit(&bars[B_Z]);
          }
          __syncwarp();
          ok = mbar_wait(&bars[B_Z], q & 1) && ok;                 // my MMA has read my A tile: tell the peer it may overwrite its half
          if (lane == 0) mbar_arrive_remote(peer_PZ);
          __syncwarp();
        }
      } else {
        // =========================== epilogue warps ===========================
        for (int t = 0; t < T; ++t, ++q) {
          // the layer below's h two steps ahead: into L2 now, read at the end of the next round
          if (l > 0 && valid && t + 2 < T) {
            if (FAST) {
              prefetch_l2(h16 + idx16(sp, (size_t)(l - 1) * T + t + 2, hh * 32));
              prefetch_l2(h16 + idx16(sp, (size_t)(l - 1) * T + t + 2, hh * 32 + 16));
            } else {
              prefetch_l2(stash + sidx<FAST>(d, sp, seq, l - 1, t + 2, 5, hh * 32));
            }
          }
          if (l > 0 && t + 1 < T) {        // step t+1 of the layer below (read at the end of this round): polled here, under the MMA
            if (lane == 0) ok = prog_wait(prog_below, PROG_WARPS * (unsigned)(t + 2)) && ok;
            __syncwarp();
          }
          ok = mbar_wait(&bars[B_Z], q & 1) && ok;
          fence_after();
          // gates of this row's 16 units: columns g*64 + hh*16 + j of the accumulator
          uint32_t v[UT];
          float ig[UT], hout[UT];
          const int ub = UH * (int)rank + hh * UT;
          auto put = [&](int comp, const float (&a)[UT]) {
            if (FAST && comp < 4) {
              // all-tensor-core step: the gates go out as packed halves into the array the BPTT later overwrites with dz (same
              // thread owns the same (sequence, units) there); 11 bits on values in (-1, 1), the precision dz travels with anyway
              if (valid) {
#pragma unroll
                for (int c8 = 0; c8 < UT / 8; ++c8)
                  *reinterpret_cast<uint4*>(g16 + idx16(sp, ((size_t)l * T + t) * 4 + comp, ub + c8 * 8)) =
                      make_uint4(pack_half2(a[c8 * 8], a[c8 * 8 + 1]), pack_half2(a[c8 * 8 + 2], a[c8 * 8 + 3]),
                                 pack_half2(a[c8 * 8 + 4], a[c8 * 8 + 5]), pack_half2(a[c8 * 8 + 6], a[c8 * 8 + 7]));
              }
            } else if (valid) {
#pragma unroll
              for (int j4 = 0; j4 < UT / 4; ++j4)
                *reinterpret_cast<float4*>(stash + sidx<FAST>(d, sp, seq, l, t, comp, ub + j4 * 4)) = make_float4(a[j4 * 4], a[j4 * 4 + 1], a[j4 * 4 + 2], a[j4 * 4 + 3]);
            }
          };
          const float* bb = bias + hh * UT;
          tmem_ld16(tm + lane_base + 0 * 64 + hh * UT, v);           // i
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < UT; ++j) ig[j] = sigm(__uint_as_float(v[j]) + bb[j]);
          put(0, ig);
          tmem_ld16(tm + lane_base + 2 * 64 + hh * UT, v);           // g
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < UT; ++j) hout[j] = tanh_(__uint_as_float(v[j]) + bb[128 + j]);
          put(2, hout);
#pragma unroll
          for (int j = 0; j < UT; ++j) ig[j] *= hout[j];
          tmem_ld16(tm + lane_base + 1 * 64 + hh * UT, v);           // f
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < UT; ++j) hout[j] = sigm(__uint_as_float(v[j]) + bb[64 + j]);
          put(1, hout);
#pragma unroll
          for (int j = 0; j < UT; ++j) c_state[j] = fmaf(hout[j], c_state[j], ig[j]);
          put(4, c_state);
          tmem_ld16(tm + lane_base + 3 * 64 + hh * UT, v);           // o
          tmem_ld_wait();
          fence_before();
#pragma unroll
          for (int j = 0; j < UT; ++j) ig[j] = sigm(__uint_as_float(v[j]) + bb[192 + j]);
          put(3, ig);
#pragma unroll
          for (int j = 0; j < UT; ++j) hout[j] = ig[j] * tanh_(c_state[j]);
          uint4 hp[UT / 8];               // h_t as packed halves: the A tiles' recurrent part and (FAST) the h16 array
#pragma unroll
          for (int c8 = 0; c8 < UT / 8; ++c8)
            hp[c8] = make_uint4(pack_half2(hout[c8 * 8], hout[c8 * 8 + 1]), pack_half2(hout[c8 * 8 + 2], hout[c8 * 8 + 3]),
                                pack_half2(hout[c8 * 8 + 4], hout[c8 * 8 + 5]), pack_half2(hout[c8 * 8 + 6], hout[c8 * 8 + 7]));
          if (FAST) {
            if (valid) {
#pragma unroll
              for (int c8 = 0; c8 < UT / 8; ++c8) *reinterpret_cast<uint4*>(h16 + idx16(sp, (size_t)l * T + t, ub + c8 * 8)) = hp[c8];
            }
            if (l == d.L - 1 && t == T - 1) put(5, hout);     // the head reads fp32
          } else {
            put(5, hout);
          }
          // the peer's MMA of this step has completed (waited every round so that no phase is ever skipped)
          ok = mbar_wait_cluster(&bars[B_PZ], q & 1) && ok;
          if (t + 1 < T) {
            // h_t -> recurrent part of both A tiles (fp16), then the next step's input part
#pragma unroll
            for (int c8 = 0; c8 < UT / 8; ++c8) {
              const uint4 o = hp[c8];
              const uint32_t off = tile_chunk_off(TS, row, xw + UH * rank + hh * UT + c8 * 8);
              sts128(sA + off, o);
              st_async_v4(peerA + off, o, peer_PH);
            }
            x_store(t + 1);
            fence_async_smem();   // own sts128 -> the MMA's operand reads (the remote half goes by st.async)
            __syncwarp();
            if (lane == 0) { mbar_arrive(&bars[B_A]); mbar_arrive_remote_relaxed(peer_PH); }
          }
          if (l + 1 < d.L) {               // publish h_t to the pair of the layer above; the release waits for this warp's stash
            __syncwarp();                  // stores, after the recurrence's own hand-over so that it does not sit in its way
            if (lane == 0) prog_arrive(prog_mine);
          }
        }
      }
      // everybody of this CTA is past the layer's last round before the weights / A tile are rewritten
      fence_before();
      __syncthreads();
      fence_after();
    }
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  // ---- teardown: no remote store or arrival may target a CTA that has exited ----
  fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// head: Dense(H -> 1) on the last hidden state of the top layer (fp32, one warp per sequence)
__global__ void __launch_bounds__(256)
lstm_head_kernel(const LstmDesc d, const float* __restrict__ params, const float* __restrict__ stash, const size_t n_seq,
                 const unsigned* __restrict__ err_flag, const unsigned call_id, float* __restrict__ yhat) {
  const size_t s = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= n_seq) return;
  float a = 0.f;
  for (int u = lane; u < d.H; u += 32) a += stash[sidx(d, s, d.L - 1, d.T - 1, 5, u)] * __ldg(params + d.hw_off + u);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0) yhat[s] = (err_flag && *err_flag == call_id) ? __int_as_float(0x7fc00000) : a + __ldg(params + d.hb_off);
}

}  // namespace

extern unsigned g_lstm_call_id;   // lstm_tc.cu

// forward over all layers and steps: fills the stash (i, f, g, o, c, h per sequence, layer, step) and yhat
int lstm_tc_fwd(const lstm::LstmDesc& d, const float* params, const float* X, size_t n_seq, float* stash, float* yhat,
                unsigned* err_flag, unsigned* prog, uint16_t* h16, uint16_t* g16, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(lstm_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    DLAS_CUDA(cudaFuncSetAttribute(lstm_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  ++g_lstm_call_id;
  const size_t n_tiles = (n_seq + TS - 1) / TS;
  const int max_lanes = (sm_count() / 2) / d.L;          // a lane = the L pairs (one per layer) that share a tile
  if (max_lanes < 1) { set_error("lstm (tc): %d layers need more SM pairs than this GPU has", d.L); return DLAS_ERR_UNSUPPORTED_SHAPE; }
  const int pairs = (int)(n_tiles < (size_t)max_lanes ? n_tiles : (size_t)max_lanes) * d.L;
  DLAS_CUDA(cudaMemsetAsync(prog, 0, n_tiles * d.L * sizeof(unsigned), st));
  if (d.fast) lstm_fwd_tc_kernel<true><<<2 * pairs, NT, SMEM_BYTES, st>>>(d, params, X, n_seq, stash, err_flag, g_lstm_call_id, prog, h16, g16);
  else lstm_fwd_tc_kernel<false><<<2 * pairs, NT, SMEM_BYTES, st>>>(d, params, X, n_seq, stash, err_flag, g_lstm_call_id, prog, h16, g16);
  DLAS_LAUNCH_CHECK();
  lstm_head_kernel<<<(unsigned)((n_seq + 7) / 8), 256, 0, st>>>(d, params, stash, n_seq, err_flag, g_lstm_call_id, yhat);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

}  // namespace dlas
