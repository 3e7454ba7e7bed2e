// LSTM backward recurrence (BPTT) on the tensor cores (DLAS_MODE_TC stage "backward"; Readme.md:21, BASELINE.json
// configs[4]).  The mirror of lstm_tc_fwd.cu: a CTA PAIR walks a tile of 128 sequences backwards through time, layer by
// layer from the top; CTA `rank` owns hidden units [64 rank, 64 rank + 64) and keeps their dc in registers.
//   per step t:  dz (4 gates of the own units) from the stash and dh_t, dc_{t+1}   (epilogue warps, fp32; dz overwrites the
//                gates in the stash for the weight-gradient GEMM and goes to the A tile sDZ as fp16)
//                d[x_t | h_{t-1}] (partial over the own 256 gate columns) = dz K_l[:, own columns]^T
//                                                             (tcgen05: M = 128 sequences, N = 144 / 256, K = 256)
//                the partial's columns of the PEER's units go to the peer through distributed shared memory (fp32), the
//                own units' columns are read back from TMEM at step t-1; the dx columns (layers above 0) go to one of two
//                partial dseq buffers in global memory, which the layer below adds when it reads dh_t.
// The weight tile is the forward kernel's ([x | h] rows x own gate columns): tcgen05 reads it K-major here, MN-major there.
// mbarriers: B_A (sDZ of this step in place), B_Z (tcgen05.commit), B_RFULL (the peer's partial for my units has arrived:
// st.async stores that complete their bytes on the receiver's barrier, so no releasing arrive has to wait for the sender's
// global stores; plus one relaxed arrival per peer warp and round, which keeps the two CTAs within one round of each other),
// B_RFREE (the peer has read what I sent last: its receive buffer may be overwritten).  Every barrier completes exactly one
// phase per round (arrivals without data where a round has nothing to send), so no waiter ever skips a phase and no barrier
// gets two phases ahead of one.
#include "common.cuh"
#include "lstm_common.cuh"
#include "tc_common.cuh"

namespace dlas {
namespace {
using namespace tc;
using namespace lstm;

constexpr int HT = 128, TS = 128, XW0 = 16, UH = 64;
__host__ __device__ inline int xw_of(int l) { return l == 0 ? XW0 : HT; }
constexpr int N_EPI = 512, ISSUER_WARP = 16, NT = 544;   // 16 epilogue warps: the step is bound by the latency of the stash reads
constexpr int UT = 16;                                     // units per epilogue thread
constexpr uint32_t S_W = 0 /*[256 rows][256 gate cols] fp16, 128 KB*/, S_DZ = 131072 /*[128][256] fp16, 64 KB*/,
                   S_RECV = 196608 /*[128][64] fp32, 32 KB*/, S_BAR = S_RECV + 32768;
enum { B_A = 0, B_Z, B_RFULL, B_RFREE, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8, SMEM_BYTES = S_TMEM + 16, TM_COLS = 256;

__device__ __forceinline__ float sigm(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_(float x) { return 2.0f * sigm(2.0f * x) - 1.0f; }

// receive buffer [128 rows][16 chunks of 4 floats], chunks XOR-swizzled with the row so that a quarter-warp's 16-byte accesses
// (8 consecutive rows, same chunk) fall into different banks
__device__ __forceinline__ uint32_t recv_off(int row, int c16) { return (uint32_t)(row * 256 + ((c16 ^ (row & 15)) << 4)); }

// dy (d loss / d yhat, unscaled) of every sequence overwrites yhat; head gradient, head bias gradient and the loss sum go
// to the per-block partial vectors in the layout lstm_bwd_kernel uses (block b: sequences b, b + grid, ... in order)
__global__ void __launch_bounds__(256)
lstm_head_grad_kernel(const LstmDesc d, const int loss, const float* __restrict__ y, float* __restrict__ yhat_dy,
                      const float* __restrict__ stash, const size_t n_seq, float* __restrict__ partials) {
  float* my = partials + (size_t)blockIdx.x * pstride(d.n_params);
  const int u = threadIdx.x;
  __shared__ float g_sh[256], lo_sh[256];
  float acc = 0.f;
  // the block's sequences in order, 256 at a time: thread j first turns yhat of sequence j into dy (and keeps g, the loss
  // term), then thread u < H adds h[u] * g over them -- loads that do not depend on each other
  for (size_t base = blockIdx.x; base < n_seq; base += (size_t)gridDim.x * 256) {
    const size_t s = base + (size_t)u * gridDim.x;
    float g = 0.f, lo = 0.f;
    if (s < n_seq) {
      const float dd = yhat_dy[s] - y[s];
      if (loss == DLAS_LOSS_MSE) { lo = dd * dd; g = 2.0f * dd; }
      else { const float a = fabsf(dd); lo = fminf(a, 6.0f); g = (a <= 6.0f) ? (dd > 0.0f ? 1.0f : (dd < 0.0f ? -1.0f : 0.0f)) : 0.0f; }
      yhat_dy[s] = g;
    }
    g_sh[u] = g; lo_sh[u] = lo;
    __syncthreads();
    const size_t left = (n_seq - base + gridDim.x - 1) / gridDim.x;
    const int cnt = (int)(left < 256 ? left : 256);
    if (u < d.H) {
#pragma unroll 4
      for (int j = 0; j < cnt; ++j) acc += stash[sidx(d, base + (size_t)j * gridDim.x, d.L - 1, d.T - 1, 5, u)] * g_sh[j];
    } else if (u == d.H) {
      for (int j = 0; j < cnt; ++j) acc += g_sh[j];
    } else if (u == d.H + 1) {
      for (int j = 0; j < cnt; ++j) acc += lo_sh[j];
    }
    __syncthreads();
  }
  if (u < d.H) my[d.hw_off + u] = acc;
  else if (u == d.H) my[d.hb_off] = acc;
  else if (u == d.H + 1) my[d.n_params] = acc;
}

template <bool FAST>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1)
lstm_bwd_tc_kernel(const LstmDesc d, const float* __restrict__ params, float* __restrict__ stash, const float* __restrict__ dy,
                   float* __restrict__ dseqP, const size_t n_seq, unsigned* __restrict__ err_flag, const unsigned call_id,
                   unsigned* __restrict__ prog, uint16_t* __restrict__ dz16) {
  // FAST: dz leaves as fp16 (dz16, lstm_common.cuh) -- the operand the weight-gradient GEMM takes -- instead of overwriting
  // the fp32 gates of the stash.
  // Layer pipeline (as in lstm_tc_fwd.cu, top layer first): pair p works on layer L-1 - p / n_lanes of the tiles lane,
  // lane + n_lanes, ...; a layer reads the dx partials of the layer above as soon as that pair has published the step.
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sW = sbase + S_W, sDZ = sbase + S_DZ, sRecv = sbase + S_RECV;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int T = d.T;
  const size_t n_tiles = (n_seq + TS - 1) / TS;
  const size_t pstr = (FAST ? d.n_pad : n_seq) * (size_t)T * HT;   // floats per partial dseq buffer
  // dseqP: [set][CTA rank][sequence][t][unit].  A layer writes the set of its own parity and reads the other one, so a middle
  // layer (3 layers and more) never overwrites a partial its peer may still be reading; up to 2 layers use one set.
  const int n_sets = d.L >= 3 ? 2 : 1;
  bool ok = (sbase & 1023u) == 0;
  if (tid == 0) {
    mbar_init(&bars[B_A], N_EPI / 32); mbar_init(&bars[B_Z], 1);
    mbar_init(&bars[B_RFULL], 1 + N_EPI / 32); mbar_init(&bars[B_RFREE], N_EPI / 32);   // B_RFULL: 1 local arrival (expect_tx) + the peer's 16 warps + 32 KB of st.async bytes
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  fence_before();
  __syncthreads();
  fence_after();
  cluster_sync_all();
  const uint32_t tm = *tmem_slot;
  const uint32_t peerRecv = map_to_cta(sRecv, rank ^ 1u);
  const uint32_t peer_RFULL = map_to_cta(smem_u32(&bars[B_RFULL]), rank ^ 1u), peer_RFREE = map_to_cta(smem_u32(&bars[B_RFREE]), rank ^ 1u);
  uint32_t q = 0;

  const int q4 = warp & 3, hh = (warp >> 2) & 3;
  const int row = q4 * 32 + lane;
  const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
  const int u0 = UH * (int)rank + hh * UT;                // first of this thread's 16 units

  const int n_lanes = n_pairs / d.L, my_layer = d.L - 1 - pair / n_lanes;
  for (size_t tile = pair % n_lanes; tile < n_tiles; tile += n_lanes) {
    const size_t seq = tile * TS + row;
    const bool valid = (warp < ISSUER_WARP) && seq < n_seq;
    const size_t sq = valid ? seq : 0;
    const StashPos sp = stash_pos(d, sq);
    unsigned* const prog_mine = prog + tile * d.L + my_layer;                           // dx steps this pair has published
    const unsigned* const prog_above = prog_mine + (my_layer + 1 < d.L ? 1 : 0);        // ... and the layer above
    for (int l = my_layer; l == my_layer; --l) {
      const int dl = d.d_l[l], xw = xw_of(l), nk = xw + HT;
      float* dsW = dseqP + (size_t)(n_sets == 2 ? (l & 1) : 0) * 2 * pstr;              // this layer's dx partials
      const float* dsR = dseqP + (size_t)(n_sets == 2 ? ((l + 1) & 1) : 0) * 2 * pstr;  // the layer above's
      {   // weights of this layer -> the forward kernel's tile: row n of [x | h], column g*64 + u
        const float* K = params + d.k_off[l];
        for (int idx = tid; idx < nk * 32; idx += NT) {
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
      }
      fence_async_smem();
      __syncthreads();
      if (warp == ISSUER_WARP) {
        // =========================== MMA issuer ===========================
        const bool leader = elect_one();
        const uint32_t idesc = make_idesc(128, nk, false, false);
        const uint64_t dA = desc_kmajor(sDZ, TS, 0, 0), dB = desc_kmajor(sW, 256, 0, 0);
        for (int t = T - 1; t >= 0; --t, ++q) {
          ok = mbar_wait(&bars[B_A], q & 1) && ok;
          fence_after();
          if (leader) {
#pragma unroll
            for (int i = 0; i < 16; ++i)     // K = the 256 own gate columns
              mma_f16(tm, dA + koff_kmajor(TS, i), dB + koff_kmajor(256, i), idesc, i > 0);
            mma_commit(&bars[B_Z]);
          }
          __syncwarp();
        }
      } else {
        // =========================== epilogue warps ===========================
        float dc_next[UT], dh_rec[UT];
#pragma unroll
        for (int j = 0; j < UT; ++j) dc_next[j] = 0.f;
        const float* hw = params + d.hw_off + u0;
        for (int t = T - 1; t >= 0; --t, ++q) {
          const bool first = (t == T - 1);
          if (valid && t > 0) {   // step t-1's stash lines (and the layer above's dx): into L2 a round ahead
            if (FAST) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                prefetch_l2(dz16 + idx16(sp, ((size_t)l * T + t - 1) * 4 + g, u0));
                prefetch_l2(dz16 + idx16(sp, ((size_t)l * T + t - 1) * 4 + g, u0 + 8));
              }
              prefetch_l2(stash + sidx<FAST>(d, sp, sq, l, t - 1, 4, u0));
            } else {
#pragma unroll
              for (int c = 0; c < 5; ++c) prefetch_l2(stash + sidx<FAST>(d, sp, sq, l, t - 1, c, u0));
            }
            if (t > 1) prefetch_l2(stash + sidx<FAST>(d, sp, sq, l, t - 2, 4, u0));
            if (l < d.L - 1) { prefetch_l2(dsR + didx<FAST>(d, sp, sq, t - 1, u0)); prefetch_l2(dsR + pstr + didx<FAST>(d, sp, sq, t - 1, u0)); }
          }
          if (l < d.L - 1) {     // dx of step t from the layer above (read below, after the exchange): polled here, under the MMA
            if (lane == 0) ok = prog_wait(prog_above, PROG_WARPS * (unsigned)(T - t)) && ok;
            __syncwarp();
          }
          // ---- what the previous round's MMA (step t+1) produced: dx partial, my units' dh partial, the peer's units'
          if (!first) {
            ok = mbar_wait(&bars[B_Z], (q - 1) & 1) && ok;
            fence_after();
            uint32_t v[16];
            tmem_ld16(tm + lane_base + xw + UH * (rank ^ 1u) + hh * UT, v);   // the peer's units: send
            tmem_ld_wait();
            if (q >= 1) ok = mbar_wait_cluster(&bars[B_RFREE], (q - 1) & 1) && ok;   // the peer has read what I sent last round
#pragma unroll
            for (int j4 = 0; j4 < UT / 4; ++j4)
              st_async_v4(peerRecv + recv_off(row, hh * 4 + j4), make_uint4(v[j4 * 4], v[j4 * 4 + 1], v[j4 * 4 + 2], v[j4 * 4 + 3]), peer_RFULL);
            tmem_ld16(tm + lane_base + xw + u0, v);                    // my units: keep
            tmem_ld_wait();
            fence_before();
#pragma unroll
            for (int j = 0; j < UT; ++j) dh_rec[j] = __uint_as_float(v[j]);
          } else {
            if (q >= 1) ok = mbar_wait_cluster(&bars[B_RFREE], (q - 1) & 1) && ok;   // nothing to send: phases stay in step
#pragma unroll
            for (int j = 0; j < UT; ++j) dh_rec[j] = 0.f;
          }
          // the peer's partial for my units: 32 KB of asynchronous stores that complete their bytes on this barrier (none in a
          // layer's first round).  Every warp also arrives on the peer's barrier (relaxed: the bytes carry their own completion):
          // that keeps the pair in step, so that no barrier ever gets two phases ahead of a waiter
          __syncwarp();
          if (lane == 0) mbar_arrive_remote_relaxed(peer_RFULL);
          if (tid == 0) mbar_arrive_expect_tx(&bars[B_RFULL], first ? 0u : (uint32_t)(TS * UH * 4));
          ok = mbar_wait_cluster(&bars[B_RFULL], q & 1) && ok;
          if (!first) {
#pragma unroll
            for (int j4 = 0; j4 < UT / 4; ++j4) {
              const float4 r = lds128f(sRecv + recv_off(row, hh * 4 + j4));
              dh_rec[j4 * 4] += r.x; dh_rec[j4 * 4 + 1] += r.y; dh_rec[j4 * 4 + 2] += r.z; dh_rec[j4 * 4 + 3] += r.w;
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_remote_relaxed(peer_RFREE);
          // ---- the dx columns of step t+1 (layers above 0) -> this CTA's partial buffer.  Here, not before the exchange: no
          // fence of the recurrence ever waits for these stores (warp-uniform: tcgen05.ld is warp-collective; only the stores
          // depend on the row being valid)
          if (!first && l > 0) {
            uint32_t w[32];
            float* dst = dsW + (size_t)rank * pstr;
            tmem_ld32(tm + lane_base + hh * 32, w);
            tmem_ld_wait();
            fence_before();
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4)
              if (valid) *reinterpret_cast<float4*>(dst + didx<FAST>(d, sp, sq, t + 1, hh * 32 + j4 * 4)) = make_float4(__uint_as_float(w[j4 * 4]), __uint_as_float(w[j4 * 4 + 1]),
                                                                             __uint_as_float(w[j4 * 4 + 2]), __uint_as_float(w[j4 * 4 + 3]));
          }
          // ---- gate gradients of step t for this row's 16 units, 8 units at a time ----

          const float dyr = (l == d.L - 1 && first && valid) ? __ldg(dy + sq) : 0.f;

#pragma unroll
          for (int b8 = 0; b8 < UT / 8; ++b8) {
            float gi[8], gf[8], gg[8], go[8], cc[8], cpv[8], dhv[8];
#pragma unroll
            for (int h4 = 0; h4 < 2; ++h4) {
              const int o = b8 * 8 + h4 * 4;
              float4 a = make_float4(0.f, 0.f, 0.f, 0.f), f4 = a, g4 = a, o4 = a, c4 = a, p4 = a, d4 = a;
              if (valid) {
                if (FAST) {   // gates as packed halves (written by the forward kernel into the dz array): 4 of them per 8-byte word pair
                  const size_t sl4 = ((size_t)l * T + t) * 4;
                  const uint2 pi = __ldcg(reinterpret_cast<const uint2*>(dz16 + idx16(sp, sl4 + 0, u0 + b8 * 8) + h4 * 4));
                  const uint2 pf = __ldcg(reinterpret_cast<const uint2*>(dz16 + idx16(sp, sl4 + 1, u0 + b8 * 8) + h4 * 4));
                  const uint2 pg = __ldcg(reinterpret_cast<const uint2*>(dz16 + idx16(sp, sl4 + 2, u0 + b8 * 8) + h4 * 4));
                  const uint2 po = __ldcg(reinterpret_cast<const uint2*>(dz16 + idx16(sp, sl4 + 3, u0 + b8 * 8) + h4 * 4));
                  auto un = [](uint2 p) {
                    const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&p.x)), hi = __half22float2(*reinterpret_cast<const __half2*>(&p.y));
                    return make_float4(lo.x, lo.y, hi.x, hi.y);
                  };
                  a = un(pi); f4 = un(pf); g4 = un(pg); o4 = un(po);
                } else {
                  a = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 0, u0 + o)));
                  f4 = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 1, u0 + o)));
                  g4 = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 2, u0 + o)));
                  o4 = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 3, u0 + o)));
                }
                c4 = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 4, u0 + o)));
                if (t > 0) p4 = __ldg(reinterpret_cast<const float4*>(stash + sidx<FAST>(d, sp, sq, l, t - 1, 4, u0 + o)));
                if (l < d.L - 1) {
                  const float4 d0 = __ldcg(reinterpret_cast<const float4*>(dsR + didx<FAST>(d, sp, sq, t, u0 + o)));
                  const float4 d1 = __ldcg(reinterpret_cast<const float4*>(dsR + pstr + didx<FAST>(d, sp, sq, t, u0 + o)));
                  d4 = make_float4(d0.x + d1.x, d0.y + d1.y, d0.z + d1.z, d0.w + d1.w);
                } else if (first) {
                  const float4 w4 = __ldg(reinterpret_cast<const float4*>(hw + o));
                  d4 = make_float4(dyr * w4.x, dyr * w4.y, dyr * w4.z, dyr * w4.w);
                }
              }
              const int e = h4 * 4;
              gi[e] = a.x; gi[e + 1] = a.y; gi[e + 2] = a.z; gi[e + 3] = a.w;
              gf[e] = f4.x; gf[e + 1] = f4.y; gf[e + 2] = f4.z; gf[e + 3] = f4.w;
              gg[e] = g4.x; gg[e + 1] = g4.y; gg[e + 2] = g4.z; gg[e + 3] = g4.w;
              go[e] = o4.x; go[e + 1] = o4.y; go[e + 2] = o4.z; go[e + 3] = o4.w;
              cc[e] = c4.x; cc[e + 1] = c4.y; cc[e + 2] = c4.z; cc[e + 3] = c4.w;
              cpv[e] = p4.x; cpv[e + 1] = p4.y; cpv[e + 2] = p4.z; cpv[e + 3] = p4.w;
              dhv[e] = d4.x; dhv[e + 1] = d4.y; dhv[e + 2] = d4.z; dhv[e + 3] = d4.w;
            }
            float zi[8], zf[8], zg[8], zo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int j = b8 * 8 + e;
              const float dh = dhv[e] + dh_rec[j];
              const float tc_ = tanh_(cc[e]);
              const float dc = dc_next[j] + dh * go[e] * (1.0f - tc_ * tc_);
              zi[e] = dc * gg[e] * gi[e] * (1.0f - gi[e]);
              zf[e] = dc * cpv[e] * gf[e] * (1.0f - gf[e]);
              zg[e] = dc * gi[e] * (1.0f - gg[e] * gg[e]);
              zo[e] = dh * tc_ * go[e] * (1.0f - go[e]);
              dc_next[j] = dc * gf[e];
            }
            if (valid && !FAST) {   // dz overwrites the gates in the stash (fp32): the weight-gradient pass reads it
#pragma unroll
              for (int h4 = 0; h4 < 2; ++h4) {
                const int o = b8 * 8 + h4 * 4, e = h4 * 4;
                *reinterpret_cast<float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 0, u0 + o)) = make_float4(zi[e], zi[e + 1], zi[e + 2], zi[e + 3]);
                *reinterpret_cast<float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 1, u0 + o)) = make_float4(zf[e], zf[e + 1], zf[e + 2], zf[e + 3]);
                *reinterpret_cast<float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 2, u0 + o)) = make_float4(zg[e], zg[e + 1], zg[e + 2], zg[e + 3]);
                *reinterpret_cast<float4*>(stash + sidx<FAST>(d, sp, sq, l, t, 3, u0 + o)) = make_float4(zo[e], zo[e + 1], zo[e + 2], zo[e + 3]);
              }
            }
            // fp16 A tile of the next MMA: column g*64 + (unit within this CTA)
            const int cu = hh * UT + b8 * 8;
            const uint4 pz[4] = {make_uint4(pack_half2(zi[0], zi[1]), pack_half2(zi[2], zi[3]), pack_half2(zi[4], zi[5]), pack_half2(zi[6], zi[7])),
                                 make_uint4(pack_half2(zf[0], zf[1]), pack_half2(zf[2], zf[3]), pack_half2(zf[4], zf[5]), pack_half2(zf[6], zf[7])),
                                 make_uint4(pack_half2(zg[0], zg[1]), pack_half2(zg[2], zg[3]), pack_half2(zg[4], zg[5]), pack_half2(zg[6], zg[7])),
                                 make_uint4(pack_half2(zo[0], zo[1]), pack_half2(zo[2], zo[3]), pack_half2(zo[4], zo[5]), pack_half2(zo[6], zo[7]))};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              sts128(sDZ + tile_chunk_off(TS, row, g * 64 + cu), pz[g]);
              if (FAST && valid) *reinterpret_cast<uint4*>(dz16 + idx16(sp, ((size_t)l * T + t) * 4 + g, u0 + b8 * 8)) = pz[g];
            }
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[B_A]);
          if (l > 0 && !first) {   // dx of step t+1 (stored at the top of this round) -> the pair of the layer below
            if (lane == 0) prog_arrive(prog_mine);
          }
        }
        // ---- the layer's last MMA (step 0): only its dx columns are needed (layers above 0) ----
        ok = mbar_wait(&bars[B_Z], (q - 1) & 1) && ok;
        fence_after();
        if (l > 0) {
          uint32_t v[32];
          float* dst = dsW + (size_t)rank * pstr;
#pragma unroll
          for (int c = 0; c < 1; ++c) {
            tmem_ld32(tm + lane_base + hh * 32 + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4)
              if (valid) *reinterpret_cast<float4*>(dst + didx<FAST>(d, sp, sq, 0, hh * 32 + j4 * 4)) = make_float4(__uint_as_float(v[j4 * 4]), __uint_as_float(v[j4 * 4 + 1]),
                                                                             __uint_as_float(v[j4 * 4 + 2]), __uint_as_float(v[j4 * 4 + 3]));
          }
          __syncwarp();
          if (lane == 0) prog_arrive(prog_mine);       // dx of step 0: T arrivals per warp in all
        }
        fence_before();
      }
      // every warp of this CTA is past the layer before the weight tile and sDZ are rewritten
      __syncthreads();
      fence_after();
    }
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  fence_before();
  __syncthreads();
  cluster_sync_all();     // no remote store or arrival may target a CTA that has exited
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

}  // namespace

extern unsigned g_lstm_call_id;   // lstm_tc.cu

// loss, head gradient and full BPTT: dz into the stash (comps 0..3), head entries + loss sum into the per-block partials
int lstm_tc_bwd(const lstm::LstmDesc& d, int loss, const float* params, const float* y, float* yhat_dy, size_t n_seq,
                float* stash, float* dseqP, float* partials, int n_blocks, unsigned* err_flag, unsigned* prog, uint16_t* dz16, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(lstm_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    DLAS_CUDA(cudaFuncSetAttribute(lstm_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  ++g_lstm_call_id;
  lstm_head_grad_kernel<<<n_blocks, 256, 0, st>>>(d, loss, y, yhat_dy, stash, n_seq, partials);
  DLAS_LAUNCH_CHECK();
  const size_t n_tiles = (n_seq + TS - 1) / TS;
  const int max_lanes = (sm_count() / 2) / d.L;
  if (max_lanes < 1) { set_error("lstm (tc): %d layers need more SM pairs than this GPU has", d.L); return DLAS_ERR_UNSUPPORTED_SHAPE; }
  const int pairs = (int)(n_tiles < (size_t)max_lanes ? n_tiles : (size_t)max_lanes) * d.L;
  DLAS_CUDA(cudaMemsetAsync(prog, 0, n_tiles * d.L * sizeof(unsigned), st));
  if (d.fast) lstm_bwd_tc_kernel<true><<<2 * pairs, NT, SMEM_BYTES, st>>>(d, params, stash, yhat_dy, dseqP, n_seq, err_flag, g_lstm_call_id, prog, dz16);
  else lstm_bwd_tc_kernel<false><<<2 * pairs, NT, SMEM_BYTES, st>>>(d, params, stash, yhat_dy, dseqP, n_seq, err_flag, g_lstm_call_id, prog, dz16);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

}  // namespace dlas
