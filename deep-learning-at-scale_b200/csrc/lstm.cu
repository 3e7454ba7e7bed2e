// LSTM regressor (Readme.md:21 names the model; BASELINE.json config 5 fixes 2 layers, hidden 128,
// seq 48): fp32 forward, full BPTT and weight gradients on CUDA cores.
//
// Sequences are independent, so the recurrence needs no grid-wide step barrier: each CTA walks a
// tile of BT=8 sequences through all T steps of a layer with the hidden and cell state in shared
// memory (persistent recurrence), reading the gate weights through L2 (they are 0.8 MB and stay
// resident there).  Three kernels per training step:
//   1. lstm_fwd_kernel   forward over layers and time; stashes i,f,g,o,c,h per (seq,layer,t)
//   2. lstm_bwd_kernel   loss + head gradient, then reverse time per layer: gate gradients dz
//                        overwrite i,f,g,o in the stash, dx feeds the layer below
//   3. lstm_dw_kernel    dK = [x_t, h_{t-1}]^T dz as one long-K reduction over all (seq,t) rows,
//                        per-CTA partials, added in CTA order by lstm_reduce_kernel
// Cell (Keras gate order i,f,c,o; sigmoid recurrent activation, tanh cell/output activation):
//   z = [x_t, h_{t-1}] K + b;  c_t = s(z_f) c_{t-1} + s(z_i) tanh(z_c);  h_t = s(z_o) tanh(c_t)
// Flat layout: per layer K[(d_l+H)][4H] (input rows then recurrent rows), b[4H]; head w[H], b[1].
#include "common.cuh"
#include "simt_tiles.cuh"
#include "lstm_common.cuh"

namespace dlas {
// lstm_tc.cu: tcgen05 stages, interchangeable with the fp32 kernels below (same stash and parameter layout)
bool lstm_tc_supported(const lstm::LstmDesc& d);
size_t lstm_tc_dw_partial_floats(const lstm::LstmDesc& d);
int lstm_tc_fwd(const lstm::LstmDesc& d, const float* params, const float* X, size_t n_seq, float* stash, float* yhat,
                unsigned* err_flag, unsigned* prog, uint16_t* h16, uint16_t* g16, cudaStream_t st);
int lstm_tc_bwd(const lstm::LstmDesc& d, int loss, const float* params, const float* y, float* yhat_dy, size_t n_seq,
                float* stash, float* dseqP, float* partials, int n_blocks, unsigned* err_flag, unsigned* prog, uint16_t* dz16, cudaStream_t st);
int lstm_tc_dw(const lstm::LstmDesc& d, const float* X, const float* stash, size_t n_seq, float* grad_sum, float* tc_partials,
               unsigned* err_flag, const uint16_t* h16, const uint16_t* dz16, cudaStream_t st);

namespace {
using namespace lstm;
// stages of DLAS_MODE_TC (bits of the internal mode word; DLAS_MODE_TC = every stage the shape supports; the tests pass
// 0x100 | bits to exchange single stages)
enum { TC_FWD = 1, TC_BWD = 2, TC_DW = 4 };

constexpr int BT = 8;      // sequences per recurrence tile
constexpr int BS = 12;     // padded row stride of the transposed recurrence tiles
constexpr int NT = 512;
constexpr int DR = 32;     // (seq,t) rows per weight-gradient tile
constexpr int DRS = 36;
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// smem: xh[(dmax+H)][BS] | z[4H][BS] | c[H][BS] | aux[H][BS]
struct RecSmem { float *xh, *z, *c, *aux; };
__device__ __forceinline__ RecSmem carve(float* smem, const LstmDesc& d) {
  RecSmem s;
  int dmax = d.d_in > d.H ? d.d_in : d.H;
  s.xh = smem;
  s.z = s.xh + (dmax + d.H) * BS;
  s.c = s.z + 4 * d.H * BS;
  s.aux = s.c + d.H * BS;
  return s;
}
__host__ __device__ inline size_t rec_smem_bytes(const LstmDesc& d) {
  int dmax = d.d_in > d.H ? d.d_in : d.H;
  return (size_t)((dmax + d.H) + 4 * d.H + 2 * d.H) * BS * sizeof(float) + 64 * sizeof(float);
}

__global__ void __launch_bounds__(NT, 1)
lstm_fwd_kernel(const LstmDesc d, const float* __restrict__ params, const float* __restrict__ X,
                const size_t n_seq, float* __restrict__ stash, float* __restrict__ yhat) {
  extern __shared__ __align__(16) float smem[];
  RecSmem s = carve(smem, d);
  const int H = d.H, T = d.T;
  const size_t n_tiles = (n_seq + BT - 1) / BT;
  for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const size_t s0 = tile * BT;
    const int ns = (int)((n_seq - s0) < (size_t)BT ? (n_seq - s0) : (size_t)BT);
    for (int l = 0; l < d.L; ++l) {
      const int dl = d.d_l[l];
      const float* K = params + d.k_off[l];
      const float* bias = params + d.b_off[l];
      for (int idx = threadIdx.x; idx < H * BS; idx += NT) { s.c[idx] = 0.0f; s.xh[dl * BS + idx] = 0.0f; }
      for (int t = 0; t < T; ++t) {
        // input rows of this step, transposed into xh[0..dl)
        for (int idx = threadIdx.x; idx < BT * dl; idx += NT) {
          int r = idx / dl, k = idx - r * dl;
          float v = 0.0f;
          if (r < ns) v = (l == 0) ? X[((s0 + r) * T + t) * (size_t)dl + k] : stash[stash_idx(d, s0 + r, l - 1, t, 5) + k];
          s.xh[k * BS + r] = v;
        }
        __syncthreads();
        gemm_dispatch<BT, BS, NT>(s.xh, K, 4 * H, dl + H, 4 * H, [&](int j, int r0, auto& acc) {
          const float b = __ldg(bias + j);
          constexpr int RT = sizeof(acc) / sizeof(float);
#pragma unroll
          for (int i = 0; i < RT; ++i) s.z[j * BS + r0 + i] = acc[i] + b;
        });
        __syncthreads();
        for (int idx = threadIdx.x; idx < BT * H; idx += NT) {
          int r = idx / H, u = idx - r * H;
          float gi = sigmoidf_(s.z[u * BS + r]);
          float gf = sigmoidf_(s.z[(H + u) * BS + r]);
          float gg = tanhf(s.z[(2 * H + u) * BS + r]);
          float go = sigmoidf_(s.z[(3 * H + u) * BS + r]);
          float c = gf * s.c[u * BS + r] + gi * gg;
          float h = go * tanhf(c);
          s.c[u * BS + r] = c;
          s.xh[(dl + u) * BS + r] = h;
          if (r < ns) {
            size_t b0 = stash_idx(d, s0 + r, l, t, 0) + u;
            stash[b0] = gi; stash[b0 + H] = gf; stash[b0 + 2 * (size_t)H] = gg;
            stash[b0 + 3 * (size_t)H] = go; stash[b0 + 4 * (size_t)H] = c; stash[b0 + 5 * (size_t)H] = h;
          }
        }
        __syncthreads();
      }
    }
    if (yhat) {  // head: Dense(H -> 1) on the last hidden state of the top layer
      const int dl = d.d_l[d.L - 1];
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      for (int r = warp; r < ns; r += NT / 32) {
        float a = 0.0f;
        for (int u = lane; u < H; u += 32) a += s.xh[(dl + u) * BS + r] * __ldg(params + d.hw_off + u);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
        if (lane == 0) yhat[s0 + r] = a + __ldg(params + d.hb_off);
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(NT, 1)
lstm_bwd_kernel(const LstmDesc d, const int loss, const float* __restrict__ params,
                const float* __restrict__ KT, const float* __restrict__ y, const float* __restrict__ yhat,
                const size_t n_seq, float* __restrict__ stash, float* __restrict__ dseq,
                float* __restrict__ partials) {
  extern __shared__ __align__(16) float smem[];
  RecSmem s = carve(smem, d);
  float* dy = smem + (rec_smem_bytes(d) / sizeof(float)) - 64;  // [BT] + reduction scratch
  const int H = d.H, T = d.T;
  float* my_partial = partials + (size_t)blockIdx.x * pstride(d.n_params);
  float acc_hw = 0.0f;              // thread u < H: head weight gradient; thread H: head bias; thread H+1: loss
  float acc_misc = 0.0f;
  const size_t n_tiles = (n_seq + BT - 1) / BT;
  for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const size_t s0 = tile * BT;
    const int ns = (int)((n_seq - s0) < (size_t)BT ? (n_seq - s0) : (size_t)BT);
    if (threadIdx.x < BT) {
      float g = 0.0f, lo = 0.0f;
      if ((int)threadIdx.x < ns) {
        float dd = yhat[s0 + threadIdx.x] - y[s0 + threadIdx.x];
        if (loss == DLAS_LOSS_MSE) { lo = dd * dd; g = 2.0f * dd; }
        else { float a = fabsf(dd); lo = fminf(a, 6.0f); g = (a <= 6.0f) ? (dd > 0.0f ? 1.0f : (dd < 0.0f ? -1.0f : 0.0f)) : 0.0f; }
      }
      dy[threadIdx.x] = g;
      dy[BT + threadIdx.x] = lo;
    }
    __syncthreads();
    if ((int)threadIdx.x < H) {
      for (int r = 0; r < ns; ++r) acc_hw += stash[stash_idx(d, s0 + r, d.L - 1, T - 1, 5) + threadIdx.x] * dy[r];
    } else if ((int)threadIdx.x == H) {
      for (int r = 0; r < ns; ++r) acc_misc += dy[r];
    } else if ((int)threadIdx.x == H + 1) {
      for (int r = 0; r < ns; ++r) acc_misc += dy[BT + r];
    }
    for (int l = d.L - 1; l >= 0; --l) {
      const int dl = d.d_l[l];
      const float* KTl = KT + d.k_off[l];       // [4H][dl+H]
      // aux = dh_next, c = dc_next
      for (int idx = threadIdx.x; idx < H * BS; idx += NT) { s.aux[idx] = 0.0f; s.c[idx] = 0.0f; }
      __syncthreads();
      for (int t = T - 1; t >= 0; --t) {
        for (int idx = threadIdx.x; idx < BT * H; idx += NT) {
          int r = idx / H, u = idx - r * H;
          float dzi = 0.0f, dzf = 0.0f, dzg = 0.0f, dzo = 0.0f, dcn = 0.0f;
          if (r < ns) {
            size_t b0 = stash_idx(d, s0 + r, l, t, 0) + u;
            float gi = stash[b0], gf = stash[b0 + H], gg = stash[b0 + 2 * (size_t)H], go = stash[b0 + 3 * (size_t)H];
            float c = stash[b0 + 4 * (size_t)H];
            float cprev = t ? stash[stash_idx(d, s0 + r, l, t - 1, 4) + u] : 0.0f;
            float dh_in;
            if (l == d.L - 1) dh_in = (t == T - 1) ? dy[r] * __ldg(params + d.hw_off + u) : 0.0f;
            else dh_in = dseq[((s0 + r) * T + t) * (size_t)H + u];
            float dh = dh_in + s.aux[u * BS + r];
            float tc = tanhf(c);
            float dc = s.c[u * BS + r] + dh * go * (1.0f - tc * tc);
            dzi = dc * gg * gi * (1.0f - gi);
            dzf = dc * cprev * gf * (1.0f - gf);
            dzg = dc * gi * (1.0f - gg * gg);
            dzo = dh * tc * go * (1.0f - go);
            dcn = dc * gf;
            stash[b0] = dzi; stash[b0 + H] = dzf; stash[b0 + 2 * (size_t)H] = dzg; stash[b0 + 3 * (size_t)H] = dzo;
          }
          s.c[u * BS + r] = dcn;
          s.z[u * BS + r] = dzi; s.z[(H + u) * BS + r] = dzf; s.z[(2 * H + u) * BS + r] = dzg; s.z[(3 * H + u) * BS + r] = dzo;
        }
        __syncthreads();
        // d[x_t, h_{t-1}] = dz K^T
        gemm_dispatch<BT, BS, NT>(s.z, KTl, dl + H, 4 * H, dl + H, [&](int k, int r0, auto& acc) {
          constexpr int RT = sizeof(acc) / sizeof(float);
#pragma unroll
          for (int i = 0; i < RT; ++i) {
            int r = r0 + i;
            if (k >= dl) s.aux[(k - dl) * BS + r] = acc[i];
            else if (l > 0 && r < ns) dseq[((s0 + r) * T + t) * (size_t)H + k] = acc[i];
          }
        });
        __syncthreads();
      }
    }
  }
  if ((int)threadIdx.x < H) my_partial[d.hw_off + threadIdx.x] = acc_hw;
  else if ((int)threadIdx.x == H) my_partial[d.hb_off] = acc_misc;
  else if ((int)threadIdx.x == H + 1) my_partial[d.n_params] = acc_misc;
}

// dK_l, db_l: rows are (seq,t) pairs; A row = [input(seq,t), h_l(seq,t-1)], dz row from the stash.
__global__ void __launch_bounds__(NT, 1)
lstm_dw_kernel(const LstmDesc d, const float* __restrict__ X, const float* __restrict__ stash,
               const size_t n_seq, float* __restrict__ partials) {
  extern __shared__ __align__(16) float smem[];
  const int H = d.H, T = d.T;
  const int dmax = d.d_in > H ? d.d_in : H;
  float* aT = smem;                       // [(dmax+H)][DRS]
  float* dzT = smem + (dmax + H) * DRS;   // [4H][DRS]
  float* my_partial = partials + (size_t)blockIdx.x * pstride(d.n_params);
  const size_t n_rows = n_seq * T;
  const size_t n_tiles = (n_rows + DR - 1) / DR;
  for (int l = 0; l < d.L; ++l) {
    const int dl = d.d_l[l];
    bool first = true;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const size_t r0 = tile * DR;
      const int nr = (int)((n_rows - r0) < (size_t)DR ? (n_rows - r0) : (size_t)DR);
      for (int idx = threadIdx.x; idx < DR * (dl + H); idx += NT) {
        int r = idx / (dl + H), k = idx - r * (dl + H);
        float v = 0.0f;
        if (r < nr) {
          size_t row = r0 + r, seq = row / T;
          int t = (int)(row - seq * T);
          if (k < dl) v = (l == 0) ? X[row * (size_t)dl + k] : stash[stash_idx(d, seq, l - 1, t, 5) + k];
          else v = t ? stash[stash_idx(d, seq, l, t - 1, 5) + (k - dl)] : 0.0f;
        }
        aT[k * DRS + r] = v;
      }
      for (int idx = threadIdx.x; idx < DR * 4 * H; idx += NT) {
        int r = idx / (4 * H), j = idx - r * (4 * H);
        float v = 0.0f;
        if (r < nr) {
          size_t row = r0 + r, seq = row / T;
          int t = (int)(row - seq * T);
          v = stash[stash_idx(d, seq, l, t, 0) + j];   // comps 0..3 are contiguous: dz_i, dz_f, dz_g, dz_o
        }
        dzT[j * DRS + r] = v;
      }
      __syncthreads();
      accumulate_dw<DR, DRS, NT>(aT, dzT, dl + H, 4 * H, my_partial + d.k_off[l], my_partial + d.b_off[l], first);
      first = false;
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(256)
transpose_kernel(const LstmDesc d, const float* __restrict__ params, float* __restrict__ KT) {
  for (int l = 0; l < d.L; ++l) {
    const int K = d.d_l[l] + d.H, N = 4 * d.H;
    const float* W = params + d.k_off[l];
    float* Tm = KT + d.k_off[l];
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < K * N; idx += gridDim.x * 256) {
      int j = idx / K, k = idx - j * K;
      Tm[idx] = W[(size_t)k * N + j];
    }
  }
}

__global__ void __launch_bounds__(256)
lstm_reduce_kernel(const float* __restrict__ partials, size_t stride, int n_active, size_t i0, size_t count,
                   float* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= count) return;
  float s = 0.0f;
  for (int c = 0; c < n_active; ++c) s += partials[(size_t)c * stride + i0 + i];
  out[i0 + i] = s;
}

struct Ws { float *stash, *dseq, *KT, *partials, *yhat, *tcp, *dseqP; unsigned *status, *prog; uint16_t *h16, *dz16; size_t bytes; };
Ws carve_ws(const LstmDesc& d, size_t n_seq, void* base) {
  Ws w;
  size_t o = 0;
  auto take = [&](size_t n) { float* p = base ? (float*)base + o : nullptr; o += (n + 63) & ~(size_t)63; return p; };
  const size_t n_pad = (n_seq + 31) & ~(size_t)31;       // the tensor-core kernels' stash layout blocks sequences by 32
  w.stash = take(n_pad * d.L * d.T * 6 * (size_t)d.H);
  w.dseq = take(d.L > 1 ? n_seq * d.T * (size_t)d.H : 0);
  w.KT = take(d.n_params);
  w.partials = take((size_t)sm_count() * pstride(d.n_params));
  w.yhat = take(n_seq);
  w.tcp = take(lstm_tc_supported(d) ? lstm_tc_dw_partial_floats(d) : 0);      // tensor-core weight-gradient partials
  // tensor-core BPTT: the two CTAs of a pair write partial dx of a layer into separate buffers (x 2 sets from 3 layers up)
  w.dseqP = take(lstm_tc_supported(d) && d.L > 1 ? (size_t)(d.L >= 3 ? 4 : 2) * n_pad * d.T * (size_t)d.H : 0);
  w.status = reinterpret_cast<unsigned*>(take(64));                              // status words of the tensor-core kernels
  // layer pipeline of the tensor-core recurrences: one progress counter per (128-sequence tile, layer)
  w.prog = reinterpret_cast<unsigned*>(take(lstm_tc_supported(d) ? ((n_seq + 127) / 128) * (size_t)d.L : 0));
  // all-tensor-core step: h and dz as fp16 operands (lstm_common.cuh idx16), halves counted as half floats
  w.h16 = reinterpret_cast<uint16_t*>(take(lstm_tc_supported(d) ? n_pad * d.L * d.T * (size_t)d.H / 2 : 0));
  w.dz16 = reinterpret_cast<uint16_t*>(take(lstm_tc_supported(d) ? n_pad * d.L * d.T * 4 * (size_t)d.H / 2 : 0));
  w.bytes = o * sizeof(float);
  return w;
}

template <class Kern>
int set_smem(Kern k, size_t bytes) {
  if (bytes > 227 * 1024) {
    set_error("lstm: shape needs %zu bytes of shared memory per CTA", bytes);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  DLAS_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  return DLAS_OK;
}

}  // namespace
}  // namespace dlas

using namespace dlas;

namespace dlas {
namespace {
// mode word -> stage bits
int stage_bits(int mode, const LstmDesc& d, const char* fn, int* bits) {
  *bits = 0;
  if (mode == DLAS_MODE_FP32) return DLAS_OK;
  if (mode != DLAS_MODE_TC && (mode & ~0x7) != 0x100) { set_error("%s: unknown mode %d", fn, mode); return DLAS_ERR_INVALID; }
  if (!lstm_tc_supported(d)) {
    set_error("%s: DLAS_MODE_TC needs hidden = 128 and d_in <= 16 (got hidden %d, d_in %d)", fn, d.H, d.d_in);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  *bits = (mode == DLAS_MODE_TC) ? (TC_FWD | TC_BWD | TC_DW) : (mode & 0x7);
  return DLAS_OK;
}
}  // namespace
}  // namespace dlas

extern "C" size_t dlas_lstm_n_params(int d_in, int hidden, int n_layers) {
  LstmDesc d;
  if (make_desc(d_in, hidden, n_layers, 1, d) != DLAS_OK) return 0;
  return (size_t)d.n_params;
}

extern "C" size_t dlas_lstm_workspace_bytes(int d_in, int hidden, int n_layers, int T, size_t n_seq) {
  LstmDesc d;
  if (make_desc(d_in, hidden, n_layers, T, d) != DLAS_OK) return 0;
  return carve_ws(d, n_seq, nullptr).bytes;
}

extern "C" int dlas_lstm_predict(int d_in, int hidden, int n_layers, int T, const float* params, const float* X,
                                 size_t n_seq, float* yhat, void* workspace, size_t workspace_bytes, void* stream) {
  return dlas_lstm_predict_mode(DLAS_MODE_FP32, d_in, hidden, n_layers, T, params, X, n_seq, yhat, workspace, workspace_bytes, stream);
}

extern "C" int dlas_lstm_predict_mode(int mode, int d_in, int hidden, int n_layers, int T, const float* params, const float* X,
                                      size_t n_seq, float* yhat, void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  LstmDesc d;
  int rc = make_desc(d_in, hidden, n_layers, T, d);
  if (rc != DLAS_OK) return rc;
  int bits = 0;
  rc = stage_bits(mode, d, "dlas_lstm_predict_mode", &bits);
  if (rc != DLAS_OK) return rc;
  if (bits & TC_FWD) { d.fast = 1; d.n_pad = (n_seq + 31) & ~(size_t)31; }   // only tensor-core kernels touch the stash here
  if (n_seq == 0) return DLAS_OK;
  DLAS_REQUIRE(params && X && yhat, "dlas_lstm_predict: null pointer");
  Ws w = carve_ws(d, n_seq, workspace);
  if (!workspace || workspace_bytes < w.bytes) {
    set_error("dlas_lstm_predict: workspace of %zu bytes given, %zu needed", workspace_bytes, w.bytes);
    return DLAS_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  size_t smem = rec_smem_bytes(d);
  rc = set_smem(lstm_fwd_kernel, smem);
  if (rc != DLAS_OK) return rc;
  if (bits & TC_FWD) return lstm_tc_fwd(d, params, X, n_seq, w.stash, yhat, w.status, w.prog, w.h16, w.dz16, st);
  lstm_fwd_kernel<<<sm_count(), NT, smem, st>>>(d, params, X, n_seq, w.stash, yhat);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_lstm_mode_supported(int d_in, int hidden, int n_layers, int mode) {
  LstmDesc d;
  if (make_desc(d_in, hidden, n_layers, 1, d) != DLAS_OK) return 0;
  return mode == DLAS_MODE_FP32 || (mode == DLAS_MODE_TC && lstm_tc_supported(d)) ? 1 : 0;
}

extern "C" int dlas_lstm_grad(int d_in, int hidden, int n_layers, int T, int loss, const float* params,
                              const float* X, const float* y, size_t n_seq, float* grad_sum, void* workspace,
                              size_t workspace_bytes, void* stream) {
  return dlas_lstm_grad_mode(DLAS_MODE_FP32, d_in, hidden, n_layers, T, loss, params, X, y, n_seq, grad_sum, workspace,
                             workspace_bytes, stream);
}

extern "C" int dlas_lstm_grad_mode(int mode, int d_in, int hidden, int n_layers, int T, int loss, const float* params,
                                   const float* X, const float* y, size_t n_seq, float* grad_sum, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  LstmDesc d;
  int rc = make_desc(d_in, hidden, n_layers, T, d);
  if (rc != DLAS_OK) return rc;
  int bits = 0;
  rc = stage_bits(mode, d, "dlas_lstm_grad_mode", &bits);
  if (rc != DLAS_OK) return rc;
  // every stage on the tensor cores: the stash may use their own (coalesced) layout; a mix keeps the fp32 kernels' layout
  if ((bits & (TC_FWD | TC_BWD | TC_DW)) == (TC_FWD | TC_BWD | TC_DW)) { d.fast = 1; d.n_pad = (n_seq + 31) & ~(size_t)31; }
  DLAS_REQUIRE(loss == DLAS_LOSS_MSE || loss == DLAS_LOSS_MAE_CLIP, "dlas_lstm_grad: unknown loss %d", loss);
  DLAS_REQUIRE(params && grad_sum && (n_seq == 0 || (X && y)), "dlas_lstm_grad: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_seq == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)d.n_params + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  Ws w = carve_ws(d, n_seq, workspace);
  if (!workspace || workspace_bytes < w.bytes) {
    set_error("dlas_lstm_grad: workspace of %zu bytes given, %zu needed", workspace_bytes, w.bytes);
    return DLAS_ERR_WORKSPACE;
  }
  const int grid = sm_count();
  size_t smem = rec_smem_bytes(d);
  rc = set_smem(lstm_fwd_kernel, smem);
  if (rc != DLAS_OK) return rc;
  rc = set_smem(lstm_bwd_kernel, smem);
  if (rc != DLAS_OK) return rc;
  const int dmax = d.d_in > d.H ? d.d_in : d.H;
  size_t smem_dw = (size_t)((dmax + d.H) + 4 * d.H) * DRS * sizeof(float);
  rc = set_smem(lstm_dw_kernel, smem_dw);
  if (rc != DLAS_OK) return rc;
  if (!(bits & TC_BWD)) {       // K^T for the fp32 backward recurrence
    transpose_kernel<<<64, 256, 0, st>>>(d, params, w.KT);
    DLAS_LAUNCH_CHECK();
  }
  if (bits & TC_FWD) {   // persistent recurrence on the tensor cores
    rc = lstm_tc_fwd(d, params, X, n_seq, w.stash, w.yhat, w.status, w.prog, w.h16, w.dz16, st);
    if (rc != DLAS_OK) return rc;
  } else {
    lstm_fwd_kernel<<<grid, NT, smem, st>>>(d, params, X, n_seq, w.stash, w.yhat);
    DLAS_LAUNCH_CHECK();
  }
  if (bits & TC_BWD) {   // loss + head gradient, then the backward recurrence on the tensor cores
    rc = lstm_tc_bwd(d, loss, params, y, w.yhat, n_seq, w.stash, w.dseqP, w.partials, grid, w.status, w.prog, w.dz16, st);
    if (rc != DLAS_OK) return rc;
  } else {
    lstm_bwd_kernel<<<grid, NT, smem, st>>>(d, loss, params, w.KT, y, w.yhat, n_seq, w.stash, w.dseq, w.partials);
    DLAS_LAUNCH_CHECK();
  }
  size_t n_layer_params = (size_t)d.hw_off;
  if (bits & TC_DW) {   // weight gradient as a tcgen05 GEMM over all (sequence, t) rows
    rc = lstm_tc_dw(d, X, w.stash, n_seq, grad_sum, w.tcp, w.status, w.h16, w.dz16, st);
    if (rc != DLAS_OK) return rc;
  } else {
    lstm_dw_kernel<<<grid, NT, smem_dw, st>>>(d, X, w.stash, n_seq, w.partials);
    DLAS_LAUNCH_CHECK();
    // layer parameters: CTAs that owned at least one (seq,t) tile; head + loss: every CTA wrote its slot
    size_t n_tiles = (n_seq * (size_t)T + DR - 1) / DR;
    int n_active = (int)(n_tiles < (size_t)grid ? n_tiles : (size_t)grid);
    lstm_reduce_kernel<<<(unsigned)((n_layer_params + 255) / 256), 256, 0, st>>>(w.partials, pstride(d.n_params), n_active,
                                                                               0, n_layer_params, grad_sum);
    DLAS_LAUNCH_CHECK();
  }
  size_t n_head = (size_t)d.H + 2;  // w[H], b, loss
  lstm_reduce_kernel<<<(unsigned)((n_head + 255) / 256), 256, 0, st>>>(w.partials, pstride(d.n_params), grid,
                                                                     n_layer_params, n_head, grad_sum);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}
