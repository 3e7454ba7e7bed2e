// The CNN of the reference script (cnn.py:110-114), fp32 on CUDA cores:
//   Convolution1D(1, F=100, K=13, relu, valid) -> Dropout(0.5) -> Flatten -> Dense(T*F=3600, 12)
// trained with mae_clip / SGD at batch 20 (cnn.py:117-128).  At that batch a step is launch- and latency-bound
// (20 rows x 180 Kflop), so the path is six small deterministic kernels, no atomics, each shaped to have thousands of
// threads even at 20 rows:
//   conv+relu+dropout -> dense forward + loss (block per row) -> dWd partials -> dflat/dconv -> dWc partials (warp per
//   entry) -> reduce (+ the optimiser update in dlas_cnn_fit_steps, which runs a whole epoch's batch loop in one call).
// Row groups give the weight-gradient kernels parallelism at large batches; partials are added in
// group order.  Dropout is "inverted" (x * mask / keep at train time, identity at test time); the mask
// bit of (sample, unit) is the counter hash the oracle uses, so both draw the same mask.
// Flat layout: Wc[K][F] (tap-major), bc[F], Wd[T*F][n_out] (Flatten is time-major: t*F + f), bd[n_out].
#include "common.cuh"
#include "optim_math.cuh"

namespace dlas {
namespace {

constexpr int MAXO = 16;            // max Dense outputs (cnn.py: 12)
constexpr uint64_t STREAM_DROPOUT = 300;

struct CnnDesc {
  int L, K, F, O, T, TF;
  int wc, bc, wd, bd, n_params;
};

int make_desc(int L, int K, int F, int O, CnnDesc& d) {
  DLAS_REQUIRE(L >= 1 && K >= 1 && K <= L && F >= 1 && O >= 1 && O <= MAXO, "cnn: bad shape L=%d K=%d F=%d n_out=%d (n_out <= %d)", L, K, F, O, MAXO);
  d.L = L; d.K = K; d.F = F; d.O = O; d.T = L - K + 1; d.TF = d.T * F;
  d.wc = 0; d.bc = K * F; d.wd = d.bc + F; d.bd = d.wd + d.TF * O; d.n_params = d.bd + O;
  return DLAS_OK;
}

// a[b][t*F+f] = dropout(relu(bc[f] + sum_k x[b][t+k] Wc[k][f]))
__global__ void __launch_bounds__(256)
conv_kernel(const CnnDesc d, const float* __restrict__ p, const float* __restrict__ X, size_t n_rows, float keep,
            uint64_t drop_key, uint64_t sample0, float* __restrict__ a) {
  const size_t total = n_rows * (size_t)d.TF, stride = (size_t)gridDim.x * 256;
  const uint32_t thresh = (uint32_t)(keep * 16777216.0f);
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < total; e += stride) {
    const size_t b = e / d.TF;
    const int j = (int)(e - b * d.TF), t = j / d.F, f = j - t * d.F;
    const float* x = X + b * d.L + t;
    float s = __ldg(p + d.bc + f);
    for (int k = 0; k < d.K; ++k) s = fmaf(__ldg(x + k), __ldg(p + d.wc + k * d.F + f), s);
    s = fmaxf(s, 0.f);
    if (keep < 1.0f) s = (hash_u24(drop_key, (sample0 + b) * (uint64_t)d.TF + (uint64_t)j) < thresh) ? s / keep : 0.f;
    a[e] = s;
  }
}

// one block per row (the script trains on 20 rows: a warp per row would leave 20 warps on the whole GPU): the eight warps
// split the T*F inputs, warp sums are added in warp order; out = a Wd + bd; per-row loss; dout = d(sum of losses)/d(out)
__global__ void __launch_bounds__(256)
dense_kernel(const CnnDesc d, int loss, const float* __restrict__ p, const float* __restrict__ a, const float* __restrict__ y,
             size_t n_rows, float* __restrict__ out, float* __restrict__ dout, float* __restrict__ lossrow) {
  __shared__ float part[8][MAXO];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (size_t b = blockIdx.x; b < n_rows; b += gridDim.x) {
    float acc[MAXO];
#pragma unroll
    for (int o = 0; o < MAXO; ++o) acc[o] = 0.f;
    const float* ar = a + b * d.TF;
    for (int j = threadIdx.x; j < d.TF; j += 256) {
      const float av = ar[j];
      const float* w = p + d.wd + (size_t)j * d.O;
#pragma unroll
      for (int o = 0; o < MAXO; ++o)
        if (o < d.O) acc[o] = fmaf(av, __ldg(w + o), acc[o]);
    }
#pragma unroll
    for (int o = 0; o < MAXO; ++o)
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], s);
    if (lane == 0) {
#pragma unroll
      for (int o = 0; o < MAXO; ++o) part[warp][o] = acc[o];
    }
    __syncthreads();
    if (warp == 0) {
      float lo = 0.f;
      if (lane < d.O) {
        float v = 0.f;
        for (int w8 = 0; w8 < 8; ++w8) v += part[w8][lane];
        v += __ldg(p + d.bd + lane);
        out[b * d.O + lane] = v;
        if (y) {
          const float dd = v - y[b * d.O + lane], inv = 1.0f / (float)d.O;
          float g;
          if (loss == DLAS_LOSS_MSE) { lo = dd * dd * inv; g = 2.0f * dd * inv; }
          else { const float ab = fabsf(dd); lo = fminf(ab, 6.0f) * inv; g = (ab <= 6.0f) ? (dd > 0.f ? inv : (dd < 0.f ? -inv : 0.f)) : 0.f; }  // cnn.py:29-32
          dout[b * d.O + lane] = g;
        }
      }
      if (y) {
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) lo += __shfl_xor_sync(0xffffffffu, lo, s);
        if (lane == 0) lossrow[b] = lo;
      }
    }
    __syncthreads();
  }
}

// partial[g][j*O+o] = sum over the group's rows of a[b][j] dout[b][o]; entries TF*O.. hold dbd
__global__ void __launch_bounds__(256)
dwd_kernel(const CnnDesc d, const float* __restrict__ a, const float* __restrict__ dout, size_t n_rows, int rows_per_group,
           float* __restrict__ partial, size_t pstride) {
  const int g = blockIdx.y;
  const size_t b0 = (size_t)g * rows_per_group, b1 = min(n_rows, b0 + (size_t)rows_per_group);
  const int e = blockIdx.x * 256 + threadIdx.x, n_w = d.TF * d.O;
  if (e >= n_w + d.O) return;
  float s = 0.f;
  if (e < n_w) {
    const int j = e / d.O, o = e - j * d.O;
    for (size_t b = b0; b < b1; ++b) s = fmaf(a[b * d.TF + j], dout[b * d.O + o], s);
    partial[(size_t)g * pstride + d.wd + e] = s;
  } else {
    const int o = e - n_w;
    for (size_t b = b0; b < b1; ++b) s += dout[b * d.O + o];
    partial[(size_t)g * pstride + d.bd + o] = s;
  }
}

// in place: a[b][j] <- dconv = (a > 0) * (dout[b] . Wd[j]) / keep   (a > 0 implies relu' = 1 and mask = 1)
__global__ void __launch_bounds__(256)
dflat_kernel(const CnnDesc d, const float* __restrict__ p, const float* __restrict__ dout, size_t n_rows, float keep,
             float* __restrict__ a) {
  const size_t total = n_rows * (size_t)d.TF, stride = (size_t)gridDim.x * 256;
  const float scale = keep < 1.0f ? 1.0f / keep : 1.0f;
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < total; e += stride) {
    const size_t b = e / d.TF;
    const int j = (int)(e - b * d.TF);
    float s = 0.f;
    if (a[e] > 0.f) {
      const float* w = p + d.wd + (size_t)j * d.O;
      for (int o = 0; o < d.O; ++o) s = fmaf(dout[b * d.O + o], __ldg(w + o), s);
      s *= scale;
    }
    a[e] = s;
  }
}

// partial[g][k*F+f] = sum_b sum_t dconv[b][t][f] x[b][t+k]; entries K*F.. hold dbc.  One WARP per entry: the lanes split
// the (row, time) pairs of the group and their sums meet in a shuffle tree (1,400 entries only: a thread per entry would
// walk 720 dependent steps at the script's batch of 20)
__global__ void __launch_bounds__(256)
dwc_kernel(const CnnDesc d, const float* __restrict__ dconv, const float* __restrict__ X, size_t n_rows, int rows_per_group,
           float* __restrict__ partial, size_t pstride) {
  const int g = blockIdx.y, lane = threadIdx.x & 31;
  const size_t b0 = (size_t)g * rows_per_group, b1 = min(n_rows, b0 + (size_t)rows_per_group);
  const int e = blockIdx.x * 8 + (threadIdx.x >> 5), n_w = d.K * d.F;
  if (e >= n_w + d.F) return;
  const int n_idx = (int)(b1 - b0) * d.T;
  float s = 0.f;
  if (e < n_w) {
    const int k = e / d.F, f = e - k * d.F;
    for (int idx = lane; idx < n_idx; idx += 32) {
      const size_t b = b0 + idx / d.T;
      const int t = idx % d.T;
      s = fmaf(dconv[b * d.TF + t * d.F + f], __ldg(X + b * d.L + t + k), s);
    }
  } else {
    const int f = e - n_w;
    for (int idx = lane; idx < n_idx; idx += 32) {
      const size_t b = b0 + idx / d.T;
      const int t = idx % d.T;
      s += dconv[b * d.TF + t * d.F + f];
    }
  }
#pragma unroll
  for (int w = 16; w > 0; w >>= 1) s += __shfl_xor_sync(0xffffffffu, s, w);
  if (lane == 0) partial[(size_t)g * pstride + (e < n_w ? d.wc + e : d.bc + (e - n_w))] = s;
}

// FUSE: the optimiser update is applied in the same pass (single-GPU step loop, dlas_cnn_fit_steps) and the batch-mean
// loss goes to *loss_out; otherwise the sums go to grad_sum for the caller's collective.
template <bool FUSE>
__global__ void __launch_bounds__(256)
cnn_reduce_kernel(const CnnDesc d, const float* __restrict__ partial, size_t pstride, int n_groups,
                  const float* __restrict__ lossrow, size_t n_rows, float* __restrict__ grad_sum, OptScalars opt,
                  float grad_scale, float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                  float* __restrict__ loss_out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < d.n_params) {
    float s = 0.f;
    for (int g = 0; g < n_groups; ++g) s += partial[(size_t)g * pstride + i];
    if (FUSE) {
      float p = params[i], mi = m[i], vi = (opt.kind == 0) ? 0.f : v[i];
      opt_update(opt, s * grad_scale, p, mi, vi);
      params[i] = p;
      m[i] = mi;
      if (opt.kind != 0) v[i] = vi;
    } else {
      grad_sum[i] = s;
    }
  }
  if (blockIdx.x == 0) {   // loss sum: 256 strided partial sums, then added in thread order
    __shared__ float red[256];
    float s = 0.f;
    for (size_t b = threadIdx.x; b < n_rows; b += 256) s += lossrow[b];
    red[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int k = 0; k < 256; ++k) t += red[k];
      if (FUSE) { if (loss_out) *loss_out = t * grad_scale; }
      else grad_sum[d.n_params] = t;
    }
  }
}

struct Ws { float *a, *out, *dout, *lossrow, *partial; int n_groups, rows_per_group; size_t pstride, bytes; };
Ws carve(const CnnDesc& d, size_t n_rows, void* base) {
  Ws w;
  size_t o = 0;
  auto take = [&](size_t n) { float* p = base ? (float*)base + o : nullptr; o += (n + 63) & ~(size_t)63; return p; };
  w.n_groups = (int)((n_rows + 63) / 64);
  if (w.n_groups > 64) w.n_groups = 64;
  if (w.n_groups < 1) w.n_groups = 1;
  w.rows_per_group = (int)((n_rows + w.n_groups - 1) / w.n_groups);
  w.pstride = ((size_t)d.n_params + 63) & ~(size_t)63;
  w.a = take(n_rows * d.TF);
  w.out = take(n_rows * d.O);
  w.dout = take(n_rows * d.O);
  w.lossrow = take(n_rows);
  w.partial = take((size_t)w.n_groups * w.pstride);
  w.bytes = o * sizeof(float);
  return w;
}

unsigned blocks_for(size_t n) {
  size_t need = (n + 255) / 256, cap = (size_t)sm_count() * 16;
  if (need < 1) need = 1;
  return (unsigned)(need < cap ? need : cap);
}

}  // namespace
}  // namespace dlas

using namespace dlas;

extern "C" size_t dlas_cnn_n_params(int L, int K, int F, int n_out) {
  CnnDesc d;
  return make_desc(L, K, F, n_out, d) == DLAS_OK ? (size_t)d.n_params : 0;
}

extern "C" size_t dlas_cnn_workspace_bytes(int L, int K, int F, int n_out, size_t n_rows) {
  CnnDesc d;
  if (make_desc(L, K, F, n_out, d) != DLAS_OK) return 0;
  return carve(d, n_rows, nullptr).bytes;
}

static int cnn_forward(const CnnDesc& d, int loss, const float* params, const float* X, const float* y, size_t n_rows, float keep,
                       uint64_t seed, uint64_t sample0, const Ws& w, cudaStream_t st) {
  conv_kernel<<<blocks_for(n_rows * d.TF), 256, 0, st>>>(d, params, X, n_rows, keep, stream_key(seed, STREAM_DROPOUT), sample0, w.a);
  DLAS_LAUNCH_CHECK();
  dense_kernel<<<(unsigned)(n_rows < 65535 ? n_rows : 65535), 256, 0, st>>>(d, loss, params, w.a, y, n_rows, w.out, w.dout, w.lossrow);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

namespace dlas { int make_opt_scalars(const dlas_opt_cfg* cfg, uint64_t iteration, OptScalars& o); }   // optim.cu

// forward + backward + (sums -> grad_sum | fused update when upd != nullptr)
struct CnnUpdate { OptScalars opt; float grad_scale; float *params, *m, *v, *loss_out; };
static int cnn_grad_impl(int L, int K, int F, int n_out, int loss, const float* params, const float* X, const float* y,
                         size_t n_rows, float keep, uint64_t dropout_seed, uint64_t sample0, float* grad_sum,
                         void* workspace, size_t workspace_bytes, void* stream, const CnnUpdate* upd) {
  CnnDesc d;
  int rc = make_desc(L, K, F, n_out, d);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(loss == DLAS_LOSS_MSE || loss == DLAS_LOSS_MAE_CLIP, "dlas_cnn_grad: unknown loss %d", loss);
  DLAS_REQUIRE(keep > 0.0f, "dlas_cnn_grad: keep probability %g must be > 0 (>= 1 disables dropout)", keep);
  DLAS_REQUIRE(params && (grad_sum || upd) && (n_rows == 0 || (X && y)), "dlas_cnn_grad: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows == 0 && upd) return DLAS_OK;
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)d.n_params + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  Ws w = carve(d, n_rows, workspace);
  if (!workspace || workspace_bytes < w.bytes) {
    set_error("dlas_cnn_grad: workspace of %zu bytes given, %zu needed", workspace_bytes, w.bytes);
    return DLAS_ERR_WORKSPACE;
  }
  rc = cnn_forward(d, loss, params, X, y, n_rows, keep, dropout_seed, sample0, w, st);
  if (rc != DLAS_OK) return rc;
  dim3 g1((d.TF * d.O + d.O + 255) / 256, w.n_groups);
  dwd_kernel<<<g1, 256, 0, st>>>(d, w.a, w.dout, n_rows, w.rows_per_group, w.partial, w.pstride);
  DLAS_LAUNCH_CHECK();
  dflat_kernel<<<blocks_for(n_rows * d.TF), 256, 0, st>>>(d, params, w.dout, n_rows, keep, w.a);
  DLAS_LAUNCH_CHECK();
  dim3 g2((d.K * d.F + d.F + 7) / 8, w.n_groups);   // a warp per entry
  dwc_kernel<<<g2, 256, 0, st>>>(d, w.a, X, n_rows, w.rows_per_group, w.partial, w.pstride);
  DLAS_LAUNCH_CHECK();
  if (upd)
    cnn_reduce_kernel<true><<<(d.n_params + 255) / 256, 256, 0, st>>>(d, w.partial, w.pstride, w.n_groups, w.lossrow, n_rows, nullptr,
                                                                      upd->opt, upd->grad_scale, upd->params, upd->m, upd->v,
                                                                      upd->loss_out);
  else
    cnn_reduce_kernel<false><<<(d.n_params + 255) / 256, 256, 0, st>>>(d, w.partial, w.pstride, w.n_groups, w.lossrow, n_rows, grad_sum,
                                                                       OptScalars{}, 0.f, nullptr, nullptr, nullptr, nullptr);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_cnn_grad(int L, int K, int F, int n_out, int loss, const float* params, const float* X, const float* y,
                             size_t n_rows, float keep, uint64_t dropout_seed, uint64_t sample0, float* grad_sum,
                             void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  return cnn_grad_impl(L, K, F, n_out, loss, params, X, y, n_rows, keep, dropout_seed, sample0, grad_sum, workspace,
                       workspace_bytes, stream, nullptr);
}

extern "C" int dlas_cnn_fit_steps(int L, int K, int F, int n_out, int loss, const dlas_opt_cfg* cfg, float* params, float* m,
                                  float* v, const float* X, const float* y, size_t n_rows, int batch,
                                  const int64_t* batch_order_host, int64_t first_batch, int n_steps, uint64_t iteration0,
                                  float keep, uint64_t dropout_seed, float* step_losses, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  DLAS_REQUIRE(cfg && params && m && X && y, "dlas_cnn_fit_steps: null pointer");
  DLAS_REQUIRE(batch >= 1 && n_steps >= 0, "dlas_cnn_fit_steps: batch %d, n_steps %d", batch, n_steps);
  const size_t n_batches = n_rows / (size_t)batch;
  DLAS_REQUIRE(n_batches >= 1 || n_steps == 0, "dlas_cnn_fit_steps: n_rows %zu < batch %d", n_rows, batch);
  for (int s = 0; s < n_steps; ++s) {
    const int64_t b = batch_order_host ? batch_order_host[s] : first_batch + s;
    DLAS_REQUIRE(b >= 0 && (size_t)b < n_batches, "dlas_cnn_fit_steps: batch index %lld out of range (%zu batches)", (long long)b,
                 n_batches);
    CnnUpdate u;
    int rc = make_opt_scalars(cfg, iteration0 + (uint64_t)s, u.opt);
    if (rc != DLAS_OK) return rc;
    DLAS_REQUIRE(u.opt.kind == DLAS_OPT_SGD || v != nullptr, "dlas_cnn_fit_steps: Adam needs the second-moment buffer v");
    u.grad_scale = 1.0f / (float)batch;
    u.params = params; u.m = m; u.v = v;
    u.loss_out = step_losses ? step_losses + s : nullptr;
    // the mask of step s is drawn for global samples [(iteration0 + s) * batch, ...): same bits as train_on_batch
    rc = cnn_grad_impl(L, K, F, n_out, loss, params, X + (size_t)b * batch * L, y + (size_t)b * batch * n_out, (size_t)batch, keep,
                       dropout_seed, (iteration0 + (uint64_t)s) * (uint64_t)batch, nullptr, workspace, workspace_bytes, stream, &u);
    if (rc != DLAS_OK) return rc;
  }
  return DLAS_OK;
}

extern "C" int dlas_cnn_predict(int L, int K, int F, int n_out, const float* params, const float* X, size_t n_rows, float* yhat,
                                void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  CnnDesc d;
  int rc = make_desc(L, K, F, n_out, d);
  if (rc != DLAS_OK) return rc;
  if (n_rows == 0) return DLAS_OK;
  DLAS_REQUIRE(params && X && yhat, "dlas_cnn_predict: null pointer");
  Ws w = carve(d, n_rows, workspace);
  if (!workspace || workspace_bytes < w.bytes) {
    set_error("dlas_cnn_predict: workspace of %zu bytes given, %zu needed", workspace_bytes, w.bytes);
    return DLAS_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  rc = cnn_forward(d, DLAS_LOSS_MSE, params, X, nullptr, n_rows, 1.0f, 0, 0, w, st);   // test time: no dropout
  if (rc != DLAS_OK) return rc;
  DLAS_CUDA(cudaMemcpyAsync(yhat, w.out, n_rows * d.O * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return DLAS_OK;
}
