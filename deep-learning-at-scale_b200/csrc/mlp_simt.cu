// Fused fp32 MLP training / inference on CUDA cores (DLAS_MODE_FP32): forward, loss, backward and
// per-CTA weight-gradient accumulation in ONE persistent kernel, any layer widths.
// Stands in for the per-batch body of Keras `model.fit` / `test_on_batch` (cnn.py:127-128, :132),
// with the loss of cnn.py:29-32 and MSE.
//
// Layout per CTA (one CTA per SM, grid = SM count, tiles of R=32 rows handed out round-robin so the
// summation order is fixed):
//   smem  actT[l][feature][RS]   activations of every layer, transposed (rows contiguous, RS=36
//                                floats so that float4 row loads stay aligned and column reads
//                                spread over banks)
//         dzT[2][max_w][RS]      ping-pong pre-activation gradients
//   HBM   partial[cta][n_params+1]  this CTA's running gradient sums (+ loss sum); private, no atomics
//         WT                     per-layer transposed weights for the backward data pass
// A second tiny kernel adds the per-CTA partials in CTA order (deterministic).
// Algorithmic work: 2*sum(in*out) flop forward and 4*sum(in*out) (minus the first layer's dX)
// backward per row; 4*(in0+out) bytes per row from HBM.
#include "common.cuh"
#include <mutex>
#include "simt_tiles.cuh"

namespace dlas {
namespace {

constexpr int R = 32;      // rows per tile
constexpr int RS = 36;     // padded row stride of transposed tiles (floats)
constexpr int NT = 512;    // threads per CTA
constexpr int MAXL = 8;    // max Dense layers

struct MlpDesc {
  int n_layers;
  int dims[MAXL + 1];
  int w_off[MAXL];
  int b_off[MAXL];
  int act_off[MAXL + 1];  // float offset of actT[l] in dynamic smem
  int dz_off[2];
  int n_params;
  int max_w;
  int smem_floats;
  int ldx;                // floats between consecutive rows of X (dims[0] for a dense matrix; the raw feature count for
                          // lag windows laid over a series: window i = dims[0] contiguous floats starting at row i)
};

int make_desc(const int* dims, int n_layers, MlpDesc& d) {
  DLAS_REQUIRE(dims != nullptr, "mlp: dims is null");
  DLAS_REQUIRE(n_layers >= 1 && n_layers <= MAXL, "mlp: n_layers %d not in 1..%d", n_layers, MAXL);
  d.n_layers = n_layers;
  int o = 0, so = 0, mw = 0;
  for (int l = 0; l <= n_layers; ++l) {
    DLAS_REQUIRE(dims[l] >= 1 && dims[l] <= 4096, "mlp: dims[%d]=%d not in 1..4096", l, dims[l]);
    d.dims[l] = dims[l];
    d.act_off[l] = so;
    so += dims[l] * RS;
    if (l > 0 && dims[l] > mw) mw = dims[l];
  }
  for (int l = 0; l < n_layers; ++l) {
    d.w_off[l] = o;
    o += dims[l] * dims[l + 1];
    d.b_off[l] = o;
    o += dims[l + 1];
  }
  d.n_params = o;
  d.ldx = dims[0];
  d.max_w = mw;
  d.dz_off[0] = so;
  so += mw * RS;
  d.dz_off[1] = so;
  so += mw * RS;
  d.smem_floats = so + 32;  // + block-reduce scratch
  if ((size_t)d.smem_floats * 4 > 227 * 1024) {
    set_error("mlp fp32 mode: layer widths need %zu bytes of shared memory per CTA (> 227 KB)",
              (size_t)d.smem_floats * 4);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  return DLAS_OK;
}

__host__ __device__ inline size_t partial_stride(int n_params) { return ((size_t)n_params + 1 + 3) & ~(size_t)3; }

__device__ __forceinline__ float block_sum(float v, float* scratch) {
  // fixed-order tree: warp shuffle, then lane 0 of each warp through smem
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.0f;
  if (threadIdx.x == 0)
    for (int w = 0; w < NT / 32; ++w) s += scratch[w];
  return s;  // valid on thread 0
}

enum { KIND_TRAIN = 0, KIND_EVAL = 1, KIND_PREDICT = 2 };

template <int KIND>
__global__ void __launch_bounds__(NT, 1)
mlp_simt_kernel(const MlpDesc d, const int loss, const float* __restrict__ params,
                const float* __restrict__ WT, const float* __restrict__ X_base, const float* __restrict__ y_base,
                const size_t n_rows, float* __restrict__ partials, float* __restrict__ yhat,
                const long long* __restrict__ cursor) {
  // device-resident batch cursor (replayable steps): rows [cursor*n_rows, (cursor+1)*n_rows)
  const size_t batch0 = cursor ? (size_t)cursor[0] * n_rows : 0;
  const float* __restrict__ X = X_base + batch0 * d.ldx;
  const float* __restrict__ y = y_base ? y_base + batch0 * d.dims[d.n_layers] : nullptr;
  extern __shared__ __align__(16) float smem[];
  const int L = d.n_layers;
  const int d0 = d.dims[0], dL = d.dims[L];
  const size_t n_tiles = (n_rows + R - 1) / R;
  float* my_partial = partials ? partials + (size_t)blockIdx.x * partial_stride(d.n_params) : nullptr;
  float loss_acc = 0.0f;
  bool first = true;

  for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const size_t row0 = tile * R;
    const int nr = (int)((n_rows - row0) < (size_t)R ? (n_rows - row0) : (size_t)R);
    // ---- X tile -> actT[0] (coalesced global reads, transposed smem writes)
    {
      float* a0 = smem + d.act_off[0];
      const float* xg = X + row0 * (size_t)d.ldx;
      if (d.ldx == d0) {
        for (int idx = threadIdx.x; idx < R * d0; idx += NT) {
          int r = idx / d0, k = idx - r * d0;
          a0[k * RS + r] = (r < nr) ? __ldcs(xg + idx) : 0.0f;
        }
      } else {   // overlapping windows: neighbouring rows share floats, keep them in cache (no streaming hint)
        for (int idx = threadIdx.x; idx < R * d0; idx += NT) {
          int r = idx / d0, k = idx - r * d0;
          a0[k * RS + r] = (r < nr) ? __ldg(xg + (size_t)r * d.ldx + k) : 0.0f;
        }
      }
    }
    __syncthreads();
    // ---- forward
    for (int l = 0; l < L; ++l) {
      const float* inT = smem + d.act_off[l];
      float* outT = smem + d.act_off[l + 1];
      const float* bias = params + d.b_off[l];
      const bool relu = (l < L - 1);
      gemm_dispatch<R, RS, NT>(inT, params + d.w_off[l], d.dims[l + 1], d.dims[l], d.dims[l + 1],
                    [&](int j, int r0, auto& acc) {
                      const float b = __ldg(bias + j);
                      constexpr int RT = sizeof(acc) / sizeof(float);
#pragma unroll
                      for (int i = 0; i < RT; ++i) {
                        float v = acc[i] + b;
                        outT[j * RS + r0 + i] = relu ? fmaxf(v, 0.0f) : v;
                      }
                    });
      __syncthreads();
    }
    const float* outT = smem + d.act_off[L];
    if (KIND == KIND_PREDICT) {
      float* yo = yhat + row0 * (size_t)dL;
      for (int idx = threadIdx.x; idx < nr * dL; idx += NT) {
        int r = idx / dL, j = idx - r * dL;
        yo[idx] = outT[j * RS + r];
      }
      __syncthreads();
      continue;
    }
    // ---- loss and d(loss sum)/d(yhat) -> dzT[cur]
    int cur = 0;
    float* dz = smem + d.dz_off[cur];
    {
      const float inv_out = 1.0f / (float)dL;
      const float* yg = y + row0 * (size_t)dL;
      for (int idx = threadIdx.x; idx < R * dL; idx += NT) {
        int r = idx / dL, j = idx - r * dL;
        float g = 0.0f;
        if (r < nr) {
          float dd = outT[j * RS + r] - __ldcs(yg + idx);
          if (loss == DLAS_LOSS_MSE) {
            loss_acc += dd * dd * inv_out;
            g = 2.0f * dd * inv_out;
          } else {  // cnn.py:29-32 mae_clip, CLIP_VALUE = 6
            float a = fabsf(dd);
            loss_acc += fminf(a, 6.0f) * inv_out;
            g = (a <= 6.0f) ? (dd > 0.0f ? inv_out : (dd < 0.0f ? -inv_out : 0.0f)) : 0.0f;
          }
        }
        if (KIND == KIND_TRAIN) dz[j * RS + r] = g;
      }
    }
    if (KIND == KIND_TRAIN) {
      __syncthreads();
      // ---- backward
      for (int l = L - 1; l >= 0; --l) {
        const float* aT = smem + d.act_off[l];
        const float* dzc = smem + d.dz_off[cur];
        accumulate_dw<R, RS, NT>(aT, dzc, d.dims[l], d.dims[l + 1], my_partial + d.w_off[l], my_partial + d.b_off[l], first);
        if (l > 0) {
          float* dzn = smem + d.dz_off[cur ^ 1];
          // dA[k][r] = sum_j dz[j][r] * W[k][j] = sum_j dzT[j][r] * WT[j][k]; masked by relu'(A)
          gemm_dispatch<R, RS, NT>(dzc, WT + d.w_off[l], d.dims[l], d.dims[l + 1], d.dims[l],
                        [&](int k, int r0, auto& acc) {
                          constexpr int RT = sizeof(acc) / sizeof(float);
#pragma unroll
                          for (int i = 0; i < RT; ++i)
                            dzn[k * RS + r0 + i] = aT[k * RS + r0 + i] > 0.0f ? acc[i] : 0.0f;
                        });
        }
        __syncthreads();
        cur ^= 1;
      }
      first = false;
    } else {
      __syncthreads();
    }
  }
  if (KIND != KIND_PREDICT) {
    float s = block_sum(loss_acc, smem + d.smem_floats - 32);
    if (threadIdx.x == 0) my_partial[d.n_params] = s;
  }
}

__global__ void __launch_bounds__(256)
transpose_weights_kernel(const MlpDesc d, const float* __restrict__ params, float* __restrict__ WT) {
  for (int l = 1; l < d.n_layers; ++l) {  // layer 0 never needs dX
    const int K = d.dims[l], N = d.dims[l + 1];
    const float* W = params + d.w_off[l];
    float* T = WT + d.w_off[l];
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < K * N; idx += gridDim.x * 256) {
      int j = idx / K, k = idx - j * K;  // T[j][k] = W[k][j]
      T[idx] = W[(size_t)k * N + j];
    }
  }
}

// out[i] = sum over active CTAs (in CTA order) of partial[c][i], i in [i0, i0+count)
__global__ void __launch_bounds__(256)
reduce_partials_kernel(const float* __restrict__ partials, size_t stride, int n_active, size_t i0,
                       size_t count, float* __restrict__ out, const long long* __restrict__ state = nullptr,
                       size_t parity_stride = 0) {
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= count) return;
  float s = 0.0f;
  for (int c = 0; c < n_active; ++c) s += partials[(size_t)c * stride + i0 + i];
  out[(state && (state[1] & 1)) ? parity_stride + i : i] = s;   // state[1] = iteration: double-buffered bucket
}

int grid_size() { return sm_count(); }

template <int KIND>
int launch_main(const MlpDesc& d, int loss, const float* params, const float* WT, const float* X,
                const float* y, size_t n_rows, float* partials, float* yhat, cudaStream_t st,
                const long long* cursor = nullptr) {
  static bool attr_set[64] = {};   // per kernel instantiation AND per device: the attribute belongs to the device
  size_t smem = (size_t)d.smem_floats * sizeof(float);
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(mlp_simt_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  mlp_simt_kernel<KIND><<<grid_size(), NT, smem, st>>>(d, loss, params, WT, X, y, n_rows, partials, yhat, cursor);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

}  // namespace

// ---- entry points used by mlp_api.cu ---------------------------------------------------------
size_t simt_workspace_bytes(const int* dims, int n_layers) {
  MlpDesc d;
  if (make_desc(dims, n_layers, d) != DLAS_OK) return 0;
  return ((size_t)grid_size() * partial_stride(d.n_params) + (size_t)d.n_params) * sizeof(float);
}

int simt_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y,
              size_t n_rows, float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st,
              const long long* cursor, size_t parity_stride, size_t ldx) {
  MlpDesc d;
  int rc = make_desc(dims, n_layers, d);
  if (rc != DLAS_OK) return rc;
  if (ldx) d.ldx = (int)ldx;
  size_t need = simt_workspace_bytes(dims, n_layers);
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("mlp: workspace of %zu bytes given, %zu needed", workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)d.n_params + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  float* partials = (float*)workspace;
  float* WT = partials + (size_t)grid_size() * partial_stride(d.n_params);
  transpose_weights_kernel<<<64, 256, 0, st>>>(d, params, WT);
  DLAS_LAUNCH_CHECK();
  int pi = prof_begin(DLAS_PROF_MLP_GRAD, st);
  rc = launch_main<KIND_TRAIN>(d, loss, params, WT, X, y, n_rows, partials, nullptr, st, cursor);
  prof_end(pi, st);
  if (rc != DLAS_OK) return rc;
  size_t n_tiles = (n_rows + R - 1) / R;
  int n_active = (int)(n_tiles < (size_t)grid_size() ? n_tiles : (size_t)grid_size());
  size_t count = (size_t)d.n_params + 1;
  reduce_partials_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(partials, partial_stride(d.n_params),
                                                                        n_active, 0, count, grad_sum,
                                                                        parity_stride ? cursor : nullptr, parity_stride);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

int simt_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X,
                   const float* y, size_t n_rows, float* loss_sum, void* workspace, size_t workspace_bytes,
                   cudaStream_t st, size_t ldx) {
  MlpDesc d;
  int rc = make_desc(dims, n_layers, d);
  if (rc != DLAS_OK) return rc;
  if (ldx) d.ldx = (int)ldx;
  size_t need = simt_workspace_bytes(dims, n_layers);
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("mlp: workspace of %zu bytes given, %zu needed", workspace_bytes, need);
    return DLAS_ERR_WORKSPACE;
  }
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
    return DLAS_OK;
  }
  float* partials = (float*)workspace;
  rc = launch_main<KIND_EVAL>(d, loss, params, nullptr, X, y, n_rows, partials, nullptr, st);
  if (rc != DLAS_OK) return rc;
  // CTAs without a tile still write their (zero) loss partial, so all CTAs are summed
  reduce_partials_kernel<<<1, 256, 0, st>>>(partials, partial_stride(d.n_params), grid_size(),
                                            (size_t)d.n_params, 1, loss_sum);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

int simt_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows,
                 float* yhat, cudaStream_t st, size_t ldx) {
  MlpDesc d;
  int rc = make_desc(dims, n_layers, d);
  if (rc != DLAS_OK) return rc;
  if (ldx) d.ldx = (int)ldx;
  if (n_rows == 0) return DLAS_OK;
  return launch_main<KIND_PREDICT>(d, DLAS_LOSS_MSE, params, nullptr, X, nullptr, n_rows, nullptr, yhat, st);
}

}  // namespace dlas

// ---------------------------------------------------------------------------------------------
// dlas_mlp_grad_noalloc: the same forward + loss + backward as dlas_mlp_grad (fp32 mode), for
// hosts that cannot give the library a scratch buffer because they are not allowed to call
// cudaMalloc where they train -- for example an executor plug-in whose device memory belongs
// entirely to the executor's own pool and is fixed once tasks start, or a latency-bound serving
// loop that bans allocator calls because cudaMalloc stalls the whole device.  The per-CTA gradient
// partials and the transposed weights live in the file-scope device array below instead.
//
// * The array is DLAS_NOALLOC_SCRATCH_BYTES (4 MiB) and every GPU that loads the library keeps it
//   until the process exits.  Networks whose scratch need is larger (roughly: more than ~7,000
//   parameters on a 148-SM part) are refused with DLAS_ERR_WORKSPACE -- use dlas_mlp_grad.
// * In host code the array's name is only a stand-in, not a device address.  The first call on each
//   device (with that device current) asks the runtime for the real address with
//   cudaGetSymbolAddress and caches it; that first call should therefore be made before any stream
//   capture begins.  Because CUDA loads modules lazily, that first call is also when the driver
//   places this array and the kernels on the GPU: make it (n_rows = 0 suffices) before the host's
//   pool has taken all device memory, or set CUDA_MODULE_LOADING=EAGER.  Later calls only launch
//   kernels and a memset and leave free device memory unchanged.
// * All calls on a device share the one array and it keeps whatever the previous call left.  Each
//   call clears the part it is about to use, and every CTA overwrites (not adds to) its slots on
//   its first tile, so stale numbers are never read.  Two calls that overlap on the GPU (different
//   streams, or different host threads) would write the same scratch and race: issue them on one
//   stream, or serialise them.
// * Nothing here calls cudaMalloc, cudaFree, or any routine that allocates (no events, no torch).
// ---------------------------------------------------------------------------------------------
namespace dlas {
namespace {
constexpr size_t kNoAllocFloats = DLAS_NOALLOC_SCRATCH_BYTES / sizeof(float);
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];
float* g_noalloc_ptr[64];  // per device, filled on first use
std::mutex g_noalloc_mu;   // two host threads making the very first call on one device
}  // namespace
}  // namespace dlas

extern "C" int dlas_mlp_grad_noalloc(const int* dims, int n_layers, int loss, const float* params,
                                     const float* X, const float* y, size_t n_rows, float* grad_sum,
                                     void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  MlpDesc d;
  int rc = make_desc(dims, n_layers, d);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(loss == DLAS_LOSS_MSE || loss == DLAS_LOSS_MAE_CLIP, "dlas_mlp_grad_noalloc: unknown loss %d", loss);
  DLAS_REQUIRE(params && grad_sum && (n_rows == 0 || (X && y)), "dlas_mlp_grad_noalloc: null pointer");
  const int grid = grid_size();
  const size_t stride = partial_stride(d.n_params);
  const size_t need_floats = (size_t)grid * stride + (size_t)d.n_params;
  if (need_floats > kNoAllocFloats) {
    set_error("dlas_mlp_grad_noalloc: these layer widths need %zu bytes of scratch but the built-in array holds "
              "%zu (DLAS_NOALLOC_SCRATCH_BYTES); use dlas_mlp_grad with a caller-provided workspace",
              need_floats * sizeof(float), (size_t)DLAS_NOALLOC_SCRATCH_BYTES);
    return DLAS_ERR_WORKSPACE;
  }
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  DLAS_REQUIRE(dev >= 0 && dev < 64, "dlas_mlp_grad_noalloc: device ordinal %d out of range", dev);
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {  // once per device, with that device current
      float* p = nullptr;
      DLAS_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows == 0) {
    DLAS_CUDA(cudaMemsetAsync(grad_sum, 0, ((size_t)d.n_params + 1) * sizeof(float), st));
    return DLAS_OK;
  }
  DLAS_CUDA(cudaMemsetAsync(scratch, 0, need_floats * sizeof(float), st));
  float* partials = scratch;
  float* WT = scratch + (size_t)grid * stride;
  transpose_weights_kernel<<<64, 256, 0, st>>>(d, params, WT);
  DLAS_LAUNCH_CHECK();
  rc = launch_main<KIND_TRAIN>(d, loss, params, WT, X, y, n_rows, partials, nullptr, st);
  if (rc != DLAS_OK) return rc;
  const size_t n_tiles = (n_rows + R - 1) / R;
  const int n_active = (int)(n_tiles < (size_t)grid ? n_tiles : (size_t)grid);
  const size_t count = (size_t)d.n_params + 1;
  reduce_partials_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(partials, stride, n_active, 0, count, grad_sum);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}
