// Optimiser update kernels: Keras-1 SGD with momentum / Nesterov / time decay (cnn.py:117) and
// Keras Adam (epsilon outside the bias-corrected sqrt).  One fused elementwise pass over the flat
// parameter vector; the mean over the (global) batch is folded in as grad_scale, so that the
// all-reduced gradient SUM is consumed as is.  Step-dependent scalars are computed on the host
// in double and passed by value.
#include "common.cuh"
#include "optim_math.cuh"
#include <math.h>

namespace dlas {
namespace {

__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, float* __restrict__ vel, const float* __restrict__ gsum, size_t n,
           float grad_scale, float lr_t, float momentum, int nesterov, float* __restrict__ loss_out) {
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i == 0 && loss_out) *loss_out = gsum[n] * grad_scale;
  if (i >= n) return;
  float g = gsum[i] * grad_scale;
  float v = momentum * vel[i] - lr_t * g;  // v <- m*v - lr_t*g
  vel[i] = v;
  p[i] += nesterov ? (momentum * v - lr_t * g) : v;
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
            const float* __restrict__ gsum, size_t n, float grad_scale, float lr_t, float b1, float b2,
            float eps, float* __restrict__ loss_out) {
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i == 0 && loss_out) *loss_out = gsum[n] * grad_scale;
  if (i >= n) return;
  float g = gsum[i] * grad_scale;
  float mi = b1 * m[i] + (1.0f - b1) * g;
  float vi = b2 * v[i] + (1.0f - b2) * g * g;
  m[i] = mi;
  v[i] = vi;
  p[i] -= lr_t * mi / (sqrtf(vi) + eps);
}

}  // namespace

int opt_step_impl(const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* grad_sum, size_t n,
                  float grad_scale, uint64_t iteration, float* loss_out, cudaStream_t st) {
  DLAS_REQUIRE(cfg != nullptr, "dlas_opt_step: cfg is null");
  DLAS_REQUIRE(params && m && grad_sum, "dlas_opt_step: null pointer");
  if (n == 0) return DLAS_OK;
  const double lr_t = (double)cfg->lr / (1.0 + (double)cfg->decay * (double)iteration);
  unsigned grid = (unsigned)((n + 255) / 256);
  if (cfg->kind == DLAS_OPT_SGD) {
    sgd_kernel<<<grid, 256, 0, st>>>(params, m, grad_sum, n, grad_scale, (float)lr_t, cfg->momentum,
                                     cfg->nesterov, loss_out);
  } else if (cfg->kind == DLAS_OPT_ADAM) {
    DLAS_REQUIRE(v != nullptr, "dlas_opt_step: Adam needs the second-moment buffer v");
    const double t = (double)iteration + 1.0;
    const double lr_a = lr_t * sqrt(1.0 - pow((double)cfg->beta2, t)) / (1.0 - pow((double)cfg->beta1, t));
    adam_kernel<<<grid, 256, 0, st>>>(params, m, v, grad_sum, n, grad_scale, (float)lr_a, cfg->beta1,
                                      cfg->beta2, cfg->eps, loss_out);
  } else {
    set_error("dlas_opt_step: unknown optimiser kind %d", cfg->kind);
    return DLAS_ERR_INVALID;
  }
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

namespace {
// Optimiser update whose step-dependent scalars come from the DEVICE-resident iteration counter, so that a
// captured (CUDA-graph) step can be replayed: lr_t = lr/(1+decay*it), Adam bias correction from t = it+1.
__global__ void __launch_bounds__(256)
opt_at_kernel(const dlas_opt_cfg cfg, float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
              const float* __restrict__ gsum, size_t n, float grad_scale, const dlas_step_state* __restrict__ st,
              float* __restrict__ loss_ring, long long ring_len) {
  __shared__ OptScalars o;
  const long long it = st->iteration;
  if (threadIdx.x == 0) {
    double lr_t = (double)cfg.lr / (1.0 + (double)cfg.decay * (double)it);
    if (cfg.kind == DLAS_OPT_ADAM) {
      const double t = (double)it + 1.0;
      lr_t = lr_t * sqrt(1.0 - pow((double)cfg.beta2, t)) / (1.0 - pow((double)cfg.beta1, t));
    }
    o.kind = cfg.kind; o.lr_t = (float)lr_t; o.momentum = cfg.momentum; o.nesterov = cfg.nesterov;
    o.beta1 = cfg.beta1; o.beta2 = cfg.beta2; o.eps = cfg.eps;
  }
  __syncthreads();
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i == 0 && loss_ring) loss_ring[it % ring_len] = gsum[n] * grad_scale;
  if (i >= n) return;
  float pi = p[i], mi = m[i], vi = (o.kind == 0) ? 0.f : v[i];
  opt_update(o, gsum[i] * grad_scale, pi, mi, vi);
  p[i] = pi;
  m[i] = mi;
  if (o.kind != 0) v[i] = vi;
}

__global__ void step_advance_kernel(dlas_step_state* st, long long n_batches) {
  st->cursor = (st->cursor + 1) % n_batches;
  st->iteration += 1;
}
}  // namespace

// Step-dependent scalars for kernels that fuse the update into their own epilogue (optim_math.cuh).
int make_opt_scalars(const dlas_opt_cfg* cfg, uint64_t iteration, OptScalars& o) {
  DLAS_REQUIRE(cfg != nullptr, "optimiser cfg is null");
  DLAS_REQUIRE(cfg->kind == DLAS_OPT_SGD || cfg->kind == DLAS_OPT_ADAM, "unknown optimiser kind %d", cfg->kind);
  double lr_t = (double)cfg->lr / (1.0 + (double)cfg->decay * (double)iteration);
  if (cfg->kind == DLAS_OPT_ADAM) {
    const double t = (double)iteration + 1.0;
    lr_t = lr_t * sqrt(1.0 - pow((double)cfg->beta2, t)) / (1.0 - pow((double)cfg->beta1, t));
  }
  o.kind = cfg->kind;
  o.lr_t = (float)lr_t;
  o.momentum = cfg->momentum;
  o.nesterov = cfg->nesterov;
  o.beta1 = cfg->beta1;
  o.beta2 = cfg->beta2;
  o.eps = cfg->eps;
  return DLAS_OK;
}

}  // namespace dlas

extern "C" int dlas_opt_step(const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                             const float* grad_sum, size_t n, float grad_scale, uint64_t iteration,
                             float* loss_out, void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  return opt_step_impl(cfg, params, m, v, grad_sum, n, grad_scale, iteration, loss_out, (cudaStream_t)stream);
}

extern "C" int dlas_opt_step_at(const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* grad_sum,
                                size_t n, float grad_scale, const dlas_step_state* state_dev, float* loss_ring,
                                int64_t ring_len, void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  DLAS_REQUIRE(cfg && params && m && grad_sum && state_dev, "dlas_opt_step_at: null pointer");
  DLAS_REQUIRE(cfg->kind == DLAS_OPT_SGD || (cfg->kind == DLAS_OPT_ADAM && v), "dlas_opt_step_at: bad optimiser / missing v");
  DLAS_REQUIRE(loss_ring == nullptr || ring_len >= 1, "dlas_opt_step_at: ring_len < 1");
  if (n == 0) return DLAS_OK;
  opt_at_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*cfg, params, m, v, grad_sum, n, grad_scale,
                                                                            state_dev, loss_ring, (long long)ring_len);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}

extern "C" int dlas_step_advance(dlas_step_state* state_dev, int64_t n_batches, void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  DLAS_REQUIRE(state_dev && n_batches >= 1, "dlas_step_advance: bad arguments");
  step_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state_dev, (long long)n_batches);
  DLAS_LAUNCH_CHECK();
  return DLAS_OK;
}
