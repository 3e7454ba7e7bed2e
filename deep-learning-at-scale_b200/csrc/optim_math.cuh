// Per-element optimiser updates shared by optim.cu and the fused reduce+update kernels.
// Keras-1 SGD (cnn.py:117): v <- m v - lr_t g; Nesterov p += m v - lr_t g, else p += v.
// Keras Adam: epsilon outside the bias-corrected square root (lr_t carries the bias correction).
#pragma once

namespace dlas {

struct OptScalars {   // step-dependent scalars computed on the host in double
  int kind;           // DLAS_OPT_*
  float lr_t;         // SGD: lr/(1+decay*it); Adam: that times sqrt(1-b2^t)/(1-b1^t)
  float momentum;
  int nesterov;
  float beta1, beta2, eps;
};

__device__ __forceinline__ void opt_update(const OptScalars& o, float g, float& p, float& m, float& v) {
  if (o.kind == 0) {
    const float vel = o.momentum * m - o.lr_t * g;
    m = vel;
    p += o.nesterov ? (o.momentum * vel - o.lr_t * g) : vel;
  } else {
    const float mi = o.beta1 * m + (1.0f - o.beta1) * g;
    const float vi = o.beta2 * v + (1.0f - o.beta2) * g * g;
    m = mi;
    v = vi;
    p -= o.lr_t * mi / (sqrtf(vi) + o.eps);
  }
}

}  // namespace dlas
