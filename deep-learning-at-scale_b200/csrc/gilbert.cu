// Gilbert's choke equation  q = P * S^1.89 / (435 * R^0.546)   (BASELINE.json north_star;
// the reference names the model at Readme.md:7-8 and ships no code for it).
//
// One streaming pass, 16 algorithmic bytes per row (three fp32 reads, one fp32 write), so the
// kernel is bounded by HBM.  The arithmetic is arranged so that it stays under the issue budget
// that bound leaves (~100 instructions per row at 3.5e11 rows/s) while keeping <=1e-6 relative
// error against the fp64 formula:
//
//   q = 2^t,  t = log2 P + 1.89 log2 S - 0.546 log2 R - log2 435
//
// * every input is split exactly into exponent e and mantissa m in [sqrt(1/2), sqrt(2));
// * the exponent part  eP + 1.89 eS - 0.546 eR - log2 435  is carried as th + tl with th an exact
//   multiple of 1/64 (constants split hi+lo, hi = k/64) -- a plain fp32 sum would lose ~2e-6
//   at |t| ~ 20;
// * log2 m = f (c0 + f Q(f)), f = m-1, Q of degree 7 (abs error 3.8e-8), or MUFU.LG2 on m;
// * 2^r, |r| <= 1/2, degree-6 polynomial; the integer part goes straight into the exponent field.
//
// tools/gilbert_poly_fit.py fits the polynomials and emulates this exact fp32 sequence on the CPU
// (max relative error 2.9e-7 over 32M samples, including 2^+-30 dynamic range).
#include "common.cuh"

namespace dlas {
namespace {

constexpr float kC0 = 1.4426950408889634f;  // 1/ln 2 (fp32-rounded; Q is fitted against this value)
constexpr float kQ0 = -0.7213473916053772f, kQ1 = 0.4809028208255768f, kQ2 = -0.3607034683227539f,
                kQ3 = 0.2881705164909363f, kQ4 = -0.2389395534992218f, kQ5 = 0.21315859258174896f,
                kQ6 = -0.20705921947956085f, kQ7 = 0.13311053812503815f;
constexpr float kE0 = 1.0f, kE1 = 0.6931471824645996f, kE2 = 0.24022652208805084f,
                kE3 = 0.05550327152013779f, kE4 = 0.009617994539439678f,
                kE5 = 0.001340043148957193f, kE6 = 0.00015478018030989915f;
constexpr float kA = 1.89f, kB = 0.546f;
constexpr float kAh = 121.0f / 64.0f, kBh = 35.0f / 64.0f, kLh = 561.0f / 64.0f;
constexpr float kAl = (float)(1.89 - 121.0 / 64.0);
constexpr float kBl = (float)(0.546 - 35.0 / 64.0);
constexpr float kLl = (float)(8.764871590736092 - 561.0 / 64.0);  // log2(435) - 561/64

// x = 2^e * (1+f), 1+f in [sqrt(1/2), sqrt(2)); e returned as float.  Valid for positive normals.
template <bool MUFU>
__device__ __forceinline__ void split_log2(float x, float& e, float& L) {
  int i = __float_as_int(x);
  int ei = (i - 0x3f3504f3) >> 23;
  float m = __int_as_float(i - (ei << 23));
  e = __int_as_float(ei + 0x4B400000) - 12582912.0f;  // exact int->float for |ei| < 2^22
  if (MUFU) {
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(L) : "f"(m));
  } else {
    float f = m - 1.0f;
    float q = kQ7;
    q = __fmaf_rn(q, f, kQ6);
    q = __fmaf_rn(q, f, kQ5);
    q = __fmaf_rn(q, f, kQ4);
    q = __fmaf_rn(q, f, kQ3);
    q = __fmaf_rn(q, f, kQ2);
    q = __fmaf_rn(q, f, kQ1);
    q = __fmaf_rn(q, f, kQ0);
    float u = __fmaf_rn(f, q, kC0);
    L = __fmul_rn(f, u);
  }
}

__device__ __forceinline__ bool pos_normal(float x) {
  return (uint32_t)(__float_as_int(x) - 0x00800000) < (uint32_t)(0x7f800000 - 0x00800000);
}

// Any input: IEEE powf evaluation of the formula (rare rows only).
__device__ __noinline__ float gilbert_slow(float P, float S, float R) {
  if (P == 0.0f && S > 0.0f && R > 0.0f && S < INFINITY) return 0.0f;
  return P * powf(S, kA) / (435.0f * powf(R, kB));
}

// Fast path, valid when P,S,R are positive normal floats and |n| <= 125 (checked by the caller,
// once per thread, over all of its rows).  Returns q; n_out = the integer part of t.
template <int MODE>
__device__ __forceinline__ float gilbert_fast(float P, float S, float R, float& n_out) {
  float eP, eS, eR, LP, LS, LR;
  split_log2<(MODE & 1) != 0>(P, eP, LP);
  split_log2<(MODE & 2) != 0>(S, eS, LS);
  split_log2<(MODE & 4) != 0>(R, eR, LR);
  float th = __fmaf_rn(-kBh, eR, __fmaf_rn(kAh, eS, __fadd_rn(eP, -kLh)));  // exact, k/64
  float tl = __fmaf_rn(-kBl, eR, __fmaf_rn(kAl, eS, -kLl));
  float tf = __fmaf_rn(-kB, LR, __fmaf_rn(kA, LS, LP));
  float ts = __fadd_rn(tl, tf);
  float nf = __fadd_rn(__fadd_rn(th, ts), 12582912.0f);  // round-to-nearest-integer by magic add
  float n = __fadd_rn(nf, -12582912.0f);
  float r = __fadd_rn(__fadd_rn(th, -n), ts);
  float p = kE6;
  p = __fmaf_rn(p, r, kE5);
  p = __fmaf_rn(p, r, kE4);
  p = __fmaf_rn(p, r, kE3);
  p = __fmaf_rn(p, r, kE2);
  p = __fmaf_rn(p, r, kE1);
  p = __fmaf_rn(p, r, kE0);
  n_out = n;
  // low bits of nf hold n (0x4B400000 << 23 == 0): add n to the exponent field of p
  return __int_as_float(__float_as_int(p) + (__float_as_int(nf) << 23));
}

template <int MODE>
__device__ __forceinline__ float gilbert_one(float P, float S, float R) {
  float n;
  float q = gilbert_fast<MODE>(P, S, R, n);
  bool ok = pos_normal(P) & pos_normal(S) & pos_normal(R) & (fabsf(n) <= 125.0f);
  return ok ? q : gilbert_slow(P, S, R);
}

// Cold path for one row that fell outside the fast path; arguments by value so that the hot loop
// keeps everything in registers (a by-reference float4 forces a local-memory round trip).
__device__ __noinline__ float patch1(float p, float s, float r, float n, float q) {
  bool ok = pos_normal(p) & pos_normal(s) & pos_normal(r) & (fabsf(n) <= 125.0f);
  return ok ? q : gilbert_slow(p, s, r);
}

__device__ __forceinline__ uint32_t umin3(uint32_t a, uint32_t b, uint32_t c) { return min(min(a, b), c); }
__device__ __forceinline__ uint32_t umax3(uint32_t a, uint32_t b, uint32_t c) { return max(max(a, b), c); }

constexpr int kThreads = 256;
constexpr int kUnroll = 2;  // float4 groups per array per thread

// One-shot grid: every thread owns kUnroll float4 groups (8 rows), all six 128-bit loads issued
// before the arithmetic.  Measured on B200 (tools/gilbert_tune.cu, 200M rows): one-shot 7.0 TB/s
// against 6.6 TB/s for a grid-stride loop over 16 CTAs/SM; default / streaming / no-allocate load
// policies and 256-bit loads are within 1% of each other.
template <int MODE>
__global__ void __launch_bounds__(kThreads)
gilbert_vec_kernel(const float4* __restrict__ P, const float4* __restrict__ S,
                   const float4* __restrict__ R, float4* __restrict__ q, size_t n4) {
  const size_t base = (size_t)blockIdx.x * kThreads * kUnroll + threadIdx.x;
  float4 p[kUnroll], s[kUnroll], r[kUnroll];
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    size_t i = base + (size_t)u * kThreads;
    if (i < n4) {
      p[u] = __ldg(P + i);
      s[u] = __ldg(S + i);
      r[u] = __ldg(R + i);
    } else {
      p[u] = s[u] = r[u] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    }
  }
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    size_t i = base + (size_t)u * kThreads;
    float4 o;
    float n0, n1, n2, n3;
    o.x = gilbert_fast<MODE>(p[u].x, s[u].x, r[u].x, n0);
    o.y = gilbert_fast<MODE>(p[u].y, s[u].y, r[u].y, n1);
    o.z = gilbert_fast<MODE>(p[u].z, s[u].z, r[u].z, n2);
    o.w = gilbert_fast<MODE>(p[u].w, s[u].w, r[u].w, n3);
    // one validity test per float4 group: all 12 inputs positive normal, all |n| <= 125
    uint32_t lo = umin3(umin3(__float_as_uint(p[u].x), __float_as_uint(p[u].y), __float_as_uint(p[u].z)),
                        umin3(__float_as_uint(p[u].w), __float_as_uint(s[u].x), __float_as_uint(s[u].y)),
                        umin3(umin3(__float_as_uint(s[u].z), __float_as_uint(s[u].w), __float_as_uint(r[u].x)),
                              umin3(__float_as_uint(r[u].y), __float_as_uint(r[u].z), __float_as_uint(r[u].w)),
                              0xffffffffu));
    uint32_t hi = umax3(umax3(__float_as_uint(p[u].x), __float_as_uint(p[u].y), __float_as_uint(p[u].z)),
                        umax3(__float_as_uint(p[u].w), __float_as_uint(s[u].x), __float_as_uint(s[u].y)),
                        umax3(umax3(__float_as_uint(s[u].z), __float_as_uint(s[u].w), __float_as_uint(r[u].x)),
                              umax3(__float_as_uint(r[u].y), __float_as_uint(r[u].z), __float_as_uint(r[u].w)), 0u));
    float nm = fmaxf(fmaxf(fabsf(n0), fabsf(n1)), fmaxf(fabsf(n2), fabsf(n3)));
    if (!((lo >= 0x00800000u) & (hi <= 0x7f7fffffu) & (nm <= 125.0f))) {
      o.x = patch1(p[u].x, s[u].x, r[u].x, n0, o.x);
      o.y = patch1(p[u].y, s[u].y, r[u].y, n1, o.y);
      o.z = patch1(p[u].z, s[u].z, r[u].z, n2, o.z);
      o.w = patch1(p[u].w, s[u].w, r[u].w, n3, o.w);
    }
    if (i < n4) q[i] = o;
  }
}

template <int MODE>
__global__ void __launch_bounds__(kThreads)
gilbert_scalar_kernel(const float* __restrict__ P, const float* __restrict__ S,
                      const float* __restrict__ R, float* __restrict__ q, size_t n) {
  const size_t stride = (size_t)gridDim.x * kThreads;
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride)
    q[i] = gilbert_one<MODE>(P[i], S[i], R[i]);
}

template <int MODE>
int launch(const float* P, const float* S, const float* R, float* q, size_t n, cudaStream_t st) {
  const int sms = sm_count();
  const bool aligned = ((((uintptr_t)P) | ((uintptr_t)S) | ((uintptr_t)R) | ((uintptr_t)q)) & 15) == 0;
  size_t n4 = aligned ? n / 4 : 0;
  if (n4) {
    size_t need = (n4 + (size_t)kThreads * kUnroll - 1) / ((size_t)kThreads * kUnroll);
    DLAS_REQUIRE(need <= 0x7fffffffu, "dlas_gilbert_f32: n too large for one launch");
    unsigned grid = (unsigned)need;
    int pi = prof_begin(DLAS_PROF_GILBERT, st);
    gilbert_vec_kernel<MODE><<<grid, kThreads, 0, st>>>((const float4*)P, (const float4*)S,
                                                        (const float4*)R, (float4*)q, n4);
    prof_end(pi, st);
    DLAS_LAUNCH_CHECK();
  }
  size_t done = n4 * 4;
  if (done < n) {
    size_t rem = n - done;
    size_t need = (rem + kThreads - 1) / kThreads;
    size_t cap = (size_t)sms * 16;
    unsigned grid = (unsigned)(need < cap ? need : cap);
    gilbert_scalar_kernel<MODE><<<grid, kThreads, 0, st>>>(P + done, S + done, R + done, q + done, rem);
    DLAS_LAUNCH_CHECK();
  }
  return DLAS_OK;
}

}  // namespace
}  // namespace dlas

#ifndef GILBERT_TUNE
extern "C" int dlas_gilbert_f32(const float* P, const float* S, const float* R, float* q, size_t n,
                                int variant, void* stream) {
  using namespace dlas;
  DLAS_CHECK_DEVICE();
  if (n == 0) return DLAS_OK;
  DLAS_REQUIRE(P && S && R && q, "dlas_gilbert_f32: null pointer");
  DLAS_REQUIRE(variant >= 0 && variant <= 7, "dlas_gilbert_f32: variant %d not in 0..7", variant);
  cudaStream_t st = (cudaStream_t)stream;
  switch (variant) {
    case 0: return launch<0>(P, S, R, q, n, st);
    case 1: return launch<1>(P, S, R, q, n, st);
    case 2: return launch<2>(P, S, R, q, n, st);
    case 3: return launch<3>(P, S, R, q, n, st);
    case 4: return launch<4>(P, S, R, q, n, st);
    case 5: return launch<5>(P, S, R, q, n, st);
    case 6: return launch<6>(P, S, R, q, n, st);
    default: return launch<7>(P, S, R, q, n, st);
  }
}
#endif  // GILBERT_TUNE
