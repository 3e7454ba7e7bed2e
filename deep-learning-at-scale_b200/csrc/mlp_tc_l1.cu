// First layer of the wide-input ANN d0-256-256-1 (dynamic ANN, Readme.md:19: d0 = 24 lags x 8 features = 192) on tcgen05,
// as two kernels around the CTA-pair kernel of mlp_tc.cu.  The fused pair kernel keeps every weight-gradient accumulator
// in TMEM; for d0 = 192 dW1 alone would need 2 x 208 of the 512 columns next to dW2's 256, and W1 as an fp16 tile is
// 104 KB next to W2, H1 and dZ2.  So layer 1 runs outside:
//   l1_fwd_tc_kernel  H1 = relu([x | 1] [W1 ; b1])  per 128-row tile: window rows -> fp16 A tile (the lag windows are read
//                     in place from the raw series: row i = d0 contiguous floats at X + i * ldx), W1 resident in shared
//                     memory (MN-major B operand), Z1 in TMEM (N = 256), epilogue -> fp16 H1 written over the A tile in the
//                     pair kernel's own swizzled layout and bulk-stored as a 64 KB tile IMAGE, which the pair kernel
//                     fetches with one bulk copy instead of computing Z1 / E1.
//   l1_dw_tc_kernel   dW1^T[unit][k] += (dZ1p0 + dZ1p1)^T [x | 1] as a long-K GEMM over all rows (the pair kernel stores
//                     each CTA's partial dZ1 tile as an image; the loader adds the two), CTA b owns 128 of the 256 units
//                     (accumulator lanes) and every 74th tile; column d0 of the accumulator is db1.
//   l1_reduce_kernel  partials -> dW1, db1 [-> optimiser update], in slice order (deterministic).
#include "common.cuh"
#include "optim_math.cuh"
#include "tc_common.cuh"

namespace dlas {
namespace {
using namespace tc;

constexpr int TILE = 128, D1 = 256;
constexpr int N_WORK = 256, ISSUER_WARP = 8, NT = 288;
constexpr uint32_t IMG_TILE_BYTES = 65536;      // [128 rows][256 columns] fp16, column-blocked SWIZZLE_128B

__host__ __device__ inline int k1_of(int d0) { return (d0 + 1 + 15) & ~15; }   // features + the ones column, padded to 16

// rows [row0, row0 + 128) of the (strided) input as the fp16 tile [x | 1 | 0] (K1 columns used): thread (r, half) converts the
// 8-column chunks half, half + 2, ...  d0 is a multiple of 8, ldx a multiple of 4 (16-byte row loads).
__device__ __forceinline__ void load_x_tile(uint32_t sXw, const float* __restrict__ X, size_t ldx, size_t row0, size_t n_rows,
                                            int d0, int k1, int tid) {
  const int r = tid & 127, half = tid >> 7;
  const bool in = row0 + r < n_rows;
  const float* src = X + (row0 + r) * ldx;
  const int n_chunks = k1 >> 3;
  constexpr int U = 7;                                     // chunks (2 loads each) in flight per round: d0 = 192 -> 2 rounds
  for (int c0 = half; c0 < n_chunks; c0 += 2 * U) {
    float4 a[U], e[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + 2 * u;
      a[u] = make_float4(0.f, 0.f, 0.f, 0.f); e[u] = a[u];
      if (c < n_chunks) {
        if (c * 8 < d0) {
          if (in) { a[u] = __ldg(reinterpret_cast<const float4*>(src + c * 8)); e[u] = __ldg(reinterpret_cast<const float4*>(src + c * 8 + 4)); }
        } else if (c * 8 == d0) {
          a[u].x = 1.0f;                                   // the ones column multiplies the bias row of [W1 ; b1 ; 0]
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + 2 * u;
      if (c < n_chunks)
        sts128(sXw + tile_chunk_off(TILE, r, c * 8), make_uint4(pack_half2(a[u].x, a[u].y), pack_half2(a[u].z, a[u].w), pack_half2(e[u].x, e[u].y), pack_half2(e[u].z, e[u].w)));
    }
  }
}

// ---- forward: H1 tile images -------------------------------------------------------------------------------------
namespace fw {
constexpr uint32_t S_W1 = 0 /*[256 rows][256] fp16 128 KB*/, S_A = 131072 /*64 KB*/, S_BAR = S_A + 65536;
enum { B_A = 0, B_Z, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8, SMEM_BYTES = S_TMEM + 16, TM_COLS = 256;
}  // namespace fw

__global__ void __launch_bounds__(NT, 1)
l1_fwd_tc_kernel(const int d0, const float* __restrict__ params, const float* __restrict__ X, const size_t ldx, const size_t n_rows,
                 uint8_t* __restrict__ h1_img, unsigned* __restrict__ err_flag, const unsigned call_id) {
  using namespace fw;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sW1 = sbase + S_W1, sA = sbase + S_A;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int k1 = k1_of(d0);
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  bool ok = (sbase & 1023u) == 0;
  if (tid == 0) {
    mbar_init(&bars[B_A], N_WORK / 32); mbar_init(&bars[B_Z], 1);
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  asm volatile("griddepcontrol.wait;" ::: "memory");           // params come from the previous step's update
  // [W1 ; b1 ; 0] -> fp16 tile: rows = k (features, bias row, zero rows), columns = units
  for (int i0 = 0; i0 < k1 * 32; i0 += NT * 4) {          // 4 chunks (8 loads) in flight per thread
    float4 a[4], e[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = i0 + u * NT + tid, n = idx >> 5, c = (idx & 31) * 8;
      a[u] = make_float4(0.f, 0.f, 0.f, 0.f); e[u] = a[u];
      if (idx < k1 * 32 && n <= d0) {                      // rows 0..d0-1 weights, row d0 the bias (b1 follows W1), then zero rows
        const float* src = params + (size_t)n * D1 + c;
        a[u] = __ldg(reinterpret_cast<const float4*>(src));
        e[u] = __ldg(reinterpret_cast<const float4*>(src + 4));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = i0 + u * NT + tid;
      if (idx < k1 * 32)
        sts128(sW1 + tile_chunk_off(256, idx >> 5, (idx & 31) * 8),
               make_uint4(pack_half2(a[u].x, a[u].y), pack_half2(a[u].z, a[u].w), pack_half2(e[u].x, e[u].y), pack_half2(e[u].z, e[u].w)));
    }
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tm = *tmem_slot;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  uint32_t it = 0;
  if (warp == ISSUER_WARP) {
    const bool leader = elect_one();
    constexpr uint32_t ID_Z = make_idesc(128, 256, false, true);
    const uint64_t dA = desc_kmajor(sA, TILE, 0, 0), dW = desc_mnmajor(sW1, 256, 0, 0);
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      ok = mbar_wait(&bars[B_A], it & 1) && ok;
      fence_after();
      if (leader) {
        for (int i = 0; i < k1 / 16; ++i) mma_f16(tm, dA + koff_kmajor(TILE, i), dW + koff_mnmajor(i), ID_Z, i > 0);
        mma_commit(&bars[B_Z]);
      }
      __syncwarp();
    }
  } else {
    const int q4 = warp & 3, hh = warp >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      load_x_tile(sA, X, ldx, tile * TILE, n_rows, d0, k1, tid);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[B_A]);
      ok = mbar_wait(&bars[B_Z], it & 1) && ok;       // Z1 (bias included) is in TMEM; the MMAs have read the A tile
      fence_after();
      // H1 = relu(Z1) as packed halves over the A tile, in the pair kernel's tile layout
#pragma unroll 1
      for (int c32 = 0; c32 < 4; ++c32) {
        uint32_t v[32];
        tmem_ld32(tm + lane_base + hh * 128 + c32 * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g8 = 0; g8 < 4; ++g8) {
          uint4 o;
          o.x = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 0]), __uint_as_float(v[g8 * 8 + 1])));
          o.y = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 2]), __uint_as_float(v[g8 * 8 + 3])));
          o.z = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 4]), __uint_as_float(v[g8 * 8 + 5])));
          o.w = relu_h2(pack_half2(__uint_as_float(v[g8 * 8 + 6]), __uint_as_float(v[g8 * 8 + 7])));
          sts128(sA + tile_chunk_off(TILE, row, hh * 128 + c32 * 32 + g8 * 8), o);
        }
      }
      fence_before();
      fence_async_smem();
      named_bar_sync(1, N_WORK);
      if (tid == 0) {            // the finished tile leaves as one image; the A tile is free once the copy engine has read it
        bulk_s2g(h1_img + tile * (size_t)IMG_TILE_BYTES, sA, IMG_TILE_BYTES);
        bulk_commit();
        bulk_wait_read0();
      }
      named_bar_sync(1, N_WORK);
    }
    if (tid == 0) bulk_wait0();  // every image has reached memory before the kernel ends
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  fence_before();
  __syncthreads();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// ---- weight gradient of layer 1 ------------------------------------------------------------------------------------
namespace dw {
constexpr uint32_t S_DZ = 0 /*2 x [128 rows][128 units] 32 KB*/, S_X = 65536 /*2 x [128][256] 64 KB*/, S_BAR = S_X + 131072;
enum { B_FULL0 = 0, B_FULL1, B_FREE0, B_FREE1, B_ACC, N_BARS };
constexpr uint32_t S_TMEM = S_BAR + N_BARS * 8, SMEM_BYTES = S_TMEM + 16, TM_COLS = 256;
}  // namespace dw
__host__ __device__ inline size_t l1_part_floats(int d0) { return (size_t)(d0 + 1) * 128; }   // per CTA: [k 0..d0][128 units]

__device__ __forceinline__ uint32_t add_h2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("add.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

__global__ void __launch_bounds__(NT, 1)
l1_dw_tc_kernel(const int d0, const float* __restrict__ X, const size_t ldx, const size_t n_rows, const uint8_t* __restrict__ dz1_img,
                const size_t img_stride /*bytes between the two CTAs' partial images*/, float* __restrict__ partials,
                unsigned* __restrict__ err_flag, const unsigned call_id) {
  using namespace dw;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S_TMEM);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sDZ = sbase + S_DZ, sX = sbase + S_X;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int jc = blockIdx.x & 1, slice = blockIdx.x >> 1, n_slices = gridDim.x >> 1;
  const int k1 = k1_of(d0);
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  const int my_tiles = (n_tiles > (size_t)slice) ? (int)((n_tiles - slice + n_slices - 1) / n_slices) : 0;
  bool ok = (sbase & 1023u) == 0;
  if (tid == 0) {
    mbar_init(&bars[B_FULL0], N_WORK / 32); mbar_init(&bars[B_FULL1], N_WORK / 32);
    mbar_init(&bars[B_FREE0], 1); mbar_init(&bars[B_FREE1], 1); mbar_init(&bars[B_ACC], 1);
    mbar_fence_init();
  }
  if (warp == ISSUER_WARP) tmem_alloc(tmem_slot, TM_COLS);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tm = *tmem_slot;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  float* my = partials + (size_t)blockIdx.x * l1_part_floats(d0);
  if (warp == ISSUER_WARP) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc(128, k1, true, true);
    for (int c = 0; c < my_tiles; ++c) {
      const int b = c & 1;
      ok = mbar_wait(&bars[B_FULL0 + b], (c >> 1) & 1) && ok;
      fence_after();
      if (leader) {
        const uint64_t dDZ = desc_mnmajor(sDZ + b * 32768, TILE, 0, 0), dXm = desc_mnmajor(sX + b * 65536, TILE, 0, 0);
#pragma unroll
        for (int i = 0; i < TILE / 16; ++i)       // acc[unit][k] += dZ1^T [x | 1]   (K = the tile's 128 rows)
          mma_f16(tm, dDZ + koff_mnmajor(i), dXm + koff_mnmajor(i), idesc, !(c == 0 && i == 0));
        mma_commit(&bars[B_FREE0 + b]);
        if (c == my_tiles - 1) mma_commit(&bars[B_ACC]);
      }
      __syncwarp();
    }
  } else {
    // dZ1 of this CTA's 128 units: column blocks 2jc, 2jc+1 of the two partial tile images, added as packed halves.  An
    // image is byte-identical to the shared-memory tile, so the copy keeps every 16-byte chunk in place.  The chunks of
    // tile c+1 are requested before tile c's rows are converted, so their latency hides under that work.
    uint4 a[8], e[8];
    auto dz_issue = [&](int c) {
      const size_t tile = (size_t)slice + (size_t)c * n_slices;
      const uint8_t* g0 = dz1_img + tile * (size_t)IMG_TILE_BYTES + (size_t)jc * 32768;
      const uint8_t* g1 = g0 + img_stride;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        a[u] = __ldcs(reinterpret_cast<const uint4*>(g0 + (size_t)(tid + u * N_WORK) * 16));
        e[u] = __ldcs(reinterpret_cast<const uint4*>(g1 + (size_t)(tid + u * N_WORK) * 16));
      }
    };
    if (my_tiles > 0) dz_issue(0);
    for (int c = 0; c < my_tiles; ++c) {
      const int b = c & 1;
      const size_t tile = (size_t)slice + (size_t)c * n_slices;
      if (c >= 2) ok = mbar_wait(&bars[B_FREE0 + b], ((c >> 1) - 1) & 1) && ok;
#pragma unroll
      for (int u = 0; u < 8; ++u)
        sts128(sDZ + b * 32768 + (uint32_t)(tid + u * N_WORK) * 16,
               make_uint4(add_h2(a[u].x, e[u].x), add_h2(a[u].y, e[u].y), add_h2(a[u].z, e[u].z), add_h2(a[u].w, e[u].w)));
      if (c + 1 < my_tiles) dz_issue(c + 1);
      load_x_tile(sX + b * 65536, X, ldx, tile * TILE, n_rows, d0, k1, tid);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[B_FULL0 + b]);
    }
    if (warp < 4) {   // drain: lane = unit jc*128 + j, column n = input feature (n < d0) or the bias (n == d0)
      const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
      const int j = warp * 32 + lane;
      if (my_tiles > 0) {
        ok = mbar_wait(&bars[B_ACC], 0) && ok;
        fence_after();
#pragma unroll 1
        for (int n0 = 0; n0 < k1; n0 += 16) {
          uint32_t v[16];
          tmem_ld16(tm + lane_base + n0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (n0 + i <= d0) my[(size_t)(n0 + i) * 128 + j] = __uint_as_float(v[i]);
        }
        fence_before();
      } else {
        for (int n = 0; n <= d0; ++n) my[(size_t)n * 128 + j] = 0.f;
      }
    }
  }
  if (!ok && lane == 0 && err_flag != nullptr) atomicExch(err_flag, call_id);
  fence_before();
  __syncthreads();
  if (warp == ISSUER_WARP) tmem_dealloc(tm, TM_COLS);
}

// dW1[k][unit], db1[unit] = sum over the slices of the unit's half, in slice order; MODE 1 applies the optimiser update
template <int MODE>
__global__ void __launch_bounds__(256)
l1_reduce_kernel(const int d0, const float* __restrict__ partials, const int n_ctas, unsigned* __restrict__ err_flag,
                 const unsigned call_id, float* __restrict__ out, const OptScalars opt, const float grad_scale,
                 float* __restrict__ params, float* __restrict__ m, float* __restrict__ v) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // 64 consecutive entries of [W1 (d0 x 256) ; b1 (256)] per block; the slices are cut into 4 contiguous parts summed by 4
  // thread groups (more loads in flight), then added in part order -> deterministic
  __shared__ float red[3][64];
  const int i = blockIdx.x * 64 + (threadIdx.x & 63), grp = threadIdx.x >> 6;   // (d0 + 1) * 256 is a multiple of 64
  const int k = i / D1, unit = i - k * D1, jc = unit >> 7, j = unit & 127;
  const size_t pf = l1_part_floats(d0);
  const int n_src = n_ctas >> 1, per = (n_src + 3) / 4, s0 = grp * per, s1 = min(n_src, s0 + per);
  float s = 0.f;
#pragma unroll 8
  for (int q = s0; q < s1; ++q) s += partials[(size_t)(2 * q + jc) * pf + (size_t)k * 128 + j];
  if (grp > 0) red[grp - 1][threadIdx.x & 63] = s;
  __syncthreads();
  if (grp > 0) return;
  s += red[0][threadIdx.x]; s += red[1][threadIdx.x]; s += red[2][threadIdx.x];
  if (*err_flag == call_id) s = __int_as_float(0x7fc00000);
  if (!(fabsf(s) <= 3.0e38f)) err_flag[1] = call_id;     // status words: see tc_reduce_kernel
  if (MODE == 0) {
    out[i] = s;
  } else {
    float p = params[i], mi = m[i], vi = (opt.kind == 0) ? 0.f : v[i];
    opt_update(opt, s * grad_scale, p, mi, vi);
    params[i] = p;
    m[i] = mi;
    if (opt.kind != 0) v[i] = vi;
  }
}

}  // namespace

// ---- entry points used by mlp_tc.cu ----------------------------------------------------------------------------------
size_t l1_partial_bytes(int d0) { return (size_t)((sm_count() / 2) * 2) * l1_part_floats(d0) * sizeof(float); }

int l1_fwd(int d0, const float* params, const float* X, size_t ldx, size_t n_rows, uint8_t* h1_img, unsigned* err_flag, unsigned call_id,
           cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DLAS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    DLAS_CUDA(cudaFuncSetAttribute(l1_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fw::SMEM_BYTES));
    DLAS_CUDA(cudaFuncSetAttribute(l1_dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dw::SMEM_BYTES));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const size_t n_tiles = (n_rows + TILE - 1) / TILE;
  const int grid = (int)(n_tiles < (size_t)sm_count() ? n_tiles : (size_t)sm_count());
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(NT);
  cfg.dynamicSmemBytes = fw::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, l1_fwd_tc_kernel, d0, params, X, ldx, n_rows, h1_img, err_flag, call_id));
  return DLAS_OK;
}

// dW1, db1 from the two partial dZ1 image sets: MODE 0 -> grad_sum[0 .. (d0+1)*256), MODE 1 -> optimiser update of W1, b1
int l1_dw(int d0, const float* X, size_t ldx, size_t n_rows, const uint8_t* dz1_img, size_t img_stride, float* partials,
          unsigned* err_flag, unsigned call_id, int mode, float* grad_sum, const OptScalars& opt, float grad_scale, float* params,
          float* m, float* v, cudaStream_t st) {
  const int grid = (sm_count() / 2) * 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(NT);
  cfg.dynamicSmemBytes = dw::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  DLAS_CUDA(cudaLaunchKernelEx(&cfg, l1_dw_tc_kernel, d0, X, ldx, n_rows, dz1_img, img_stride, partials, err_flag, call_id));
  cfg.gridDim = dim3((unsigned)(((d0 + 1) * D1) / 64));
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = 0;
  if (mode == 0)
    DLAS_CUDA(cudaLaunchKernelEx(&cfg, l1_reduce_kernel<0>, d0, (const float*)partials, grid, err_flag, call_id,
                                 grad_sum, opt, grad_scale, params, m, v));
  else
    DLAS_CUDA(cudaLaunchKernelEx(&cfg, l1_reduce_kernel<1>, d0, (const float*)partials, grid, err_flag, call_id,
                                 grad_sum, opt, grad_scale, params, m, v));
  return DLAS_OK;
}

}  // namespace dlas
