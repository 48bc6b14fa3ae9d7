// SmoothQuant / AWQ-lite support kernels.  All are single-pass and HBM-bound.
//
//  mo_scale_cols_amax    y = T(x * s[col]) (+ amax of the result per tensor / per row in the same pass): the weight
//                        fold + re-calibration of apply_pre_quant_scale_and_smooth (model_calib.py:1208-1270) and the
//                        `input * pre_quant_scale` pass of TensorQuantizer.forward (tensor_quantizer.py:1143).
//  mo_awq_weight_scale   AWQ get_weight_scale (model_calib.py:1453-1469): per column, sum over rows of
//                        |w| / (blockmax + tiny), blocks of `block` along the last dim.
//  mo_abs_col_sum        AWQ get_act_scale (model_calib.py:1471-1472): per column, sum over rows of |x|.
//  mo_sq_diff_mean       AWQ update_loss (model_calib.py:1485-1491): acc += mean((T(a - b))^2).
//
// The two column reductions write per-split partial sums [splits, cols]; the caller adds the few partial rows in a
// fixed order, so results do not depend on scheduling.
#include <float.h>

#include "common.cuh"

namespace mo {

// ------------------------------------------------------------------------------------------------
// y = T(x * s[col]); one WARP per row, lanes stride over the row's 32-byte vectors.
// amax_mode: 0 none, 1 per tensor (amax_bits[0]), 2 per row (amax_bits[row]); merged with atomicMax on the
// bit pattern of |y| (order-independent, NaN sorts above inf).
// ------------------------------------------------------------------------------------------------
template <typename T, bool kVec>
__global__ void __launch_bounds__(kThreads) scale_cols_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows,
                                                              int64_t cols, const float* __restrict__ s,
                                                              uint32_t* __restrict__ amax_bits, int amax_mode) {
  constexpr int E = Elem<T>::kPerVec;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * kThreads) >> 5;
  uint32_t m_all = 0;
  for (int64_t r = warp; r < rows; r += n_warps) {
    uint32_t m = 0;
    if constexpr (kVec) {
      const int64_t nv = cols / E;
      const V32* xv = reinterpret_cast<const V32*>(x + r * cols);
      V32* yv = reinterpret_cast<V32*>(y + r * cols);
      for (int64_t c0 = lane; c0 < nv; c0 += 64) {
        V32 v[2];
        const int64_t c1 = c0 + 32;
        v[0] = ld_stream(xv + c0);
        if (c1 < nv) v[1] = ld_stream(xv + c1);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int64_t c = u ? c1 : c0;
          if (c >= nv) break;
          float f[E];
          Elem<T>::unpack(v[u], f);
          const float* ps = s + c * E;
#pragma unroll
          for (int i = 0; i < E; i += 4) {
            const float4 s4 = __ldg(reinterpret_cast<const float4*>(ps + i));
            f[i] = __fmul_rn(f[i], s4.x); f[i + 1] = __fmul_rn(f[i + 1], s4.y);
            f[i + 2] = __fmul_rn(f[i + 2], s4.z); f[i + 3] = __fmul_rn(f[i + 3], s4.w);
          }
          Elem<T>::pack(f, v[u]);
          st_stream(yv + c, v[u]);
          if (amax_mode) {
            Elem<T>::unpack(v[u], f);
#pragma unroll
            for (int i = 0; i < E; ++i) m = max(m, __float_as_uint(f[i]) & 0x7fffffffu);
          }
        }
      }
    } else {
      for (int64_t c = lane; c < cols; c += 32) {
        T t;
        Elem<T>::store(&t, __fmul_rn(Elem<T>::load(x + r * cols + c), __ldg(s + c)));
        y[r * cols + c] = t;
        m = max(m, __float_as_uint(Elem<T>::load(&t)) & 0x7fffffffu);
      }
    }
    if (amax_mode == 2) {
      m = warp_max_u32(m);
      if (lane == 0) atomicMax(amax_bits + r, m);
    }
    m_all = max(m_all, m);
  }
  if (amax_mode == 1) {
    m_all = warp_max_u32(m_all);
    if (lane == 0 && m_all) atomicMax(amax_bits, m_all);
  }
}

// ------------------------------------------------------------------------------------------------
// Column reductions over a [rows, cols] matrix.  CTA = 8 warps on 8 different rows of the same 32-vector column
// strip; grid = (strips, splits).  Each lane keeps E fp32 partial sums; the 8 warps are added in warp order through
// shared memory and the CTA writes partial[split, strip columns].
//   kWeightScale = false: sum |x|
//   kWeightScale = true : sum T(|w| / T(blockmax + tiny)), block = lanes_per_block * E columns
// ------------------------------------------------------------------------------------------------
template <typename T>
struct Tiny;
template <> struct Tiny<float> { static constexpr float v = FLT_MIN; };
template <> struct Tiny<__nv_bfloat16> { static constexpr float v = FLT_MIN; };          // bf16 tiny == 2^-126
template <> struct Tiny<__half> { static constexpr float v = 6.103515625e-05f; };         // f16 tiny == 2^-14

template <typename T>
__device__ __forceinline__ float round_T(float f) {
  T t;
  Elem<T>::store(&t, f);
  return Elem<T>::load(&t);
}

// Tried: four rows in flight per warp (loads issued before the arithmetic) — 80-90 registers instead of ~50 cost more
// occupancy than the extra loads gained (weight_scale 0.234 -> 0.246 ms, |x| column sum 0.098 -> 0.105 ms at
// [16384, 14336] bf16), so the loop stays one row per warp per iteration.
template <typename T, bool kWeightScale>
__global__ void __launch_bounds__(kThreads) col_sum_kernel(const T* __restrict__ x, int64_t rows, int64_t cols,
                                                           int lanes_per_block, float* __restrict__ partial) {
  constexpr int E = Elem<T>::kPerVec;
  __shared__ float red[kThreads / 32][32 * E + 1];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t nv = cols / E;
  const int64_t cv = (int64_t)blockIdx.x * 32 + lane;  // this lane's vector column
  const bool live = cv < nv;
  const int64_t rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per, r1 = min(rows, r0 + rows_per);
  float acc[E];
#pragma unroll
  for (int i = 0; i < E; ++i) acc[i] = 0.f;
  for (int64_t r = r0 + w; r < r1; r += kThreads / 32) {
    float f[E];
    uint32_t m = 0;
    if (live) {
      const V32 v = ld_stream(reinterpret_cast<const V32*>(x + r * cols) + cv);
      Elem<T>::unpack(v, f);
#pragma unroll
      for (int i = 0; i < E; ++i) {
        f[i] = fabsf(f[i]);
        m = max(m, __float_as_uint(f[i]));
      }
    }
    if constexpr (kWeightScale) {
      for (int o = 1; o < lanes_per_block; o <<= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (live) {
        const float den = round_T<T>(__fadd_rn(__uint_as_float(m), Tiny<T>::v));
        const DivSeq2 dv(den);  // the division's own instruction sequence, reciprocal hoisted (common.cuh)
#pragma unroll
        for (int i = 0; i < E; i += 2) {
          float q0 = f[i], q1 = f[i + 1];
          if (dv.ok) dv(q0, q1);
          else { q0 = __fdiv_rn(q0, den); q1 = __fdiv_rn(q1, den); }
          acc[i] = __fadd_rn(acc[i], round_T<T>(q0));
          acc[i + 1] = __fadd_rn(acc[i + 1], round_T<T>(q1));
        }
      }
    } else if (live) {
#pragma unroll
      for (int i = 0; i < E; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < E; ++i) red[w][lane * E + i] = acc[i];
  __syncthreads();
  for (int j = threadIdx.x; j < 32 * E; j += kThreads) {
    const int64_t col = (int64_t)blockIdx.x * 32 * E + j;
    if (col < cols) {
      float t = red[0][j];
#pragma unroll
      for (int k = 1; k < kThreads / 32; ++k) t = __fadd_rn(t, red[k][j]);
      partial[(int64_t)blockIdx.y * cols + col] = t;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// acc[0] += sum((T(a - b))^2) / n          (double accumulator; the fp32 partials are per thread)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) sq_diff_kernel(const T* __restrict__ a, const T* __restrict__ b, int64_t n,
                                                           bool vec, double inv_n, double* __restrict__ acc) {
  constexpr int E = Elem<T>::kPerVec;
  __shared__ double part[kThreads / 32];
  double d = 0.0;
  if (vec) {
    const int64_t nv = n / E;
    const V32* av = reinterpret_cast<const V32*>(a);
    const V32* bv = reinterpret_cast<const V32*>(b);
    for (int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x; c < nv; c += (int64_t)gridDim.x * kThreads) {
      const V32 va = ld_stream(av + c), vb = ld_stream(bv + c);
      float fa[E], fb[E];
      Elem<T>::unpack(va, fa);
      Elem<T>::unpack(vb, fb);
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < E; ++i) {
        const float t = round_T<T>(__fsub_rn(fa[i], fb[i]));
        s = __fmaf_rn(t, t, s);
      }
      d += (double)s;
    }
    if (blockIdx.x == 0) {
      const int64_t i = nv * E + threadIdx.x;
      if (i < n) {
        const float t = round_T<T>(__fsub_rn(Elem<T>::load(a + i), Elem<T>::load(b + i)));
        d += (double)__fmul_rn(t, t);
      }
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
      const float t = round_T<T>(__fsub_rn(Elem<T>::load(a + i), Elem<T>::load(b + i)));
      d += (double)__fmul_rn(t, t);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = d;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kThreads / 32; ++k) t += part[k];
    atomicAdd(acc, t * inv_n);
  }
}

template <typename T>
static int scale_cols_launch(const void* x, void* y, int64_t rows, int64_t cols, const float* s, uint32_t* amax_bits,
                             int amax_mode, cudaStream_t st) {
  constexpr int E = Elem<T>::kPerVec;
  const bool vec = aligned32(x) && aligned32(y) && cols % E == 0 && (reinterpret_cast<uintptr_t>(s) & 15u) == 0;
  const int64_t tiles = (rows * 32 + kThreads - 1) / kThreads;
  if (vec)
    scale_cols_kernel<T, true><<<resident_grid((const void*)scale_cols_kernel<T, true>, tiles), kThreads, 0, st>>>(static_cast<const T*>(x), static_cast<T*>(y), rows, cols, s, amax_bits, amax_mode);
  else
    scale_cols_kernel<T, false><<<resident_grid((const void*)scale_cols_kernel<T, false>, tiles), kThreads, 0, st>>>(static_cast<const T*>(x), static_cast<T*>(y), rows, cols, s, amax_bits, amax_mode);
  return check_launch("mo_scale_cols_amax");
}

template <typename T>
static int col_sum_launch(const void* x, int64_t rows, int64_t cols, int block, float* partial, int splits,
                          cudaStream_t st, const char* what) {
  constexpr int E = Elem<T>::kPerVec;
  if (!aligned32(x) || cols % E != 0) {
    set_error("%s: needs a 32-byte aligned matrix with cols %% %d == 0", what, E);
    return MO_ERR_ARG;
  }
  const dim3 grid((unsigned)((cols / E + 31) / 32), (unsigned)splits);
  if (block > 0) {
    const int lanes = block / E;
    if (block % E != 0 || lanes < 1 || lanes > 32 || (lanes & (lanes - 1)) != 0 || cols % block != 0) {
      set_error("%s: block=%d must be %d * 2^k (k <= 5) and divide cols", what, block, E);
      return MO_ERR_ARG;
    }
    col_sum_kernel<T, true><<<grid, kThreads, 0, st>>>(static_cast<const T*>(x), rows, cols, lanes, partial);
  } else {
    col_sum_kernel<T, false><<<grid, kThreads, 0, st>>>(static_cast<const T*>(x), rows, cols, 1, partial);
  }
  return check_launch(what);
}

}  // namespace mo

using namespace mo;

extern "C" int mo_scale_cols_amax(const void* x, void* y, int64_t rows, int64_t cols, int dtype, const float* scale,
                                  uint32_t* amax_bits, int amax_mode, void* stream) {
  if (rows < 0 || cols < 0) { set_error("mo_scale_cols_amax: bad shape"); return MO_ERR_SHAPE; }
  if (rows == 0 || cols == 0) return MO_OK;
  if (!x || !y || !scale) { set_error("mo_scale_cols_amax: NULL pointer"); return MO_ERR_NULL; }
  if (amax_mode < 0 || amax_mode > 2 || (amax_mode && !amax_bits)) { set_error("mo_scale_cols_amax: bad amax_mode / amax_bits"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case MO_F32: return scale_cols_launch<float>(x, y, rows, cols, scale, amax_bits, amax_mode, s);
    case MO_F16: return scale_cols_launch<__half>(x, y, rows, cols, scale, amax_bits, amax_mode, s);
    case MO_BF16: return scale_cols_launch<__nv_bfloat16>(x, y, rows, cols, scale, amax_bits, amax_mode, s);
  }
  set_error("mo_scale_cols_amax: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

static int col_sum_entry(const void* x, int64_t rows, int64_t cols, int dtype, int block, float* partial, int splits,
                         void* stream, const char* what) {
  if (rows < 0 || cols < 0) { set_error("%s: bad shape", what); return MO_ERR_SHAPE; }
  if (!partial || splits < 1 || splits > 65535) { set_error("%s: bad partial / splits", what); return MO_ERR_ARG; }
  if (rows == 0 || cols == 0) return MO_OK;
  if (!x) { set_error("%s: NULL pointer", what); return MO_ERR_NULL; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case MO_F32: return col_sum_launch<float>(x, rows, cols, block, partial, splits, s, what);
    case MO_F16: return col_sum_launch<__half>(x, rows, cols, block, partial, splits, s, what);
    case MO_BF16: return col_sum_launch<__nv_bfloat16>(x, rows, cols, block, partial, splits, s, what);
  }
  set_error("%s: unsupported dtype %d", what, dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_awq_weight_scale(const void* w, int64_t rows, int64_t cols, int dtype, int block, float* partial,
                                   int splits, void* stream) {
  if (block < 1) { set_error("mo_awq_weight_scale: block < 1"); return MO_ERR_ARG; }
  return col_sum_entry(w, rows, cols, dtype, block, partial, splits, stream, "mo_awq_weight_scale");
}

extern "C" int mo_abs_col_sum(const void* x, int64_t rows, int64_t cols, int dtype, float* partial, int splits,
                              void* stream) {
  return col_sum_entry(x, rows, cols, dtype, 0, partial, splits, stream, "mo_abs_col_sum");
}

extern "C" int mo_sq_diff_mean(const void* a, const void* b, int64_t n, int dtype, double* acc, void* stream) {
  if (n < 0) { set_error("mo_sq_diff_mean: n < 0"); return MO_ERR_SHAPE; }
  if (!acc) { set_error("mo_sq_diff_mean: acc is NULL"); return MO_ERR_NULL; }
  if (n == 0) return MO_OK;
  if (!a || !b) { set_error("mo_sq_diff_mean: NULL pointer"); return MO_ERR_NULL; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool vec = aligned32(a) && aligned32(b);
  const double inv_n = 1.0 / (double)n;
#define MO_SQ(T) { constexpr int E = Elem<T>::kPerVec; \
    const int grid = persistent_grid(((vec ? n / E : n) + kThreads - 1) / kThreads, 8); \
    sq_diff_kernel<T><<<grid, kThreads, 0, s>>>(static_cast<const T*>(a), static_cast<const T*>(b), n, vec, inv_n, acc); \
    return check_launch("mo_sq_diff_mean"); }
  switch (dtype) {
    case MO_F32: MO_SQ(float)
    case MO_F16: MO_SQ(__half)
    case MO_BF16: MO_SQ(__nv_bfloat16)
  }
#undef MO_SQ
  set_error("mo_sq_diff_mean: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}
