// MX-format dynamic block fake quantize-dequantize (element formats E4M3 / E5M2 / INT8 / E2M1 / E1M2 / E0M3 / E3M0 /
// E3M2 / E2M3 under E8M0 power-of-two block scales, or under a low-precision scale format with an optional global
// amax), and the MXFP8 / MXFP4 real-quant packers that share its block-amax -> E8M0 step.
//
// Replaces the reference's `fused_amax_convert` CUDA extension (gemm/tensor_quant_mx.cu:239-387, tensor_quant_mx.h:39-245;
// reached from tensor_quant.py:157-195 for every preset whose block_sizes carry scale_bits (8,0): MXFP8 / MXFP6 / MXFP4 /
// MXINT8_DEFAULT_CFG) and the eager packers MXFP8QTensor / MXFP4QTensor (qtensor/mxfp8_tensor.py:41-262,
// qtensor/mxfp4_tensor.py:36-144).  HBM-bound: 4 B per bf16 element for the fake-quant, 3.03 B for the MXFP8 pack.
//
// The reference kernel is a 32x4-thread CTA with ONE ELEMENT PER THREAD and a cub block reduce per block.  Here one
// thread owns one whole block (one or two 32-byte vectors): the block max is a register fmax chain, the E8M0 exponent is
// integer arithmetic on the bits of amax / format_max, the element conversions are single F2FP instructions
// (cvt.rn.satfinite.{e4m3x2,e5m2x2,e2m1x2,e3m2x2,e2m3x2}.f32) and the scales are exact powers of two.
//
// Numerics.  The reference extension is compiled with --use_fast_math (quantization/extensions.py:69), so
// `amax / format_max` is div.approx.ftz.f32: MO_NUMERICS_GPU issues that instruction; MO_NUMERICS_CPU divides exactly
// (what the header's host functions compute).  They differ only for blocks whose amax / format_max lands within 2 ulp of
// a power of two.  The pure-torch packers take ceil(log2(.)) in fp32 (log2f here as on a CUDA tensor).
#include <math.h>

#include "common.cuh"

namespace mo {

// ---------------------------------------------------------------------------------------------------------------------
// Element formats (codes = the reference's `Types` enum order, tensor_quant_mx.h:39).
// ---------------------------------------------------------------------------------------------------------------------
__host__ __device__ inline float fmt_max(int f) {  // tensor_quant_mx.h:177-201
  switch (f) {
    case MO_FMT_E4M3: return 448.f;
    case MO_FMT_E5M2: return 57344.f;
    case MO_FMT_INT8: return 127.f;
    case MO_FMT_E2M1: return 6.f;
    case MO_FMT_E0M3: return 7.f;
    case MO_FMT_E3M0: return 16.f;
    case MO_FMT_E1M2: return 3.5f;
    case MO_FMT_E3M2: return 28.f;
    case MO_FMT_E2M3: return 7.5f;
  }
  return 0.f;
}

__device__ __forceinline__ void e5m2_roundtrip2(float& lo, float& hi) {
  uint16_t e;
  asm("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;" : "=h"(e) : "f"(hi), "f"(lo));
  uint32_t h2;
  asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h2) : "h"(e));
  const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
  lo = t.x;
  hi = t.y;
}
__device__ __forceinline__ void e3m2_roundtrip2(float& lo, float& hi) {
  uint16_t e;
  asm("cvt.rn.satfinite.e3m2x2.f32 %0, %1, %2;" : "=h"(e) : "f"(hi), "f"(lo));
  uint32_t h2;
  asm("cvt.rn.f16x2.e3m2x2 %0, %1;" : "=r"(h2) : "h"(e));
  const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
  lo = t.x;
  hi = t.y;
}
__device__ __forceinline__ void e2m3_roundtrip2(float& lo, float& hi) {
  uint16_t e;
  asm("cvt.rn.satfinite.e2m3x2.f32 %0, %1, %2;" : "=h"(e) : "f"(hi), "f"(lo));
  uint32_t h2;
  asm("cvt.rn.f16x2.e2m3x2 %0, %1;" : "=r"(h2) : "h"(e));
  const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
  lo = t.x;
  hi = t.y;
}

// E3M0 {0, .25, .5, 1, 2, 4, 8, 16}: ties AWAY from zero (tensor_quant_mx.h:84-93, :133-150)
__device__ __forceinline__ float e3m0_round(float a) {
  if (!(a < 12.f)) return 16.f;  // also inf / NaN
  return a < 0.125f ? 0.f : a < 0.375f ? 0.25f : a < 0.75f ? 0.5f : a < 1.5f ? 1.f : a < 3.f ? 2.f : a < 6.f ? 4.f : 8.f;
}

// Round a pair of NON-NEGATIVE magnitudes (inf / NaN allowed) to the element format; results as fp32
// (convert_to_types, tensor_quant_mx.h:152-175).
template <int FMT>
__device__ __forceinline__ void fmt_round2(float& a, float& b) {
  if constexpr (FMT == MO_FMT_E4M3) e4m3_roundtrip2(a, b);
  else if constexpr (FMT == MO_FMT_E5M2) e5m2_roundtrip2(a, b);
  else if constexpr (FMT == MO_FMT_E2M1) e2m1_roundtrip2(a, b);
  else if constexpr (FMT == MO_FMT_E3M2) e3m2_roundtrip2(a, b);
  else if constexpr (FMT == MO_FMT_E2M3) e2m3_roundtrip2(a, b);
  else if constexpr (FMT == MO_FMT_INT8) {  // rint, saturate at +-127; (int)NaN is 0 on the device
    a = a != a ? 0.f : fminf(rintf(a), 127.f);
    b = b != b ? 0.f : fminf(rintf(b), 127.f);
  } else if constexpr (FMT == MO_FMT_E0M3) {  // integers 0..7, ties to even; inf / NaN saturate
    a = fminf(rintf(a), 7.f);
    b = fminf(rintf(b), 7.f);
  } else if constexpr (FMT == MO_FMT_E1M2) {  // halves 0..3.5, ties to even
    a = fminf(__fmul_rn(rintf(__fmul_rn(a, 2.f)), 0.5f), 3.5f);
    b = fminf(__fmul_rn(rintf(__fmul_rn(b, 2.f)), 0.5f), 3.5f);
  } else {
    a = e3m0_round(a);
    b = e3m0_round(b);
  }
}

__device__ __forceinline__ float fmt_round_rt(float a, int fmt) {
  float z = 0.f;
  switch (fmt) {
    case MO_FMT_E4M3: fmt_round2<MO_FMT_E4M3>(a, z); break;
    case MO_FMT_E5M2: fmt_round2<MO_FMT_E5M2>(a, z); break;
    case MO_FMT_INT8: fmt_round2<MO_FMT_INT8>(a, z); break;
    case MO_FMT_E0M3: fmt_round2<MO_FMT_E0M3>(a, z); break;
    case MO_FMT_E1M2: fmt_round2<MO_FMT_E1M2>(a, z); break;
    case MO_FMT_E3M0: fmt_round2<MO_FMT_E3M0>(a, z); break;
    case MO_FMT_E2M1: fmt_round2<MO_FMT_E2M1>(a, z); break;
    case MO_FMT_E3M2: fmt_round2<MO_FMT_E3M2>(a, z); break;
    case MO_FMT_E2M3: fmt_round2<MO_FMT_E2M3>(a, z); break;
    default: a = 0.f;
  }
  return a;
}

// signed variant for scale formats (convert_to_types is called on a positive scale; sign kept for completeness)
__device__ __forceinline__ float fmt_round_signed_rt(float v, int fmt) {
  const float r = fmt_round_rt(fabsf(v), fmt);
  return v < 0.f ? -r : r;
}

// ---------------------------------------------------------------------------------------------------------------------
// Scales.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float div_approx_ftz(float a, float b) {
  float r;
  asm("div.approx.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float div_fast(float a, float b, bool fast) { return fast ? div_approx_ftz(a, b) : __fdiv_rn(a, b); }

// 2^e for e in [-149, 127] (ldexpf(1.f, e); below -126 the result is an fp32 denormal)
__device__ __forceinline__ float pow2i(int e) {
  return __uint_as_float(e >= -126 ? (uint32_t)(e + 127) << 23 : 1u << (e + 149));
}

// compute_scale_e8m0_NV (tensor_quant_mx.cu:104-132): unscale = 2^ceil(log2(amax / dmax)), on the bits of the quotient.
__device__ __forceinline__ int e8m0_unscale_exp(float amax, float dmax, bool fast) {
  const uint32_t u = __float_as_uint(div_fast(amax, dmax, fast));
  const int ef = (u >> 23) & 0xff;
  const uint32_t sf = u & 0x7fffffu;
  return (sf > 0 && ef != 0xfe && !(ef == 0 && sf <= 0x400000u)) ? ef - 126 : ef - 127;
}

struct BlockScale {
  float scale, unscale;
};

// compute_scale, E8M0 branch (tensor_quant_mx.cu:135-152)
__device__ __forceinline__ BlockScale mx_scale_e8m0(float amax, float dmax, bool fast) {
  if (amax == 0.f || isinf(amax) || isnan(amax)) return {1.f, 1.f};
  const int e = e8m0_unscale_exp(amax, dmax, fast);
  return {pow2i(-e), pow2i(e)};
}

// compute_scale general branch (:145-151) and compute_scale_with_global (:154-181): fp32 quotients (fast-math in the
// reference build), the rest in double exactly as written.
__device__ __forceinline__ BlockScale mx_scale_general(float amax, int fmt, int sfmt, const float* global_amax, bool fast) {
  const float emax = fmt_max(fmt);
  if (global_amax == nullptr) {
    if (amax == 0.f || isinf(amax) || isnan(amax)) return {1.f, 1.f};
    double scale = (double)div_fast(emax, amax, fast);
    const double inv_scale = (double)fmt_round_signed_rt((float)(1. / scale), sfmt);
    scale = 1. / inv_scale;
    return {(float)scale, (float)inv_scale};
  }
  const float ga = *global_amax;
  const bool bad_l = amax == 0.f || isinf(amax) || isnan(amax);
  const bool bad_g = ga == 0.f || isinf(ga) || isnan(ga);
  if (bad_l || bad_g) return {1.f, 1.f};
  const float smax = fmt_max(sfmt);
  const float local_unscale = div_fast(amax, emax, fast);
  const double two_level = (double)smax * (double)div_fast(emax, ga, fast);
  const double local_unscale_q = (double)fmt_round_signed_rt((float)((double)local_unscale * two_level), sfmt) / two_level;
  return {(float)(1. / local_unscale_q), (float)local_unscale_q};
}

// quantize() (tensor_quant_mx.cu:36-55) on a pair; the sign of a zero input is unspecified in the reference
// (uninitialised `sign`), here +0.
template <int FMT>
__device__ __forceinline__ void mx_qdq2(float& x0, float& x1, const BlockScale& s) {
  float a = __fmul_rn(fabsf(x0), s.scale), b = __fmul_rn(fabsf(x1), s.scale);
  fmt_round2<FMT>(a, b);
  a = __fmul_rn(a, s.unscale);
  b = __fmul_rn(b, s.unscale);
  x0 = x0 < 0.f ? -a : a;
  x1 = x1 < 0.f ? -b : b;
}

// ---------------------------------------------------------------------------------------------------------------------
// Vector path: E8M0 scales, cols % BLOCK == 0, 32-byte aligned.  One thread = one block = NV vectors.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kMxUnroll = 2;

template <typename T, int BLOCK, int FMT>
__global__ void __launch_bounds__(kThreads) mx_e8m0_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t n_blocks,
                                                           int fast) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = BLOCK / E;
  static_assert(NV >= 1 && BLOCK % E == 0, "block must be a whole number of 32-byte vectors");
  const V32* xv = reinterpret_cast<const V32*>(x);
  V32* yv = reinterpret_cast<V32*>(y);
  const float dmax = fmt_max(FMT);
  const int64_t tile_blocks = (int64_t)kThreads * kMxUnroll;
  const int64_t n_tiles = (n_blocks + tile_blocks - 1) / tile_blocks;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t base = tile * tile_blocks + threadIdx.x;
    V32 v[kMxUnroll][NV];
#pragma unroll
    for (int u = 0; u < kMxUnroll; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
#pragma unroll
        for (int k = 0; k < NV; ++k) v[u][k] = ld_stream(xv + b * NV + k);
      }
    }
#pragma unroll
    for (int u = 0; u < kMxUnroll; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
        float f[BLOCK];
#pragma unroll
        for (int k = 0; k < NV; ++k) Elem<T>::unpack(v[u][k], f + k * E);
        float t[BLOCK];  // balanced max tree (max is associative): no 31-deep dependent chain per block
#pragma unroll
        for (int i = 0; i < BLOCK; ++i) t[i] = fabsf(f[i]);
#pragma unroll
        for (int w = BLOCK / 2; w >= 1; w /= 2) {
#pragma unroll
          for (int i = 0; i < w; ++i) t[i] = fmaxf(t[i], t[i + w]);
        }
        const float m = t[0];
        const BlockScale s = mx_scale_e8m0(m, dmax, fast != 0);
#pragma unroll
        for (int i = 0; i < BLOCK; i += 2) mx_qdq2<FMT>(f[i], f[i + 1], s);
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          Elem<T>::pack(f + k * E, v[u][k]);
          st_stream(yv + b * NV + k, v[u][k]);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Generic path: any element / scale format, optional global amax, block 8 / 16 / 32 (or any), ragged last dim (the
// short block is zero-padded like the reference's in_padding threads), unaligned pointers.  One thread per block.
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) mx_generic_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows,
                                                              int64_t cols, int block, int fmt, int sfmt,
                                                              const float* __restrict__ global_amax, int fast) {
  const int64_t bpr = (cols + block - 1) / block;
  const int64_t total = rows * bpr;
  for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kThreads) {
    const int64_t r = t / bpr, c0 = (t % bpr) * block;
    const int len = (int)min((int64_t)block, cols - c0);
    const T* xp = x + r * cols + c0;
    T* yp = y + r * cols + c0;
    float m = 0.f;  // padding zeros take part; fmaxf drops NaN like the reference's reduce
    for (int i = 0; i < len; ++i) m = fmaxf(m, fabsf(Elem<T>::load(xp + i)));
    const BlockScale s = sfmt == MO_FMT_E8M0 ? mx_scale_e8m0(m, fmt_max(fmt), fast != 0)
                                             : mx_scale_general(m, fmt, sfmt, global_amax, fast != 0);
    for (int i = 0; i < len; ++i) {
      const float v = Elem<T>::load(xp + i);
      float a = fmt_round_rt(__fmul_rn(fabsf(v), s.scale), fmt);
      a = __fmul_rn(a, s.unscale);
      Elem<T>::store(yp + i, v < 0.f ? -a : a);
    }
  }
}

template <typename T, int BLOCK>
static bool launch_mx_vec(const T* x, T* y, int64_t nb, int fmt, int fast, cudaStream_t s) {
  const int64_t tiles = (nb + (int64_t)kThreads * kMxUnroll - 1) / ((int64_t)kThreads * kMxUnroll);
#define MO_MX(F)                                                                                              \
  case F:                                                                                                     \
    mx_e8m0_kernel<T, BLOCK, F><<<resident_grid((const void*)mx_e8m0_kernel<T, BLOCK, F>, tiles), kThreads, 0, s>>>( \
        x, y, nb, fast);                                                                                      \
    return true;
  switch (fmt) {
    MO_MX(MO_FMT_E4M3)
    MO_MX(MO_FMT_E2M1)
    MO_MX(MO_FMT_E3M2)
    MO_MX(MO_FMT_E2M3)
    MO_MX(MO_FMT_INT8)
    MO_MX(MO_FMT_E5M2)
  }
#undef MO_MX
  return false;
}

template <typename T>
static int launch_mx(const void* xv, void* yv, int64_t rows, int64_t cols, int block, int fmt, int sfmt,
                     const float* ga, int fast, cudaStream_t s) {
  const T* x = static_cast<const T*>(xv);
  T* y = static_cast<T*>(yv);
  constexpr int E = Elem<T>::kPerVec;
  if (sfmt == MO_FMT_E8M0 && cols % block == 0 && aligned32(xv) && aligned32(yv) && block % E == 0) {
    const int64_t nb = rows * (cols / block);
    bool done = false;
    if (block == 16) done = launch_mx_vec<T, 16>(x, y, nb, fmt, fast, s);
    else if (block == 32) done = launch_mx_vec<T, 32>(x, y, nb, fmt, fast, s);
    if (done) return check_launch("mo_qdq_mx");
  }
  const int64_t total = rows * ((cols + block - 1) / block);
  const int grid = persistent_grid((total + kThreads - 1) / kThreads, 8);
  mx_generic_kernel<T><<<grid, kThreads, 0, s>>>(x, y, rows, cols, block, fmt, sfmt, ga, fast);
  return check_launch("mo_qdq_mx(generic)");
}

// ---------------------------------------------------------------------------------------------------------------------
// MXFP8 / MXFP4 packers (pure-torch in the reference).  E8M0 byte = ceil(log2(amax / fmt_max)) + 127 taken in fp32.
// ---------------------------------------------------------------------------------------------------------------------
// mxfp8_tensor.py:41-66 (`_compute_e8m0_exponent`): descale>0 ? log2 : -127, ceil, clamp [-127,127]
// mxfp4_tensor.py:74-78: ceil(max(log2(descale), -127)) (no upper clamp; amax is finite there)
__device__ __forceinline__ int e8m0_byte_torch(float amax, float dmax, bool gpu_div) {
  // `tensor / python_scalar` multiplies by the reciprocal on a CUDA tensor and divides on a CPU tensor
  const float descale = gpu_div ? __fmul_rn(amax, __frcp_rn(dmax)) : __fdiv_rn(amax, dmax);
  float l = descale > 0.f ? log2f(descale) : -127.f;
  l = ceilf(fmaxf(l, -127.f));
  l = fminf(l, 127.f);
  return (int)l + 127;
}

// One thread per 32-element block: q = e4m3(clamp(x * 2^(127 - sf), +-448)) (mxfp8_tensor.py:141-189).
template <typename T>
__global__ void __launch_bounds__(kThreads) mxfp8_pack_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                              uint8_t* __restrict__ sf, int64_t rows, int64_t cols,
                                                              int given, int gpu_div, int vec) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = 32 / E;
  const int64_t bpr = (cols + 31) / 32;
  const int64_t total = rows * bpr;
  for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kThreads) {
    const int64_t r = t / bpr, c0 = (t % bpr) * 32;
    const int len = (int)min((int64_t)32, cols - c0);
    float f[32];
    if (vec) {
      const V32* xv = reinterpret_cast<const V32*>(x + r * cols + c0);
#pragma unroll
      for (int k = 0; k < NV; ++k) Elem<T>::unpack(ld_stream(xv + k), f + k * E);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) f[i] = i < len ? Elem<T>::load(x + r * cols + c0 + i) : 0.f;
    }
    int byte;
    if (given) byte = sf[t];
    else {
      float m = fabsf(f[0]);
#pragma unroll
      for (int i = 1; i < 32; ++i) m = fmaxf(m, fabsf(f[i]));
      byte = e8m0_byte_torch(m, 448.f, gpu_div != 0);
      sf[t] = (uint8_t)byte;
    }
    const float mul = pow2i(127 - byte);  // exp2(127 - sf)
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = fminf(fmaxf(__fmul_rn(f[4 * i + j], mul), -448.f), 448.f);
      uint16_t lo, hi;
      asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(v[1]), "f"(v[0]));
      asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(v[3]), "f"(v[2]));
      w[i] = (uint32_t)lo | ((uint32_t)hi << 16);
    }
    if (q == nullptr) continue;  // scales only (get_weights_scaling_factor)
    uint8_t* qp = q + r * cols + c0;
    if (vec) {
      V32 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) o.w[i] = w[i];
      st_stream(qp, o);
    } else {
      for (int i = 0; i < len; ++i) qp[i] = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
    }
  }
}

// One thread per block: codes = sign<<3 | #{bounds < |x / 2^e|}, two per byte, even element low
// (mxfp4_tensor.py:52-83).  Quirks kept: a zero element gets the sign bit ((2 - sign(0)) // 2 == 1), ties round DOWN.
__device__ __forceinline__ uint32_t mxfp4_code(float v) {
  const float a = fabsf(v);
  const uint32_t ord = (a > 0.25f) + (a > 0.75f) + (a > 1.25f) + (a > 1.75f) + (a > 2.5f) + (a > 3.5f) + (a > 5.f);
  return (v > 0.f ? 0u : 8u) | ord;
}

__device__ __forceinline__ float mxfp4_exponent(float amax, int gpu_div) {
  const float descale = gpu_div ? __fmul_rn(amax, __frcp_rn(6.f)) : __fdiv_rn(amax, 6.f);
  return ceilf(fmaxf(log2f(descale), -127.f));  // log2(0) = -inf -> -127
}

template <typename T>
__global__ void __launch_bounds__(kThreads) mxfp4_pack_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                              uint8_t* __restrict__ sf, int64_t n_blocks, int block,
                                                              int gpu_div) {
  for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < n_blocks; t += (int64_t)gridDim.x * kThreads) {
    const T* xp = x + t * block;
    float m = 0.f;
    for (int i = 0; i < block; ++i) m = fmaxf(m, fabsf(Elem<T>::load(xp + i)));
    const float e = mxfp4_exponent(m, gpu_div);
    const float div = exp2f(e);  // exact: integer exponent
    sf[t] = (uint8_t)(int)(e + 127.f);
    for (int i = 0; i < block; i += 2) {
      const uint32_t c0 = mxfp4_code(__fdiv_rn(Elem<T>::load(xp + i), div));
      const uint32_t c1 = mxfp4_code(__fdiv_rn(Elem<T>::load(xp + i + 1), div));
      q[(t * block + i) >> 1] = (uint8_t)(c0 | (c1 << 4));
    }
  }
}

// block == 32, aligned: the block lives in registers (NV 32-byte vectors), x / 2^e is the exact product x * 2^-e
// (e in [-127, 126]), the 16 code bytes leave as one 16-byte store.
template <typename T>
__global__ void __launch_bounds__(kThreads) mxfp4_pack32_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                                uint8_t* __restrict__ sf, int64_t n_blocks, int gpu_div) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = 32 / E;
  const V32* xv = reinterpret_cast<const V32*>(x);
  for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < n_blocks; t += (int64_t)gridDim.x * kThreads) {
    float f[32];
    V32 v[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = ld_stream(xv + t * NV + k);
#pragma unroll
    for (int k = 0; k < NV; ++k) Elem<T>::unpack(v[k], f + k * E);
    float m = fabsf(f[0]);
#pragma unroll
    for (int i = 1; i < 32; ++i) m = fmaxf(m, fabsf(f[i]));
    const int e = (int)mxfp4_exponent(m, gpu_div);
    sf[t] = (uint8_t)(e + 127);
    const float mul = pow2i(-e);
    // #{bounds < a} == the hardware's round-to-nearest-even E2M1 index, minus one at the three ties RNE sends up
    // (0.75, 1.75, 3.5); one F2FP converts a pair.  (NaN elements: unspecified in the reference too.)
    uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      const float v0 = __fmul_rn(f[i], mul), v1 = __fmul_rn(f[i + 1], mul);
      const float a = fabsf(v0), b = fabsf(v1);
      uint32_t byte = e2m1_code2(a, b);
      byte -= (a == 0.75f || a == 1.75f || a == 3.5f) ? 1u : 0u;
      byte -= (b == 0.75f || b == 1.75f || b == 3.5f) ? 16u : 0u;
      byte |= (v0 > 0.f ? 0u : 8u) | (v1 > 0.f ? 0u : 0x80u);
      w[i / 8] |= byte << (8 * ((i % 8) / 2));
    }
    reinterpret_cast<uint4*>(q)[t] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// Dequantize: e4m3 codes (kind 0) or packed E2M1 nibbles (kind 1) times 2^(sf - 127), cast to T.
// vec: a thread turns E codes into one 32-byte vector of outputs (cols % E == 0, block % E == 0).
template <typename T>
__global__ void __launch_bounds__(kThreads) mx_unpack_kernel(const uint8_t* __restrict__ q, const uint8_t* __restrict__ sf,
                                                             T* __restrict__ y, int64_t rows, int64_t cols, int block,
                                                             int kind, int vec) {
  constexpr int E = Elem<T>::kPerVec;
  const int64_t bpr = (cols + block - 1) / block;
  const int64_t n = rows * cols;
  if (vec) {
    const int64_t nv = n / E;
    const int64_t vpr = cols / E;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      const int64_t r = i / vpr, c = (i - r * vpr) * E;
      const float mul = pow2i((int)sf[r * bpr + c / block] - 127);
      float f[E];
      if (kind == 0) {
        uint32_t w[E / 4];
        if constexpr (E == 16) {
          const uint4 u = __ldg(reinterpret_cast<const uint4*>(q) + i);
          w[0] = u.x; w[1] = u.y; w[2] = u.z; w[3] = u.w;
        } else {
          const uint2 u = __ldg(reinterpret_cast<const uint2*>(q) + i);
          w[0] = u.x; w[1] = u.y;
        }
#pragma unroll
        for (int k = 0; k < E; k += 2) {
          uint32_t h2;
          asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)((w[k / 4] >> (8 * (k % 4))) & 0xFFFFu)));
          const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
          f[k] = __fmul_rn(t.x, mul);
          f[k + 1] = __fmul_rn(t.y, mul);
        }
      } else {
        uint32_t w[E / 8];
        if constexpr (E == 16) {
          const uint2 u = __ldg(reinterpret_cast<const uint2*>(q) + i);
          w[0] = u.x; w[1] = u.y;
        } else {
          w[0] = __ldg(reinterpret_cast<const uint32_t*>(q) + i);
        }
#pragma unroll
        for (int k = 0; k < E; k += 2) {
          uint32_t h2;
          asm("{ .reg .b8 t; cvt.u8.u32 t, %1; cvt.rn.f16x2.e2m1x2 %0, t; }" : "=r"(h2) : "r"((w[k / 8] >> (4 * (k % 8))) & 0xFFu));
          const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
          f[k] = __fmul_rn(t.x, mul);
          f[k + 1] = __fmul_rn(t.y, mul);
        }
      }
      V32 o;
      Elem<T>::pack(f, o);
      st_stream(reinterpret_cast<V32*>(y) + i, o);
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    const int64_t r = i / cols, c = i % cols;
    const float mul = pow2i((int)sf[r * bpr + c / block] - 127);
    float v;
    if (kind == 0) {
      uint32_t h2;
      asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)q[i]));
      v = __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
    } else {
      const uint32_t nib = (q[i >> 1] >> (4 * (i & 1))) & 0xf;
      const float mag = (nib & 7) < 4 ? 0.5f * (float)(nib & 7) : (float)(2 << (((nib & 7) - 4) >> 1)) * (((nib & 7) & 1) ? 1.5f : 1.f);
      v = (nib & 8) ? -mag : mag;
    }
    Elem<T>::store(y + i, __fmul_rn(v, mul));
  }
}

}  // namespace mo

using namespace mo;

extern "C" int mo_qdq_mx(const void* x, void* y, int64_t rows, int64_t cols, int dtype, int block, int format,
                         int scale_format, const float* global_amax, int numerics, void* stream) {
  if (rows == 0 || cols == 0) return MO_OK;
  if (!x || !y) { set_error("mo_qdq_mx: NULL pointer"); return MO_ERR_NULL; }
  if (rows < 0 || cols < 0) { set_error("mo_qdq_mx: negative size"); return MO_ERR_SHAPE; }
  if (block != 8 && block != 16 && block != 32) {
    set_error("mo_qdq_mx: block=%d not in {8,16,32} (tensor_quant_mx.cu:317-325)", block);
    return MO_ERR_ARG;
  }
  if (format < 0 || format >= MO_FMT_E8M0 || scale_format < 0 || scale_format > MO_FMT_E8M0) {
    set_error("mo_qdq_mx: bad format %d / scale format %d", format, scale_format);
    return MO_ERR_ARG;
  }
  if (scale_format == MO_FMT_E8M0) global_amax = nullptr;  // compute_scale ignores it for E8M0 (tensor_quant.py:171)
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int fast = numerics != MO_DIV_IEEE;
  switch (dtype) {
    case MO_F32: return launch_mx<float>(x, y, rows, cols, block, format, scale_format, global_amax, fast, s);
    case MO_F16: return launch_mx<__half>(x, y, rows, cols, block, format, scale_format, global_amax, fast, s);
    case MO_BF16: return launch_mx<__nv_bfloat16>(x, y, rows, cols, block, format, scale_format, global_amax, fast, s);
  }
  set_error("mo_qdq_mx: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_mxfp8_pack(const void* x, uint8_t* q, uint8_t* sf, int64_t rows, int64_t cols, int dtype,
                             int scale_given, int numerics, void* stream) {
  if (rows == 0 || cols == 0) return MO_OK;
  if (!x || !sf || (!q && scale_given)) { set_error("mo_mxfp8_pack: NULL pointer"); return MO_ERR_NULL; }
  if (rows < 0 || cols < 0) { set_error("mo_mxfp8_pack: negative size"); return MO_ERR_SHAPE; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t total = rows * ((cols + 31) / 32);
  const int vec = cols % 32 == 0 && aligned32(x) && (q == nullptr || aligned32(q));
  const int gpu_div = numerics != MO_DIV_IEEE;
#define MO_P8(T)                                                                                                  \
  mxfp8_pack_kernel<T><<<resident_grid((const void*)mxfp8_pack_kernel<T>, (total + kThreads - 1) / kThreads), kThreads, 0, \
                         s>>>(static_cast<const T*>(x), q, sf, rows, cols, scale_given, gpu_div, vec)
  switch (dtype) {
    case MO_F32: MO_P8(float); break;
    case MO_F16: MO_P8(__half); break;
    case MO_BF16: MO_P8(__nv_bfloat16); break;
    default: set_error("mo_mxfp8_pack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
#undef MO_P8
  return check_launch("mo_mxfp8_pack");
}

extern "C" int mo_mxfp4_pack(const void* x, uint8_t* q, uint8_t* sf, int64_t n_blocks, int block, int dtype,
                             int numerics, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !q || !sf) { set_error("mo_mxfp4_pack: NULL pointer"); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 2 || (block & 1)) { set_error("mo_mxfp4_pack: bad n_blocks / block"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = persistent_grid((n_blocks + kThreads - 1) / kThreads, 8);
  const int gpu_div = numerics != MO_DIV_IEEE;
  if (block == 32 && aligned32(x) && (reinterpret_cast<uintptr_t>(q) & 15u) == 0) {
    switch (dtype) {
      case MO_F32: mxfp4_pack32_kernel<float><<<grid, kThreads, 0, s>>>(static_cast<const float*>(x), q, sf, n_blocks, gpu_div); break;
      case MO_F16: mxfp4_pack32_kernel<__half><<<grid, kThreads, 0, s>>>(static_cast<const __half*>(x), q, sf, n_blocks, gpu_div); break;
      case MO_BF16: mxfp4_pack32_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(static_cast<const __nv_bfloat16*>(x), q, sf, n_blocks, gpu_div); break;
      default: set_error("mo_mxfp4_pack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
    }
    return check_launch("mo_mxfp4_pack");
  }
  switch (dtype) {
    case MO_F32: mxfp4_pack_kernel<float><<<grid, kThreads, 0, s>>>(static_cast<const float*>(x), q, sf, n_blocks, block, gpu_div); break;
    case MO_F16: mxfp4_pack_kernel<__half><<<grid, kThreads, 0, s>>>(static_cast<const __half*>(x), q, sf, n_blocks, block, gpu_div); break;
    case MO_BF16: mxfp4_pack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(static_cast<const __nv_bfloat16*>(x), q, sf, n_blocks, block, gpu_div); break;
    default: set_error("mo_mxfp4_pack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_mxfp4_pack");
}

extern "C" int mo_mx_unpack(const uint8_t* q, const uint8_t* sf, void* y, int64_t rows, int64_t cols, int block,
                            int kind, int out_dtype, void* stream) {
  if (rows == 0 || cols == 0) return MO_OK;
  if (!q || !sf || !y) { set_error("mo_mx_unpack: NULL pointer"); return MO_ERR_NULL; }
  if (rows < 0 || cols < 0 || block < 1 || (kind != 0 && kind != 1)) { set_error("mo_mx_unpack: bad argument"); return MO_ERR_ARG; }
  if (kind == 1 && (cols & 1)) { set_error("mo_mx_unpack: packed E2M1 needs an even last dim"); return MO_ERR_SHAPE; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int ev = out_dtype == MO_F32 ? 8 : 16;
  const int vec = cols % ev == 0 && block % ev == 0 && aligned32(y) && (reinterpret_cast<uintptr_t>(q) & 15u) == 0;
  const int grid = persistent_grid(((vec ? rows * cols / ev : rows * cols) + kThreads - 1) / kThreads, 16);
  switch (out_dtype) {
    case MO_F32: mx_unpack_kernel<float><<<grid, kThreads, 0, s>>>(q, sf, static_cast<float*>(y), rows, cols, block, kind, vec); break;
    case MO_F16: mx_unpack_kernel<__half><<<grid, kThreads, 0, s>>>(q, sf, static_cast<__half*>(y), rows, cols, block, kind, vec); break;
    case MO_BF16: mx_unpack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(q, sf, static_cast<__nv_bfloat16*>(y), rows, cols, block, kind, vec); break;
    default: set_error("mo_mx_unpack: unsupported dtype %d", out_dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_mx_unpack");
}
