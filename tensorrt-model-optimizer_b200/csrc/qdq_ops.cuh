// Per-element fake-quant operators shared by the streaming kernels (qdq_ew.cu) and the GEMM prologue (gemm.cu).
#pragma once

#include "common.cuh"

namespace mo {

// How amax is indexed.
enum AxisMode { kPerTensor = 0, kPerChunk = 1, kPerElem = 2 };

// n / d for every 32-bit n with one mul.hi and two shifts (Granlund-Montgomery round-up form); d >= 1.
struct FastDiv {
  uint32_t d, mul, shift;  // shift = ceil(log2 d) - 1; d == 1 is handled apart
  __host__ void init(uint32_t d_) {
    d = d_;
    uint32_t l = 0;
    while ((uint64_t(1) << l) < d_) ++l;
    mul = (uint32_t)(((uint64_t(1) << 32) * ((uint64_t(1) << l) - d_)) / d_ + 1);
    shift = l == 0 ? 0 : l - 1;
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const {
    if (d == 1) return n;
    const uint32_t t = __umulhi(n, mul);
    return (t + ((n - t) >> 1)) >> shift;
  }
};

struct AxisArgs {
  const float* amax;
  int64_t axis_size;
  int64_t inner;
  int mode;
  bool small;  // n, inner and axis_size all fit 32 bits: FastDiv path
  FastDiv div_inner, div_axis;
  // channel of element idx
  __device__ __forceinline__ int64_t channel(int64_t idx) const {
    if (small) {
      const uint32_t row = div_inner.div((uint32_t)idx);
      return row - div_axis.div(row) * (uint32_t)axis_size;
    }
    return (idx / inner) % axis_size;
  }
};

// ------------------------------------------------------------------------------------------------
// FP8: y = T( float(e4m3_rn_satfinite(float(x) * scale)) * inv_scale )
// ------------------------------------------------------------------------------------------------
struct Fp8Op {
  AxisArgs ax;
  bool plain;  // amax == NULL: cast only (torch .to(float8_e4m3fn): |x| > 464 or NaN -> NaN)
  bool eager;  // MO_NUMERICS_CPU: Python `448.0 / amax` == reciprocal(amax) * 448 (tensor_quant.py:52-53)
  float scale, inv;

  __device__ __forceinline__ void set(float amax) {
    const float safe = amax <= kEps ? 1.0f : amax;   // tensor_quant_gpu_fp8.cu:94-96
    scale = eager ? __fmul_rn(__fdiv_rn(1.0f, safe), 448.0f) : __fdiv_rn(448.0f, safe);  // :97
    inv = __fdiv_rn(1.0f, scale);                     // :98
  }
  __device__ __forceinline__ void init() {
    scale = inv = 1.0f;
    if (!plain && ax.mode == kPerTensor) set(__ldg(ax.amax));
  }
  __device__ __forceinline__ void pair(float& a, float& b) const {
    float p = __fmul_rn(a, scale), q = __fmul_rn(b, scale);
    e4m3_roundtrip2(p, q);
    if (plain) {
      if (!(fabsf(a) <= 464.0f)) p = __int_as_float(0x7fc00000);
      if (!(fabsf(b) <= 464.0f)) q = __int_as_float(0x7fc00000);
    }
    a = __fmul_rn(p, inv);
    b = __fmul_rn(q, inv);
  }
  // am: the chunk's amax when the caller prefetched it with the data (kPerChunk), else ignored
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0, float am, bool have_am) {
    if (ax.mode == kPerChunk) set(have_am ? am : __ldg(ax.amax + ax.channel(idx0)));
    if (ax.mode == kPerElem) {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = one(f[i], idx0 + i);
      return;
    }
#pragma unroll
    for (int i = 0; i < N; i += 2) pair(f[i], f[i + 1]);
  }
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0) { vec<N>(f, idx0, 0.f, false); }
  __device__ __forceinline__ float one(float v, int64_t idx) {
    if (ax.mode != kPerTensor) set(__ldg(ax.amax + ax.channel(idx)));
    float z = 0.f;
    pair(v, z);
    return v;
  }
};

// ------------------------------------------------------------------------------------------------
// INT-N: q = clamp(rint(x * s), min, max), y = T(q / s), s = bound / amax; amax < 2^-24 -> 0.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float min_nan(float a, float b) {
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// kPT: per-tensor specialisation (its own kernel instantiation): the vector loop uses packed fp32 arithmetic.  Keeping both
// forms in one kernel costs 6 registers and an occupancy step (4 -> 3 CTAs per SM) that the per-channel / per-block modes,
// which also derive a scale per vector, pay 6 % for.
template <bool kUnsigned, bool kPT = false>
struct IntOp {
  AxisArgs ax;
  float minb, maxb;
  uint32_t* status;
  // per-scale state
  float scale;
  bool zero;
  float db, dr;
  bool dfast;
  uint32_t flags;

  __device__ __forceinline__ void set(float amax) {
    if (amax < 0.f) flags |= MO_FLAG_NEGATIVE_AMAX;
    zero = amax < kEps;                       // tensor_quant_gpu.cu:46 (strict <)
    scale = __fdiv_rn(maxb, amax);            // :50
    const ExactDiv d(scale);
    db = d.b; dr = d.r; dfast = d.fast;
  }
  __device__ __forceinline__ void init() {
    flags = 0;
    if (ax.mode == kPerTensor) set(__ldg(ax.amax));
  }
  template <bool kFast>
  __device__ __forceinline__ float q1t(float x) {
    if (kUnsigned && x < 0.f) flags |= MO_FLAG_NEGATIVE_IN_UNSIGNED;
    float v = __fmul_rn(x, scale);
    // clamp first (bounds are integers, so clamp and rint commute), NaN kept like the reference's
    // `output > max ? max : output` selects (:52-53)
    v = max_nan(min_nan(v, maxb), minb);
    // rint for |v| < 2^22: add/subtract 1.5*2^23 (round-half-even in the fp32 adder)
    v = __fsub_rn(__fadd_rn(v, 12582912.0f), 12582912.0f);
    float o;
    if constexpr (kFast) {
      const float q0 = __fmul_rn(v, dr);
      const float rem = __fmaf_rn(-db, q0, v);
      o = __fmaf_rn(dr, rem, q0);
    } else {
      o = __fdiv_rn(v, scale);                 // :55
    }
    // the sign of the result always equals the sign of x (restores -0.0 lost by the two tricks)
    o = __uint_as_float((__float_as_uint(o) & 0x7fffffffu) | (__float_as_uint(x) & 0x80000000u));
    return zero ? 0.0f : o;
  }
  __device__ __forceinline__ float q1(float x) { return dfast ? q1t<true>(x) : q1t<false>(x); }
  // q1t<true> on a pair with packed fp32 arithmetic (FMUL2 / FADD2 / FFMA2: one issue slot for two elements); the same
  // IEEE operations in the same order per element, so the same bits
  __device__ __forceinline__ void q2fast(float& x0, float& x1) {
    if (kUnsigned && (x0 < 0.f || x1 < 0.f)) flags |= MO_FLAG_NEGATIVE_IN_UNSIGNED;
    float v0, v1;
    asm("{ .reg .b64 a, s, t; mov.b64 a, {%2, %3}; mov.b64 s, {%4, %4}; mul.rn.f32x2 t, a, s; mov.b64 {%0, %1}, t; }"
        : "=f"(v0), "=f"(v1) : "f"(x0), "f"(x1), "f"(scale));
    v0 = max_nan(min_nan(v0, maxb), minb);
    v1 = max_nan(min_nan(v1, maxb), minb);
    float o0, o1;
    asm("{ .reg .b64 v, m, nm, r, nb, q, e;\n"
        "  mov.b64 v, {%2, %3}; mov.b64 m, {%4, %4}; mov.b64 nm, {%5, %5}; mov.b64 r, {%6, %6}; mov.b64 nb, {%7, %7};\n"
        "  add.rn.f32x2 v, v, m; add.rn.f32x2 v, v, nm;\n"          // rint: (v + 1.5 * 2^23) - 1.5 * 2^23
        "  mul.rn.f32x2 q, v, r; fma.rn.f32x2 e, nb, q, v; fma.rn.f32x2 q, r, e, q;\n"  // v / scale (ExactDiv)
        "  mov.b64 {%0, %1}, q; }"
        : "=f"(o0), "=f"(o1)
        : "f"(v0), "f"(v1), "f"(12582912.0f), "f"(-12582912.0f), "f"(dr), "f"(-db));
    o0 = __uint_as_float((__float_as_uint(o0) & 0x7fffffffu) | (__float_as_uint(x0) & 0x80000000u));
    o1 = __uint_as_float((__float_as_uint(o1) & 0x7fffffffu) | (__float_as_uint(x1) & 0x80000000u));
    x0 = zero ? 0.0f : o0;
    x1 = zero ? 0.0f : o1;
  }
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0) { vec<N>(f, idx0, 0.f, false); }
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0, float am, bool have_am) {
    if (ax.mode == kPerChunk) set(have_am ? am : __ldg(ax.amax + ax.channel(idx0)));
    if (ax.mode == kPerElem) {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = one(f[i], idx0 + i);
      return;
    }
    // one branch per vector, not per element: the exact-division window test is uniform over the vector
    if (kPT && dfast) {
      // packed fp32 (measured: per-tensor INT8 87.5 -> 94.6 % of the copy peak)
#pragma unroll
      for (int i = 0; i < N; i += 2) q2fast(f[i], f[i + 1]);
    } else if (dfast) {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = q1t<true>(f[i]);
    } else {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = q1t<false>(f[i]);
    }
  }
  __device__ __forceinline__ float one(float v, int64_t idx) {
    if (ax.mode != kPerTensor) set(__ldg(ax.amax + ax.channel(idx)));
    return q1(v);
  }
  __device__ __forceinline__ void finish() {
    if (status != nullptr && flags != 0) atomicOr(status, flags);
  }
};

__device__ __forceinline__ void finish_op(Fp8Op&) {}
template <bool U, bool P>
__device__ __forceinline__ void finish_op(IntOp<U, P>& op) { op.finish(); }

}  // namespace mo
