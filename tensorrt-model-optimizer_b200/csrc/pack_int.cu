// Real-quant packers for the integer / FP8 / NF4 weight formats.
//
// Replaces INT4QTensor / NF4QTensor (qtensor/int4_tensor.py:39-130, qtensor/nf4_tensor.py:75-199 and their CUDA kernels
// gemm/tensor_quant_gpu.cu:142-366: one thread per BYTE, one CTA per block, default stream) and the eager
// INT8QTensor / FP8QTensor (qtensor/int8_tensor.py:35-124, qtensor/fp8_tensor.py:38-151: amax, divide, expand,
// round, clamp, cast as separate passes).  HBM-bound: 2.5 B per bf16 element for the 4-bit packers, 3 B for the
// 8-bit ones.
//
// B200 shape: a thread owns one 32-byte vector; the lanes that share a quantization block (block / 16 lanes for bf16)
// reduce the block amax with xor-shuffles, every lane derives the scale with the reference's own rounding steps and
// converts its elements; eight 4-bit codes leave as one 8-byte store.  The 4-bit packers compute the per-block scale in
// the same pass (the reference: reduce_amax + reciprocal + multiply, then the pack kernel).
//
// Numerics.  Arithmetic in the tensor dtype T is reproduced as "fp32 op, round to T" (what torch does for bf16 / fp16).
// MO_NUMERICS_CPU follows the eager formulas; MO_NUMERICS_GPU follows the reference's CUDA kernels where they exist
// (INT4 / NF4; the reference re-defines AT_DISPATCH_FLOATING_TYPES to include Half and BFloat16,
// tensor_quant_gpu.cu:28-36, so the kernels run in at::Half / at::BFloat16 arithmetic).
#include <math.h>

#include "common.cuh"
#include "qdq_ops.cuh"

namespace mo {

template <typename T>
__device__ __forceinline__ float round_to(float v);
template <>
__device__ __forceinline__ float round_to<float>(float v) { return v; }
template <>
__device__ __forceinline__ float round_to<__nv_bfloat16>(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
template <>
__device__ __forceinline__ float round_to<__half>(float v) { return __half2float(__float2half_rn(v)); }

// Round a pair to T and back with ONE packed conversion (F2FP) plus two integer ops: the scalar cvt.rn.bf16.f32 /
// cvt.f32.bf16 pair (F2F) runs on the quarter-rate conversion pipe and was what bounded these kernels.
template <typename T>
__device__ __forceinline__ void round_to2(float& a, float& b) {
  if constexpr (std::is_same<T, __nv_bfloat16>::value) {
    uint32_t d;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
    a = __uint_as_float(d << 16);
    b = __uint_as_float(d & 0xffff0000u);
  } else if constexpr (std::is_same<T, __half>::value) {
    uint32_t d;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
    const float2 t = __half22float2(*reinterpret_cast<__half2*>(&d));
    a = t.x;
    b = t.y;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// INT4 (AWQ) packer: scale = T(T(1/amax) * 7)  [`7 / tensor` is reciprocal-then-multiply in torch], two codes per byte,
// EVEN element in the HIGH nibble (int4_tensor.py:74-83).
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float int4_scale(float amax) {
  return round_to<T>(__fmul_rn(round_to<T>(__fdiv_rn(1.0f, amax)), 7.0f));
}

// eager: round(T(x * s)) clamp [-8, 7] + 8 (NaN -> 0);  CUDA kernel: roundf(T(clamp(T(x * s), -8, 7) + 8)) & 0xF
template <typename T, bool kKernelForm>
__device__ __forceinline__ uint32_t int4_code(float x, float s) {
  if constexpr (kKernelForm) {  // every step in T (at::Half / at::BFloat16 operators round after each op)
    float v = round_to<T>(__fmul_rn(x, s));
    v = fmaxf(-8.f, fminf(7.f, v));
    const float t = round_to<T>(__fadd_rn(v, 8.f));  // in [0, 15]
    // roundf (half away from zero) of a non-negative t: for a half type t + 0.5 is exact in fp32, so floor(t + 0.5)
    if constexpr (sizeof(T) == 2) return (uint32_t)__float2int_rd(__fadd_rn(t, 0.5f));
    else return (uint32_t)(int)roundf(t) & 0xFu;
  } else {
    float v = round_to<T>(__fmul_rn(x, s));
    if (v != v) return 0u;
    v = fminf(fmaxf(rintf(v), -8.f), 7.f);
    return (uint32_t)((int)v + 8);
  }
}

// byte = code(x0) << 4 | code(x1), the packed-conversion form of int4_code (same results)
template <typename T, bool kKernelForm>
__device__ __forceinline__ uint32_t int4_byte(float x0, float x1, float s) {
  if constexpr (sizeof(T) == 2) {
    float a = __fmul_rn(x0, s), b = __fmul_rn(x1, s);
    round_to2<T>(a, b);
    if constexpr (kKernelForm) {
      a = __fadd_rn(fmaxf(-8.f, fminf(7.f, a)), 8.f);
      b = __fadd_rn(fmaxf(-8.f, fminf(7.f, b)), 8.f);
      round_to2<T>(a, b);  // in [0, 15]
      // roundf = floor(t + 0.5) (exact sum for a half type); floor by a round-toward-zero add of 2^23
      const uint32_t c0 = __float_as_uint(__fadd_rz(__fadd_rn(a, 0.5f), 8388608.f)) & 0xFu;
      const uint32_t c1 = __float_as_uint(__fadd_rz(__fadd_rn(b, 0.5f), 8388608.f)) & 0xFu;
      return (c0 << 4) | c1;
    } else {
      // rint by the 1.5 * 2^23 trick (|v| <= 8 after the clamp; NaN -> code 0 like the eager cast)
      const bool n0 = a != a, n1 = b != b;
      a = fminf(fmaxf(a, -8.f), 7.f);
      b = fminf(fmaxf(b, -8.f), 7.f);
      const uint32_t c0 = n0 ? 0u : (__float_as_uint(__fadd_rn(a, 12582912.f)) + 8u) & 0xFu;
      const uint32_t c1 = n1 ? 0u : (__float_as_uint(__fadd_rn(b, 12582912.f)) + 8u) & 0xFu;
      return (c0 << 4) | c1;
    }
  } else {
    return (int4_code<T, kKernelForm>(x0, s) << 4) | int4_code<T, kKernelForm>(x1, s);
  }
}

// NF4 table: the reference holds it in bf16 (nf4_tensor.py:25-45); its CUDA kernels use the fp32 literals
__device__ __constant__ float kNf4[16] = {-1.0000f, -0.6962f, -0.5251f, -0.3949f, -0.2844f, -0.1848f, -0.0911f, 0.0000f,
                                          0.0796f,  0.1609f,  0.2461f,  0.3379f,  0.4407f,  0.5626f,  0.7230f,  1.0000f};

template <typename T, bool kKernelForm>
struct Nf4Lut {
  const float* v;  // 16 table values then 15 midpoints, in shared memory
  __device__ __forceinline__ explicit Nf4Lut(float* smem) : v(smem) {
    if (threadIdx.x < 16)
      smem[threadIdx.x] = kKernelForm ? kNf4[threadIdx.x]
                                      : round_to<T>(__bfloat162float(__float2bfloat16_rn(kNf4[threadIdx.x])));
    __syncthreads();
    if (threadIdx.x < 15) smem[16 + threadIdx.x] = 0.5f * (smem[threadIdx.x] + smem[threadIdx.x + 1]);
    __syncthreads();
  }
  __device__ __forceinline__ float diff(float x, int i) const {
    if constexpr (kKernelForm) return fabsf(__fsub_rn(v[i], x));  // find_closest_index works in fp32
    else return round_to<T>(fabsf(round_to<T>(__fsub_rn(x, v[i]))));
  }
  // first argmin_i T(|T(x - lut_i)|): the candidates are the entry nearest in exact arithmetic and its two neighbours
  // (a rounded difference can tie only between adjacent entries: spacing >= 0.08, relative rounding <= 2^-8)
  __device__ __forceinline__ uint32_t nearest(float x) const {
    if (x != x) return 0u;  // argmin over NaNs / the CUDA scan both stay at index 0
    // k = #{midpoints < x} by a 4-step binary search over the 15 sorted midpoints (was 15 compares)
    int k = x > v[16 + 7] ? 8 : 0;
    k += x > v[16 + k + 3] ? 4 : 0;
    k += x > v[16 + k + 1] ? 2 : 0;
    k += x > v[16 + k] ? 1 : 0;
    const int lo = max(k - 1, 0), hi = min(k + 1, 15);
    int best = lo;
    float bd = diff(x, lo);
    for (int i = lo + 1; i <= hi; ++i) {
      const float d = diff(x, i);
      if (d < bd) { bd = d; best = i; }
    }
    return (uint32_t)best;
  }
};

// kind 0 = INT4, 1 = NF4.  Vector path: block = L * E elements, L a power of two <= 32.
template <typename T, int KIND, bool kKernelForm>
__global__ void __launch_bounds__(kThreads) pack4_vec_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                             T* __restrict__ scales, int64_t n_vec, int lanes_log2) {
  constexpr int E = Elem<T>::kPerVec;
  const int lanes = 1 << lanes_log2;
  const V32* xv = reinterpret_cast<const V32*>(x);
  __shared__ float s_lut[32];
  Nf4Lut<T, kKernelForm> lut(s_lut);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  // n_vec is a multiple of `lanes` (whole blocks) and every trip starts a warp at a multiple of 32, so the lanes of a
  // block are all valid or all invalid and the shuffles can use the full mask
  constexpr int U = 2;  // vectors in flight per thread
  for (int64_t w0 = (int64_t)blockIdx.x * kThreads + (threadIdx.x & ~31); w0 < n_vec; w0 += stride * U) {
    V32 raw[U];
    bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = w0 + u * stride + (threadIdx.x & 31);
      valid[u] = i < n_vec;
      if (valid[u]) raw[u] = ld_stream(xv + i);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = w0 + u * stride + (threadIdx.x & 31);
      if (w0 + u * stride >= n_vec) break;  // warp-uniform
      float f[E];
      if (valid[u]) Elem<T>::unpack(raw[u], f);
      else {
#pragma unroll
        for (int k = 0; k < E; ++k) f[k] = 0.f;
      }
      float m = fabsf(f[0]);
#pragma unroll
      for (int k = 1; k < E; ++k) m = fmaxf(m, fabsf(f[k]));
      for (int o = 1; o < lanes; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (!valid[u]) continue;
      float s;
      if constexpr (KIND == 0) s = int4_scale<T>(m);
      else s = m;
      if ((i & (lanes - 1)) == 0) Elem<T>::store(scales + (i >> lanes_log2), s);
      const DivSeq2 nf_dv(s);
      const bool nf_fast = KIND == 1 && nf_dv.ok && m == m && s < 3.0e38f;  // finite block, divisor in the exact window
      uint32_t w[E / 8];
#pragma unroll
      for (int k = 0; k < E / 8; ++k) w[k] = 0;
#pragma unroll
      for (int k = 0; k < E; k += 2) {  // byte j holds elements 2j (high nibble) and 2j+1 (low nibble)
        uint32_t byte;
        if constexpr (KIND == 0) byte = int4_byte<T, kKernelForm>(f[k], f[k + 1], s);
        else {
          float a = f[k], b = f[k + 1];
          if (nf_fast) {  // |x| <= s: the quotient is in [-1, 1]; the division's own FMA sequence, reciprocal hoisted
            nf_dv(a, b);
            a = copysignf(a, f[k]);
            b = copysignf(b, f[k + 1]);
          } else {
            a = __fdiv_rn(a, s);
            b = __fdiv_rn(b, s);
          }
          round_to2<T>(a, b);
          byte = (lut.nearest(a) << 4) | lut.nearest(b);
        }
        w[k / 8] |= byte << (8 * ((k % 8) / 2));
      }
      if constexpr (E == 16) *reinterpret_cast<uint2*>(q + i * 8) = make_uint2(w[0], w[1]);
      else *reinterpret_cast<uint32_t*>(q + i * 4) = w[0];
    }
  }
}

// Generic: one thread per block, any even block size.
template <typename T, int KIND, bool kKernelForm>
__global__ void __launch_bounds__(kThreads) pack4_generic_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                                 T* __restrict__ scales, int64_t n_blocks, int block) {
  __shared__ float s_lut[32];
  Nf4Lut<T, kKernelForm> lut(s_lut);
  for (int64_t b = (int64_t)blockIdx.x * kThreads + threadIdx.x; b < n_blocks; b += (int64_t)gridDim.x * kThreads) {
    const T* xp = x + b * block;
    float m = 0.f;
    for (int k = 0; k < block; ++k) m = fmaxf(m, fabsf(Elem<T>::load(xp + k)));
    const float s = KIND == 0 ? int4_scale<T>(m) : m;
    Elem<T>::store(scales + b, s);
    for (int k = 0; k < block; k += 2) {
      uint32_t c[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float v = Elem<T>::load(xp + k + j);
        if constexpr (KIND == 0) c[j] = int4_code<T, kKernelForm>(v, s);
        else c[j] = lut.nearest(round_to<T>(__fdiv_rn(v, s)));
      }
      q[(b * block + k) >> 1] = (uint8_t)((c[0] << 4) | c[1]);
    }
  }
}

// INT4 dequantize: T((nibble - 8) / scale) (int4_tensor.py:108-129, tensor_quant_gpu.cu:265-283); S = scale/out dtype.
// vec: a thread turns E/2 bytes into one 32-byte vector of outputs (the quotient by ExactDiv: integer numerators).
template <typename S>
__global__ void __launch_bounds__(kThreads) int4_unpack_kernel(const uint8_t* __restrict__ q, const S* __restrict__ scales,
                                                               S* __restrict__ y, int64_t n_bytes, int half_block, int vec, int vshift) {
  constexpr int E = Elem<S>::kPerVec;
  if (vec) {  // half_block % (E / 2) == 0, pointers aligned
    const int64_t nv = n_bytes / (E / 2);
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      const ExactDiv dv(Elem<S>::load(scales + (vshift >= 0 ? i >> vshift : (i * (E / 2)) / half_block)));
      uint32_t w[E / 8];
      if constexpr (E == 16) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(q) + i);
        w[0] = u.x; w[1] = u.y;
      } else {
        w[0] = __ldg(reinterpret_cast<const uint32_t*>(q) + i);
      }
      float f[E];
#pragma unroll
      for (int k = 0; k < E / 2; ++k) {
        const uint32_t p = (w[k / 4] >> (8 * (k % 4))) & 0xFFu;
        f[2 * k] = dv((float)((int)(p >> 4) - 8));
        f[2 * k + 1] = dv((float)((int)(p & 0xF) - 8));
      }
      V32 o;
      Elem<S>::pack(f, o);
      st_stream(reinterpret_cast<V32*>(y) + i, o);
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_bytes; i += (int64_t)gridDim.x * kThreads) {
    const float s = Elem<S>::load(scales + i / half_block);
    const uint32_t p = q[i];
    Elem<S>::store(y + 2 * i, __fdiv_rn((float)((int)(p >> 4) - 8), s));
    Elem<S>::store(y + 2 * i + 1, __fdiv_rn((float)((int)(p & 0xF) - 8), s));
  }
}

// NF4 dequantize: lut_bf16[idx] * scale.  bf16_form: the CUDA kernel (scale rounded to bf16, bf16 product, bf16 out;
// tensor_quant_gpu.cu:146-166); otherwise the eager product in fp32 (bf16 table value * fp32 scale).
template <typename O>
__global__ void __launch_bounds__(kThreads) nf4_unpack_kernel(const uint8_t* __restrict__ q, const float* __restrict__ scales,
                                                              O* __restrict__ y, int64_t n_bytes, int half_block,
                                                              int bf16_form, int vec, int vshift) {
  constexpr int E = Elem<O>::kPerVec;
  __shared__ float lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = round_to<__nv_bfloat16>(kNf4[threadIdx.x]);
  __syncthreads();
  if (vec) {
    const int64_t nv = n_bytes / (E / 2);
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      float s = __ldg(scales + (vshift >= 0 ? i >> vshift : (i * (E / 2)) / half_block));
      if (bf16_form) s = round_to<__nv_bfloat16>(s);
      uint32_t w[E / 8];
      if constexpr (E == 16) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(q) + i);
        w[0] = u.x; w[1] = u.y;
      } else {
        w[0] = __ldg(reinterpret_cast<const uint32_t*>(q) + i);
      }
      float f[E];
#pragma unroll
      for (int k = 0; k < E / 2; ++k) {
        const uint32_t p = (w[k / 4] >> (8 * (k % 4))) & 0xFFu;
        f[2 * k] = __fmul_rn(lut[p >> 4], s);
        f[2 * k + 1] = __fmul_rn(lut[p & 0xF], s);
      }
      V32 o;
      Elem<O>::pack(f, o);
      st_stream(reinterpret_cast<V32*>(y) + i, o);
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_bytes; i += (int64_t)gridDim.x * kThreads) {
    float s = scales[i / half_block];
    if (bf16_form) s = round_to<__nv_bfloat16>(s);
    const uint32_t p = q[i];
    Elem<O>::store(y + 2 * i, __fmul_rn(lut[p >> 4], s));
    Elem<O>::store(y + 2 * i + 1, __fmul_rn(lut[p & 0xF], s));
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// INT8 / FP8 packers with per-tensor, per-axis or N-D block scales.  x is viewed [O, A, I]; the scale of element
// (o, a, i) is scales[(o / bo) * so + (a / ba) * sa + (i / bi) * si].
//   INT8: int8(clamp(round(T(x / s)), -128, 127))  (int8_tensor.py:91)
//   FP8 : e4m3(T(x / s))                           (fp8_tensor.py:109)
// ---------------------------------------------------------------------------------------------------------------------
struct ScaleMap {
  int64_t A, I, bo, ba, bi, so, sa, si;
  int small;  // every index fits 32 bits: mul.hi divisions (FastDiv) instead of 64-bit ones
  FastDiv dI, dA, dbo, dba, dbi;
  __device__ __forceinline__ int64_t index(int64_t flat) const {
    if (small) {
      const uint32_t f = (uint32_t)flat;
      const uint32_t row = dI.div(f), i = f - row * (uint32_t)I, o = dA.div(row), a = row - o * (uint32_t)A;
      return (int64_t)(dbo.div(o) * (uint32_t)so + dba.div(a) * (uint32_t)sa + dbi.div(i) * (uint32_t)si);
    }
    const int64_t i = flat % I, a = (flat / I) % A, o = flat / (I * A);
    return (o / bo) * so + (a / ba) * sa + (i / bi) * si;
  }
};

template <typename T, int KIND>
__device__ __forceinline__ uint32_t q8_code_of(float v) {  // v = T(x / s)
  if constexpr (KIND == 0) {
    if (v != v) return 0u;
    return (uint32_t)(int)fminf(fmaxf(rintf(v), -128.f), 127.f) & 0xFFu;
  } else if constexpr (KIND == 1) {
    // torch's float8_e4m3fn cast does not saturate: anything that rounds past 448 (|v| > 464, inf, NaN) is NaN (0x7f)
    if (!(fabsf(v) <= 464.f)) return 0x7Fu | ((__float_as_uint(v) >> 24) & 0x80u);
    return e4m3_code1(v);
  } else {  // KIND 2: clamp(-448, 448) first (backends/fp8_per_tensor_gemm.py:31-33) = the saturating conversion
    return e4m3_code1(v);
  }
}

// KIND 2 (`_to_fp8` under torch.compile: the quotient stays in fp32, the scale is an fp32 tensor): no rounding to T
template <typename T, int KIND>
__device__ __forceinline__ float q8_scale(const T* scales, int64_t idx) {
  if constexpr (KIND == 2) return reinterpret_cast<const float*>(scales)[idx];
  else return Elem<T>::load(scales + idx);
}

template <typename T, int KIND>
__device__ __forceinline__ uint32_t q8_code(float x, float s) {
  const float v = __fdiv_rn(x, s);
  return q8_code_of<T, KIND>(KIND == 2 ? v : round_to<T>(v));
}

// vec: a thread owns one 32-byte vector of x (E elements that share a scale): the IEEE quotients come from the
// division's own instruction sequence with the reciprocal hoisted (DivSeq2, common.cuh) - the sign of a zero quotient and
// non-finite results are patched / recomputed so every code equals the one from `x / s`.
template <typename T, int KIND>
__global__ void __launch_bounds__(kThreads) pack8_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                         const T* __restrict__ scales, int64_t n, ScaleMap map, int vec) {
  constexpr int E = Elem<T>::kPerVec;
  if (vec) {
    const int64_t nv = n / E;
    const V32* xv = reinterpret_cast<const V32*>(x);
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      const V32 v = ld_stream(xv + i);
      const float s = q8_scale<T, KIND>(scales, map.index(i * E));
      float f[E];
      Elem<T>::unpack(v, f);
      const DivSeq2 dv(s);
      float m = fabsf(f[0]);
#pragma unroll
      for (int k = 1; k < E; ++k) m = fmaxf(m, fabsf(f[k]));
      // every quotient finite and the divisor in the sequence's window (s > 0 there); NaN inputs pass through it as NaN
      const bool fast = dv.ok && __fmul_rn(m, dv.r) <= 1.0e38f;
      uint32_t w[E / 4];
#pragma unroll
      for (int k = 0; k < E / 4; ++k) w[k] = 0;
      float top = 0.f;
#pragma unroll
      for (int k = 0; k < E; k += 2) {
        float a = f[k], b = f[k + 1];
        if (fast) {
          dv(a, b);
          a = copysignf(a, f[k]);  // fma(x, r, +0) loses the sign of a zero quotient
          b = copysignf(b, f[k + 1]);
        } else {
          a = __fdiv_rn(a, s);
          b = __fdiv_rn(b, s);
        }
        if constexpr (KIND != 2) round_to2<T>(a, b);
        uint32_t pair;
        if constexpr (KIND == 0) {  // round-to-nearest-even, saturating, NaN -> 0: one conversion each
          int c0, c1;
          asm("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(c0) : "f"(a));
          asm("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(c1) : "f"(b));
          pair = ((uint32_t)c0 & 0xFFu) | (((uint32_t)c1 & 0xFFu) << 8);
        } else {
          uint16_t e;
          asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(e) : "f"(b), "f"(a));
          pair = e;
          top = fmaxf(top, fmaxf(fabsf(a), fabsf(b)));
          if (a != a || b != b) top = __int_as_float(0x7f800000);
        }
        w[k / 4] |= pair << (8 * (k % 4));
        f[k] = a;
        f[k + 1] = b;
      }
      if constexpr (KIND == 1) {
        if (!(top <= 464.f)) {  // torch's cast does not saturate: redo this vector with the NaN rule (rare)
#pragma unroll
          for (int k = 0; k < E / 4; ++k) w[k] = 0;
#pragma unroll
          for (int k = 0; k < E; ++k) w[k / 4] |= q8_code_of<T, KIND>(f[k]) << (8 * (k % 4));
        }
      }
      if constexpr (E == 16) reinterpret_cast<uint4*>(q)[i] = make_uint4(w[0], w[1], w[2], w[3]);
      else reinterpret_cast<uint2*>(q)[i] = make_uint2(w[0], w[1]);
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
      q[i] = (uint8_t)q8_code<T, KIND>(Elem<T>::load(x + i), q8_scale<T, KIND>(scales, map.index(i)));
  }
}

__device__ __forceinline__ float q8_value(uint32_t byte, int kind) {
  if (kind == 0) return (float)(int8_t)byte;
  uint32_t h2;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)byte));
  return __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
}

// dequantize: O(O(q) * O(scale))  (int8_tensor.py:124, fp8_tensor.py:151); S = scale dtype, O = output dtype.
// vec: a thread turns E bytes into one 32-byte vector of outputs.
template <typename S, typename O, int KIND>
__global__ void __launch_bounds__(kThreads) unpack8_kernel(const uint8_t* __restrict__ q, const S* __restrict__ scales,
                                                           O* __restrict__ y, int64_t n, ScaleMap map, int vec) {
  constexpr int E = Elem<O>::kPerVec;
  if (vec) {
    const int64_t nv = n / E;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      const float s = round_to<O>(Elem<S>::load(scales + map.index(i * E)));
      uint32_t w[E / 4];
      if constexpr (E == 16) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(q) + i);
        w[0] = u.x; w[1] = u.y; w[2] = u.z; w[3] = u.w;
      } else {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(q) + i);
        w[0] = u.x; w[1] = u.y;
      }
      float f[E];
#pragma unroll
      for (int k = 0; k < E; k += 2) {
        const uint32_t pair = (w[k / 4] >> (8 * (k % 4))) & 0xFFFFu;
        if constexpr (KIND == 0) {
          f[k] = (float)(int8_t)(pair & 0xFF);
          f[k + 1] = (float)(int8_t)(pair >> 8);
        } else {
          uint32_t h2;
          asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)pair));
          const float2 t = __half22float2(*reinterpret_cast<__half2*>(&h2));
          f[k] = t.x;
          f[k + 1] = t.y;
        }
        f[k] = __fmul_rn(f[k], s);
        f[k + 1] = __fmul_rn(f[k + 1], s);
      }
      V32 o;
      Elem<O>::pack(f, o);
      st_stream(reinterpret_cast<V32*>(y) + i, o);
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    const float s = round_to<O>(Elem<S>::load(scales + map.index(i)));
    Elem<O>::store(y + i, __fmul_rn(q8_value(q[i], KIND), s));
  }
}

template <typename T, int KIND>
static int launch_pack4(const void* xv, uint8_t* q, void* sv, int64_t n_blocks, int block, bool kernel_form, cudaStream_t s) {
  const T* x = static_cast<const T*>(xv);
  T* sc = static_cast<T*>(sv);
  constexpr int E = Elem<T>::kPerVec;
  const int lanes = block / E;
  const bool vec = block % E == 0 && lanes >= 1 && lanes <= 32 && (lanes & (lanes - 1)) == 0 && aligned32(xv) &&
                   (reinterpret_cast<uintptr_t>(q) & 7u) == 0;
  if (vec) {
    const int64_t n_vec = n_blocks * lanes;
    const int64_t tiles = (n_vec + kThreads - 1) / kThreads;
    int lanes_log2 = 0;
    while ((1 << lanes_log2) < lanes) ++lanes_log2;
    if (kernel_form)
      pack4_vec_kernel<T, KIND, true><<<resident_grid((const void*)pack4_vec_kernel<T, KIND, true>, tiles), kThreads, 0, s>>>(x, q, sc, n_vec, lanes_log2);
    else
      pack4_vec_kernel<T, KIND, false><<<resident_grid((const void*)pack4_vec_kernel<T, KIND, false>, tiles), kThreads, 0, s>>>(x, q, sc, n_vec, lanes_log2);
  } else {
    const int grid = persistent_grid((n_blocks + kThreads - 1) / kThreads, 8);
    if (kernel_form) pack4_generic_kernel<T, KIND, true><<<grid, kThreads, 0, s>>>(x, q, sc, n_blocks, block);
    else pack4_generic_kernel<T, KIND, false><<<grid, kThreads, 0, s>>>(x, q, sc, n_blocks, block);
  }
  return check_launch(KIND == 0 ? "mo_int4_pack" : "mo_nf4_pack");
}

template <int KIND>
static int pack4_entry(const char* name, const void* x, int64_t n_blocks, int block, int dtype, void* scales, uint8_t* q,
                       int numerics, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !scales || !q) { set_error("%s: NULL pointer", name); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 2 || (block & 1)) { set_error("%s: block must be even and positive", name); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case MO_F32: return launch_pack4<float, KIND>(x, q, scales, n_blocks, block, numerics != MO_DIV_IEEE, s);
    case MO_F16: return launch_pack4<__half, KIND>(x, q, scales, n_blocks, block, numerics != MO_DIV_IEEE, s);
    case MO_BF16: return launch_pack4<__nv_bfloat16, KIND>(x, q, scales, n_blocks, block, numerics != MO_DIV_IEEE, s);
  }
  set_error("%s: unsupported dtype %d", name, dtype);
  return MO_ERR_DTYPE;
}

static ScaleMap make_map(const int64_t* dims, int64_t n) {
  const int64_t lim = (int64_t)1 << 31;
  const int64_t n_scales = ((dims[0] + dims[3] - 1) / dims[3]) * dims[6] + dims[6] + dims[7] * dims[1] + dims[2];
  ScaleMap m{dims[1], dims[2], dims[3], dims[4], dims[5], dims[6], dims[7], dims[8], 0, {}, {}, {}, {}, {}};
  bool small = n < lim && n_scales < lim;
  for (int k = 1; k <= 5; ++k) small = small && dims[k] < lim;
  if (small) {
    m.small = 1;
    m.dI.init((uint32_t)dims[2]);
    m.dA.init((uint32_t)dims[1]);
    m.dbo.init((uint32_t)dims[3]);
    m.dba.init((uint32_t)dims[4]);
    m.dbi.init((uint32_t)dims[5]);
  }
  return m;
}

// vectors of `per_vec` bytes per block of `half_block` bytes: log2 when a power of two, else -1
static int vec_shift(int half_block, int per_vec) {
  if (half_block % per_vec) return -1;
  const int v = half_block / per_vec;
  if (v & (v - 1)) return -1;
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

static bool map_ok(const int64_t* d) {
  return d[0] > 0 && d[1] > 0 && d[2] > 0 && d[3] > 0 && d[4] > 0 && d[5] >= 0 && d[6] >= 0 && d[7] >= 0;
}

}  // namespace mo

using namespace mo;

extern "C" int mo_int4_pack(const void* x, int64_t n_blocks, int block, int dtype, void* scales, uint8_t* q,
                            int numerics, void* stream) {
  return pack4_entry<0>("mo_int4_pack", x, n_blocks, block, dtype, scales, q, numerics, stream);
}

extern "C" int mo_nf4_pack(const void* x, int64_t n_blocks, int block, int dtype, void* scales, uint8_t* q,
                           int numerics, void* stream) {
  return pack4_entry<1>("mo_nf4_pack", x, n_blocks, block, dtype, scales, q, numerics, stream);
}

extern "C" int mo_int4_unpack(const uint8_t* q, const void* scales, int64_t n_blocks, int block, int dtype, void* y,
                              void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!q || !scales || !y) { set_error("mo_int4_unpack: NULL pointer"); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 2 || (block & 1)) { set_error("mo_int4_unpack: bad block"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t nb = n_blocks * (block / 2);
  const int ev = dtype == MO_F32 ? 8 : 16;  // outputs per 32-byte vector
  const int vec = (block / 2) % (ev / 2) == 0 && aligned32(y) && (reinterpret_cast<uintptr_t>(q) & 7u) == 0;
  const int grid = persistent_grid(((vec ? nb / (ev / 2) : nb) + kThreads - 1) / kThreads, 16);
  const int vshift = vec_shift(block / 2, ev / 2);
  switch (dtype) {
    case MO_F32: int4_unpack_kernel<float><<<grid, kThreads, 0, s>>>(q, static_cast<const float*>(scales), static_cast<float*>(y), nb, block / 2, vec, vshift); break;
    case MO_F16: int4_unpack_kernel<__half><<<grid, kThreads, 0, s>>>(q, static_cast<const __half*>(scales), static_cast<__half*>(y), nb, block / 2, vec, vshift); break;
    case MO_BF16: int4_unpack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(q, static_cast<const __nv_bfloat16*>(scales), static_cast<__nv_bfloat16*>(y), nb, block / 2, vec, vshift); break;
    default: set_error("mo_int4_unpack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_int4_unpack");
}

extern "C" int mo_nf4_unpack(const uint8_t* q, const float* scales, int64_t n_blocks, int block, int out_dtype,
                             int bf16_form, void* y, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!q || !scales || !y) { set_error("mo_nf4_unpack: NULL pointer"); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 2 || (block & 1)) { set_error("mo_nf4_unpack: bad block"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t nb = n_blocks * (block / 2);
  const int ev = out_dtype == MO_F32 ? 8 : 16;
  const int vec = (block / 2) % (ev / 2) == 0 && aligned32(y) && (reinterpret_cast<uintptr_t>(q) & 7u) == 0;
  const int grid = persistent_grid(((vec ? nb / (ev / 2) : nb) + kThreads - 1) / kThreads, 16);
  const int vshift = vec_shift(block / 2, ev / 2);
  switch (out_dtype) {
    case MO_F32: nf4_unpack_kernel<float><<<grid, kThreads, 0, s>>>(q, scales, static_cast<float*>(y), nb, block / 2, bf16_form, vec, vshift); break;
    case MO_BF16: nf4_unpack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(q, scales, static_cast<__nv_bfloat16*>(y), nb, block / 2, bf16_form, vec, vshift); break;
    default: set_error("mo_nf4_unpack: out dtype must be fp32 or bf16, got %d", out_dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_nf4_unpack");
}

// dims = {O, A, I, bo, ba, bi, so, sa, si}
extern "C" int mo_q8_pack(const void* x, uint8_t* q, const void* scales, const int64_t* dims, int dtype, int kind,
                          void* stream) {
  if (!dims) { set_error("mo_q8_pack: NULL dims"); return MO_ERR_NULL; }
  const int64_t n = dims[0] * dims[1] * dims[2];
  if (n == 0) return MO_OK;
  if (!x || !q || !scales) { set_error("mo_q8_pack: NULL pointer"); return MO_ERR_NULL; }
  if (!map_ok(dims + 1) || dims[0] < 0 || kind < 0 || kind > 2) { set_error("mo_q8_pack: bad dims / kind"); return MO_ERR_ARG; }
  const ScaleMap map = make_map(dims, n);
  const int ev = dtype == MO_F32 ? 8 : 16;
  const int vec = dims[2] % ev == 0 && dims[5] % ev == 0 && aligned32(x) && (reinterpret_cast<uintptr_t>(q) & 15u) == 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t work = vec ? n / ev : n;
#define MO_Q8(T, K) pack8_kernel<T, K><<<resident_grid((const void*)pack8_kernel<T, K>, (work + kThreads - 1) / kThreads), kThreads, 0, s>>>(static_cast<const T*>(x), q, static_cast<const T*>(scales), n, map, vec)
  switch (dtype) {
    case MO_F32: if (kind == 0) MO_Q8(float, 0); else if (kind == 1) MO_Q8(float, 1); else MO_Q8(float, 2); break;
    case MO_F16: if (kind == 0) MO_Q8(__half, 0); else if (kind == 1) MO_Q8(__half, 1); else MO_Q8(__half, 2); break;
    case MO_BF16: if (kind == 0) MO_Q8(__nv_bfloat16, 0); else if (kind == 1) MO_Q8(__nv_bfloat16, 1); else MO_Q8(__nv_bfloat16, 2); break;
    default: set_error("mo_q8_pack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
#undef MO_Q8
  return check_launch("mo_q8_pack");
}

template <typename S, int KIND>
static int unpack8_out(const uint8_t* q, const void* scales, void* y, int64_t n, const ScaleMap& map, int out_dtype, cudaStream_t s) {
  const int ev = out_dtype == MO_F32 ? 8 : 16;
  const int vec = map.I % ev == 0 && map.bi % ev == 0 && aligned32(y) && (reinterpret_cast<uintptr_t>(q) & 15u) == 0;
  const int grid = persistent_grid(((vec ? n / ev : n) + kThreads - 1) / kThreads, 16);
  const S* sc = static_cast<const S*>(scales);
  switch (out_dtype) {
    case MO_F32: unpack8_kernel<S, float, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<float*>(y), n, map, vec); break;
    case MO_F16: unpack8_kernel<S, __half, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<__half*>(y), n, map, vec); break;
    case MO_BF16: unpack8_kernel<S, __nv_bfloat16, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<__nv_bfloat16*>(y), n, map, vec); break;
    default: set_error("mo_q8_unpack: unsupported out dtype %d", out_dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_q8_unpack");
}

extern "C" int mo_q8_unpack(const uint8_t* q, const void* scales, void* y, const int64_t* dims, int scale_dtype,
                            int out_dtype, int kind, void* stream) {
  if (!dims) { set_error("mo_q8_unpack: NULL dims"); return MO_ERR_NULL; }
  const int64_t n = dims[0] * dims[1] * dims[2];
  if (n == 0) return MO_OK;
  if (!q || !scales || !y) { set_error("mo_q8_unpack: NULL pointer"); return MO_ERR_NULL; }
  if (!map_ok(dims + 1) || dims[0] < 0 || (kind != 0 && kind != 1)) { set_error("mo_q8_unpack: bad dims / kind"); return MO_ERR_ARG; }
  const ScaleMap map = make_map(dims, n);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (scale_dtype) {
    case MO_F32: return kind == 0 ? unpack8_out<float, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<float, 1>(q, scales, y, n, map, out_dtype, s);
    case MO_F16: return kind == 0 ? unpack8_out<__half, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<__half, 1>(q, scales, y, n, map, out_dtype, s);
    case MO_BF16: return kind == 0 ? unpack8_out<__nv_bfloat16, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<__nv_bfloat16, 1>(q, scales, y, n, map, out_dtype, s);
  }
  set_error("mo_q8_unpack: unsupported scale dtype %d", scale_dtype);
  return MO_ERR_DTYPE;
}
