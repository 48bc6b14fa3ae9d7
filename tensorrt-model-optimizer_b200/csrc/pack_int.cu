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

namespace mo {

template <typename T>
__device__ __forceinline__ float round_to(float v);
template <>
__device__ __forceinline__ float round_to<float>(float v) { return v; }
template <>
__device__ __forceinline__ float round_to<__nv_bfloat16>(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
template <>
__device__ __forceinline__ float round_to<__half>(float v) { return __half2float(__float2half_rn(v)); }

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
    return (uint32_t)(int)roundf(round_to<T>(__fadd_rn(v, 8.f))) & 0xFu;
  } else {
    float v = round_to<T>(__fmul_rn(x, s));
    if (v != v) return 0u;
    v = fminf(fmaxf(rintf(v), -8.f), 7.f);
    return (uint32_t)((int)v + 8);
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
    int k = 0;
#pragma unroll
    for (int i = 0; i < 15; ++i) k += x > v[16 + i];
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
                                                             T* __restrict__ scales, int64_t n_vec, int lanes) {
  constexpr int E = Elem<T>::kPerVec;
  const V32* xv = reinterpret_cast<const V32*>(x);
  __shared__ float s_lut[32];
  Nf4Lut<T, kKernelForm> lut(s_lut);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  // n_vec is a multiple of `lanes` (whole blocks) and every trip starts a warp at a multiple of 32, so the lanes of a
  // block are all valid or all invalid and the shuffles can use the full mask
  for (int64_t w0 = (int64_t)blockIdx.x * kThreads + (threadIdx.x & ~31); w0 < n_vec; w0 += stride) {
    const int64_t i = w0 + (threadIdx.x & 31);
    const bool valid = i < n_vec;
    float f[E];
    if (valid) Elem<T>::unpack(ld_stream(xv + i), f);
    else {
#pragma unroll
      for (int k = 0; k < E; ++k) f[k] = 0.f;
    }
    float m = fabsf(f[0]);
#pragma unroll
    for (int k = 1; k < E; ++k) m = fmaxf(m, fabsf(f[k]));
    for (int o = 1; o < lanes; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (!valid) continue;
    float s;
    if constexpr (KIND == 0) s = int4_scale<T>(m);
    else s = m;
    if ((i & (lanes - 1)) == 0) Elem<T>::store(scales + i / lanes, s);
    uint32_t w[E / 8];
#pragma unroll
    for (int k = 0; k < E / 8; ++k) w[k] = 0;
#pragma unroll
    for (int k = 0; k < E; ++k) {
      uint32_t c;
      if constexpr (KIND == 0) c = int4_code<T, kKernelForm>(f[k], s);
      else c = lut.nearest(round_to<T>(__fdiv_rn(f[k], s)));
      // byte j holds elements 2j (high nibble) and 2j+1 (low nibble)
      w[k / 8] |= c << (8 * ((k % 8) / 2) + ((k & 1) ? 0 : 4));
    }
    if constexpr (E == 16) *reinterpret_cast<uint2*>(q + i * 8) = make_uint2(w[0], w[1]);
    else *reinterpret_cast<uint32_t*>(q + i * 4) = w[0];
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

// INT4 dequantize: T((nibble - 8) / scale) (int4_tensor.py:108-129, tensor_quant_gpu.cu:265-283); S = scale/out dtype
template <typename S>
__global__ void __launch_bounds__(kThreads) int4_unpack_kernel(const uint8_t* __restrict__ q, const S* __restrict__ scales,
                                                               S* __restrict__ y, int64_t n_bytes, int half_block) {
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
                                                              int bf16_form) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_bytes; i += (int64_t)gridDim.x * kThreads) {
    float s = scales[i / half_block];
    if (bf16_form) s = round_to<__nv_bfloat16>(s);
    const uint32_t p = q[i];
    const float a = round_to<__nv_bfloat16>(kNf4[p >> 4]), b = round_to<__nv_bfloat16>(kNf4[p & 0xF]);
    Elem<O>::store(y + 2 * i, __fmul_rn(a, s));
    Elem<O>::store(y + 2 * i + 1, __fmul_rn(b, s));
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
  __device__ __forceinline__ int64_t index(int64_t flat) const {
    const int64_t i = flat % I, a = (flat / I) % A, o = flat / (I * A);
    return (o / bo) * so + (a / ba) * sa + (i / bi) * si;
  }
};

template <typename T, int KIND>
__device__ __forceinline__ uint32_t q8_code(float x, float s) {
  const float v = round_to<T>(__fdiv_rn(x, s));
  if constexpr (KIND == 0) {
    if (v != v) return 0u;
    return (uint32_t)(int)fminf(fmaxf(rintf(v), -128.f), 127.f) & 0xFFu;
  } else {
    // torch's float8_e4m3fn cast does not saturate: anything that rounds past 448 (|v| > 464, inf, NaN) is NaN (0x7f)
    if (!(fabsf(v) <= 464.f)) return 0x7Fu | ((__float_as_uint(v) >> 24) & 0x80u);
    return e4m3_code1(v);
  }
}

template <typename T, int KIND>
__global__ void __launch_bounds__(kThreads) pack8_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                         const T* __restrict__ scales, int64_t n, ScaleMap map, int vec) {
  if (vec) {  // 8 consecutive elements share a scale, pointers 16-byte aligned
    const int64_t nv = n / 8;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += (int64_t)gridDim.x * kThreads) {
      const float s = Elem<T>::load(scales + map.index(i * 8));
      float f[8];
      if constexpr (sizeof(T) == 2) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(x) + i);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            f[2 * k] = __uint_as_float(w[k] << 16);
            f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u);
          } else {
            const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[k]));
            f[2 * k] = t.x;
            f[2 * k + 1] = t.y;
          }
        }
      } else {
        const float4 a = __ldg(reinterpret_cast<const float4*>(x) + 2 * i), b = __ldg(reinterpret_cast<const float4*>(x) + 2 * i + 1);
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
      }
      uint2 o = {0u, 0u};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        o.x |= q8_code<T, KIND>(f[k], s) << (8 * k);
        o.y |= q8_code<T, KIND>(f[4 + k], s) << (8 * k);
      }
      reinterpret_cast<uint2*>(q)[i] = o;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
      q[i] = (uint8_t)q8_code<T, KIND>(Elem<T>::load(x + i), Elem<T>::load(scales + map.index(i)));
  }
}

// dequantize: O(O(q) * O(scale))  (int8_tensor.py:124, fp8_tensor.py:151); S = scale dtype, O = output dtype
template <typename S, typename O, int KIND>
__global__ void __launch_bounds__(kThreads) unpack8_kernel(const uint8_t* __restrict__ q, const S* __restrict__ scales,
                                                           O* __restrict__ y, int64_t n, ScaleMap map) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    const float s = round_to<O>(Elem<S>::load(scales + map.index(i)));
    float v;
    if constexpr (KIND == 0) v = (float)(int8_t)q[i];
    else {
      uint32_t h2;
      asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)q[i]));
      v = __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
    }
    Elem<O>::store(y + i, __fmul_rn(v, s));
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
    if (kernel_form)
      pack4_vec_kernel<T, KIND, true><<<resident_grid((const void*)pack4_vec_kernel<T, KIND, true>, tiles), kThreads, 0, s>>>(x, q, sc, n_vec, lanes);
    else
      pack4_vec_kernel<T, KIND, false><<<resident_grid((const void*)pack4_vec_kernel<T, KIND, false>, tiles), kThreads, 0, s>>>(x, q, sc, n_vec, lanes);
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
  const int grid = persistent_grid((nb + kThreads - 1) / kThreads, 16);
  switch (dtype) {
    case MO_F32: int4_unpack_kernel<float><<<grid, kThreads, 0, s>>>(q, static_cast<const float*>(scales), static_cast<float*>(y), nb, block / 2); break;
    case MO_F16: int4_unpack_kernel<__half><<<grid, kThreads, 0, s>>>(q, static_cast<const __half*>(scales), static_cast<__half*>(y), nb, block / 2); break;
    case MO_BF16: int4_unpack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(q, static_cast<const __nv_bfloat16*>(scales), static_cast<__nv_bfloat16*>(y), nb, block / 2); break;
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
  const int grid = persistent_grid((nb + kThreads - 1) / kThreads, 16);
  switch (out_dtype) {
    case MO_F32: nf4_unpack_kernel<float><<<grid, kThreads, 0, s>>>(q, scales, static_cast<float*>(y), nb, block / 2, bf16_form); break;
    case MO_BF16: nf4_unpack_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(q, scales, static_cast<__nv_bfloat16*>(y), nb, block / 2, bf16_form); break;
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
  if (!map_ok(dims + 1) || dims[0] < 0 || (kind != 0 && kind != 1)) { set_error("mo_q8_pack: bad dims / kind"); return MO_ERR_ARG; }
  const ScaleMap map{dims[1], dims[2], dims[3], dims[4], dims[5], dims[6], dims[7], dims[8]};
  const int vec = dims[2] % 8 == 0 && (dims[5] % 8 == 0) && (reinterpret_cast<uintptr_t>(x) & 31u) == 0 &&
                  (reinterpret_cast<uintptr_t>(q) & 7u) == 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t work = vec ? n / 8 : n;
#define MO_Q8(T, K) pack8_kernel<T, K><<<resident_grid((const void*)pack8_kernel<T, K>, (work + kThreads - 1) / kThreads), kThreads, 0, s>>>(static_cast<const T*>(x), q, static_cast<const T*>(scales), n, map, vec)
  switch (dtype) {
    case MO_F32: if (kind == 0) MO_Q8(float, 0); else MO_Q8(float, 1); break;
    case MO_F16: if (kind == 0) MO_Q8(__half, 0); else MO_Q8(__half, 1); break;
    case MO_BF16: if (kind == 0) MO_Q8(__nv_bfloat16, 0); else MO_Q8(__nv_bfloat16, 1); break;
    default: set_error("mo_q8_pack: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
#undef MO_Q8
  return check_launch("mo_q8_pack");
}

template <typename S, int KIND>
static int unpack8_out(const uint8_t* q, const void* scales, void* y, int64_t n, const ScaleMap& map, int out_dtype, cudaStream_t s) {
  const int grid = persistent_grid((n + kThreads - 1) / kThreads, 16);
  const S* sc = static_cast<const S*>(scales);
  switch (out_dtype) {
    case MO_F32: unpack8_kernel<S, float, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<float*>(y), n, map); break;
    case MO_F16: unpack8_kernel<S, __half, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<__half*>(y), n, map); break;
    case MO_BF16: unpack8_kernel<S, __nv_bfloat16, KIND><<<grid, kThreads, 0, s>>>(q, sc, static_cast<__nv_bfloat16*>(y), n, map); break;
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
  const ScaleMap map{dims[1], dims[2], dims[3], dims[4], dims[5], dims[6], dims[7], dims[8]};
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (scale_dtype) {
    case MO_F32: return kind == 0 ? unpack8_out<float, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<float, 1>(q, scales, y, n, map, out_dtype, s);
    case MO_F16: return kind == 0 ? unpack8_out<__half, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<__half, 1>(q, scales, y, n, map, out_dtype, s);
    case MO_BF16: return kind == 0 ? unpack8_out<__nv_bfloat16, 0>(q, scales, y, n, map, out_dtype, s) : unpack8_out<__nv_bfloat16, 1>(q, scales, y, n, map, out_dtype, s);
  }
  set_error("mo_q8_unpack: unsupported scale dtype %d", scale_dtype);
  return MO_ERR_DTYPE;
}
