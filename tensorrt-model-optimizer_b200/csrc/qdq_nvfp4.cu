// NVFP4 (E2M1 values, E4M3 block scales, fp32 global scale) fake quantize-dequantize.
//
// Replaces the Triton kernels fp4_fake_quant_kernel (reference gemm/fp4_kernel_hopper.py:32-99),
// static_blockwise_fp4_fake_quant_kernel (gemm/fp4_kernel.py:193-214), static_blockwise_fp4_cast_kernel
// (:319-363) and the eager compute_fp4_scales (:217-251).  HBM-bound, 4 B per bf16 element.
//
// B200 shape: ONE THREAD OWNS ONE 16-ELEMENT BLOCK = one 32-byte vector (LDG.E.256): the block max is
// a register fmax chain (no shuffles, no shared memory), the scale is computed once per thread, the
// E2M1 rounding is the hardware cvt.rn.satfinite.e2m1x2.f32 (F2FP), and a warp reads / writes 1 KB
// contiguous.  Persistent grid, kUnrollB blocks in flight per thread.
//
// Numerics modes (MO_DIV_*): TRITON reproduces what the reference executes on an sm_100a GPU
// (div.full.f32 for both divisions, and torch-CUDA's `tensor / python_scalar` == multiply by the
// fp32 reciprocal for the host-side global scale); IEEE reproduces the CPU oracle (true divisions).
#include <math.h>

#include "common.cuh"

namespace mo {

constexpr int kUnrollB = 2;

template <bool kIeee>
__device__ __forceinline__ float global_scale_from_amax(float ga) {
  // fp4_kernel_hopper.py:140 (host torch op) then :73-74 (kernel)
  float gs = kIeee ? __fdiv_rn(ga, 2688.0f) : __fmul_rn(ga, 1.0f / 2688.0f);
  return gs > 0.0f ? gs : 1e-12f;
}

template <bool kIeee>
__device__ __forceinline__ float block_scale(float bmax, float gs, float six_gs, const DivFull& dsix) {
  // common/nvfp4_quant.py:121-125 and fp4_kernel_hopper.py:83-84; dsix = DivFull(six_gs), hoisted by the caller
  float t = kIeee ? __fdiv_rn(bmax, six_gs) : dsix(bmax);
  t = fminf(t, 448.0f);  // min.f32: NaN operand dropped
  float s = __fmul_rn(gs, e4m3_roundtrip1(t));
  return s >= 1e-5f ? s : 1.0f;
}

// Quantize-dequantize N values that share `s`; f in/out (fp32).
template <int N, bool kIeee>
__device__ __forceinline__ void qdq_block_values(float* f, float s) {
  if constexpr (kIeee) {
    const DivSeq2 dv(s);  // the division's own instruction sequence, reciprocal hoisted, packed fp32
#pragma unroll
    for (int i = 0; i < N; i += 2) {
      float a = fabsf(f[i]), b = fabsf(f[i + 1]);
      if (dv.ok) dv(a, b);
      else { a = __fdiv_rn(a, s); b = __fdiv_rn(b, s); }
      e2m1_roundtrip2(a, b);
      const float oa = __fmul_rn(s, a), ob = __fmul_rn(s, b);
      // reference: where(x >= 0, o, fma(-s, q, 0)); fma(-s,q,+0) == 0 - o (gives +0 for o == 0)
      f[i] = f[i] >= 0.0f ? oa : __fsub_rn(0.0f, oa);
      f[i + 1] = f[i + 1] >= 0.0f ? ob : __fsub_rn(0.0f, ob);
    }
  } else {
    const DivFull df(s);  // div.full.f32 with the range test and MUFU.RCP once per block: one FMUL per quotient
    // warp-uniform test so the common case is a real branch (a per-thread test was compiled into predication of BOTH
    // forms: 99 predicated instructions per 32 elements)
    if (__all_sync(__activemask(), df.plain())) {  // callers may be in a divergent tail: vote among the active lanes
#pragma unroll
      for (int i = 0; i < N; i += 2) {
        float a = df.fast(fabsf(f[i])), b = df.fast(fabsf(f[i + 1]));
        e2m1_roundtrip2(a, b);
        const float oa = __fmul_rn(s, a), ob = __fmul_rn(s, b);
        f[i] = f[i] >= 0.0f ? oa : __fsub_rn(0.0f, oa);
        f[i + 1] = f[i + 1] >= 0.0f ? ob : __fsub_rn(0.0f, ob);
      }
    } else {
#pragma unroll
      for (int i = 0; i < N; i += 2) {
        float a = df(fabsf(f[i])), b = df(fabsf(f[i + 1]));
        e2m1_roundtrip2(a, b);
        const float oa = __fmul_rn(s, a), ob = __fmul_rn(s, b);
        f[i] = f[i] >= 0.0f ? oa : __fsub_rn(0.0f, oa);
        f[i + 1] = f[i + 1] >= 0.0f ? ob : __fsub_rn(0.0f, ob);
      }
    }
  }
}

template <int N>
__device__ __forceinline__ float abs_max_drop_nan(const float* f) {
  // max.f32 semantics (NaN dropped); a balanced tree instead of a 15-deep dependent chain (max is associative, so the
  // result is the same): the chain's latency was exposed once per block
  float m[N];
#pragma unroll
  for (int i = 0; i < N; ++i) m[i] = fabsf(f[i]);
#pragma unroll
  for (int w = N / 2; w >= 1; w /= 2) {
#pragma unroll
    for (int i = 0; i < w; ++i) m[i] = fmaxf(m[i], m[i + w]);
  }
  return m[0];
}

// ------------------------------------------------------------------------------------------------
// Dynamic, vector path: cols % BLOCK == 0, 32-byte aligned.  NV = vectors per NVFP4 block.
// ------------------------------------------------------------------------------------------------
template <typename T, int BLOCK, bool kIeee>
__global__ void __launch_bounds__(kThreads) nvfp4_dynamic_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                 int64_t n_blocks,
                                                                 const float* __restrict__ global_amax) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = BLOCK / E;
  static_assert(NV >= 1 && BLOCK % E == 0, "block must be a whole number of 32-byte vectors");
  const float gs = global_scale_from_amax<kIeee>(__ldg(global_amax));
  const float six_gs = __fmul_rn(gs, 6.0f);
  const DivFull dsix(six_gs);
  const V32* xv = reinterpret_cast<const V32*>(x);
  V32* yv = reinterpret_cast<V32*>(y);
  const int64_t tile_blocks = (int64_t)kThreads * kUnrollB;
  const int64_t n_tiles = (n_blocks + tile_blocks - 1) / tile_blocks;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t base = tile * tile_blocks + threadIdx.x;
    V32 v[kUnrollB][NV];
#pragma unroll
    for (int u = 0; u < kUnrollB; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
#pragma unroll
        for (int k = 0; k < NV; ++k) v[u][k] = ld_stream(xv + b * NV + k);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnrollB; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
        float f[BLOCK];
#pragma unroll
        for (int k = 0; k < NV; ++k) Elem<T>::unpack(v[u][k], f + k * E);
        const float s = block_scale<kIeee>(abs_max_drop_nan<BLOCK>(f), gs, six_gs, dsix);
        qdq_block_values<BLOCK, kIeee>(f, s);
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          Elem<T>::pack(f + k * E, v[u][k]);
          st_stream(yv + b * NV + k, v[u][k]);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Generic path (ragged cols, unaligned pointers, odd block sizes): one thread per (row, block),
// scalar loads, a short last block is zero-padded like the reference's boundary-checked tile load.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float ld_any(const void* p, int64_t i) { return Elem<T>::load(static_cast<const T*>(p) + i); }

template <typename T, bool kIeee>
__global__ void __launch_bounds__(kThreads) nvfp4_dynamic_generic_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                         int64_t rows, int64_t cols, int block,
                                                                         const float* __restrict__ global_amax) {
  const float gs = global_scale_from_amax<kIeee>(__ldg(global_amax));
  const float six_gs = __fmul_rn(gs, 6.0f);
  const DivFull dsix(six_gs);
  const int64_t bpr = (cols + block - 1) / block;
  const int64_t total = rows * bpr;
  for (int64_t t = (int64_t)blockIdx.x * kThreads + threadIdx.x; t < total; t += (int64_t)gridDim.x * kThreads) {
    const int64_t r = t / bpr, c0 = (t % bpr) * block;
    const int len = (int)min((int64_t)block, cols - c0);
    const T* xp = x + r * cols + c0;
    T* yp = y + r * cols + c0;
    float m = fabsf(Elem<T>::load(xp));
    for (int i = 1; i < len; ++i) m = fmaxf(m, fabsf(Elem<T>::load(xp + i)));
    if (len < block) m = fmaxf(m, 0.0f);  // zero padding takes part in the max
    const float s = block_scale<kIeee>(m, gs, six_gs, dsix);
    for (int i = 0; i < len; ++i) {
      float f[2] = {Elem<T>::load(xp + i), 0.0f};
      qdq_block_values<2, kIeee>(f, s);
      Elem<T>::store(yp + i, f[0]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Static: per-block scale given.  Generic over (Tin, Tout, runtime block); vector fast path when
// Tin == Tout, BLOCK in {16, 32, 64} and pointers are 32-byte aligned.
// ------------------------------------------------------------------------------------------------
struct StaticScale {
  float safe;
  bool zero;
  __device__ __forceinline__ explicit StaticScale(float s) {
    zero = s == 0.0f;  // common/nvfp4_quant.py:90-96
    safe = (zero || isnan(s) || isinf(s)) ? 1.0f : s;
  }
};

// kFromAmax: `scale` holds the per-block AMAX and the scale is derived here (compute_fp4_scales folded in: no separate
// pass that reads n_blocks amax values and writes n_blocks scales)
template <typename T, int BLOCK, bool kIeee, bool kFromAmax>
__global__ void __launch_bounds__(kThreads) nvfp4_static_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                int64_t n_blocks, const float* __restrict__ scale,
                                                                const float* __restrict__ global_amax, int quantize,
                                                                float fp8_max) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = BLOCK / E;
  const Fp4ScaleOf<kIeee> of(global_amax, kFromAmax ? quantize : 0, fp8_max);
  const V32* xv = reinterpret_cast<const V32*>(x);
  V32* yv = reinterpret_cast<V32*>(y);
  const int64_t tile_blocks = (int64_t)kThreads * kUnrollB;
  const int64_t n_tiles = (n_blocks + tile_blocks - 1) / tile_blocks;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t base = tile * tile_blocks + threadIdx.x;
    V32 v[kUnrollB][NV];
    float sc[kUnrollB];
#pragma unroll
    for (int u = 0; u < kUnrollB; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
        sc[u] = __ldg(scale + b);
#pragma unroll
        for (int k = 0; k < NV; ++k) v[u][k] = ld_stream(xv + b * NV + k);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnrollB; ++u) {
      const int64_t b = base + (int64_t)u * kThreads;
      if (b < n_blocks) {
        float f[BLOCK];
#pragma unroll
        for (int k = 0; k < NV; ++k) Elem<T>::unpack(v[u][k], f + k * E);
        const StaticScale ss(kFromAmax ? of(sc[u]) : sc[u]);
        qdq_block_values<BLOCK, kIeee>(f, ss.safe);
        if (ss.zero) {
#pragma unroll
          for (int i = 0; i < BLOCK; ++i) f[i] = 0.0f;
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          Elem<T>::pack(f + k * E, v[u][k]);
          st_stream(yv + b * NV + k, v[u][k]);
        }
      }
    }
  }
}

template <typename Tin, typename Tout, bool kIeee>
__global__ void __launch_bounds__(kThreads) nvfp4_static_generic_kernel(const Tin* __restrict__ x, Tout* __restrict__ y,
                                                                        int64_t n_blocks, int block,
                                                                        const float* __restrict__ scale) {
  const int64_t n = n_blocks * block;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    const StaticScale ss(__ldg(scale + i / block));
    float f[2] = {Elem<Tin>::load(x + i), 0.0f};
    qdq_block_values<2, kIeee>(f, ss.safe);
    Elem<Tout>::store(y + i, ss.zero ? 0.0f : f[0]);
  }
}

// fp4 cast (no scale): gemm/fp4_kernel.py:319-363
template <typename Tin, typename Tout>
__global__ void __launch_bounds__(kThreads) fp4_cast_kernel(const Tin* __restrict__ x, Tout* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    const float v = Elem<Tin>::load(x + i);
    float a = fabsf(v), z = 0.0f;
    e2m1_roundtrip2(a, z);
    Elem<Tout>::store(y + i, v >= 0.0f ? a : -a);
  }
}

template <bool kIeee>
__global__ void __launch_bounds__(kThreads) fp4_scales_kernel(const float* __restrict__ amax, float* __restrict__ scale,
                                                              int64_t n, const float* __restrict__ global_amax,
                                                              int quantize, float fp8_max) {
  const Fp4ScaleOf<kIeee> of(global_amax, quantize, fp8_max);
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    scale[i] = of(amax[i]);
}

__global__ void __launch_bounds__(kThreads) absmax_f32_kernel(const float* __restrict__ a, int64_t n, uint32_t* out) {
  uint32_t m = 0;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
    m = max(m, __float_as_uint(a[i]) & 0x7fffffffu);
  m = block_max_u32(m);
  if (threadIdx.x == 0) atomicMax(out, m);
}

template <typename T, bool kIeee>
static int launch_dynamic(const void* x, void* y, int64_t rows, int64_t cols, const float* ga, int block,
                          cudaStream_t s) {
  constexpr int E = Elem<T>::kPerVec;
  const T* xp = static_cast<const T*>(x);
  T* yp = static_cast<T*>(y);
  const bool vec = cols % block == 0 && aligned32(x) && aligned32(y) && block % E == 0;
  if (vec) {
    const int64_t nb = rows * (cols / block);
    const int64_t tiles = (nb + (int64_t)kThreads * kUnrollB - 1) / ((int64_t)kThreads * kUnrollB);
#define MO_DYN(B) nvfp4_dynamic_kernel<T, B, kIeee><<<resident_grid((const void*)nvfp4_dynamic_kernel<T, B, kIeee>, tiles), kThreads, 0, s>>>(xp, yp, nb, ga)
    if (block == 16) MO_DYN(16);
    else if (block == 32) MO_DYN(32);
    else MO_DYN(64);
#undef MO_DYN
  } else {
    const int64_t total = rows * ((cols + block - 1) / block);
    const int grid = persistent_grid((total + kThreads - 1) / kThreads, 8);
    nvfp4_dynamic_generic_kernel<T, kIeee><<<grid, kThreads, 0, s>>>(xp, yp, rows, cols, block, ga);
  }
  return check_launch("mo_qdq_nvfp4_dynamic");
}

template <typename Tin, typename Tout, bool kIeee>
static int launch_static(const void* x, void* y, int64_t nb, int block, const float* scale, cudaStream_t s,
                         bool from_amax = false, const float* ga = nullptr, int quantize = 0, float fp8_max = 448.f) {
  const Tin* xp = static_cast<const Tin*>(x);
  Tout* yp = static_cast<Tout*>(y);
  if constexpr (std::is_same<Tin, Tout>::value) {
    constexpr int E = Elem<Tin>::kPerVec;
    if (aligned32(x) && aligned32(y) && (block == 16 || block == 32 || block == 64) && block % E == 0) {
      const int64_t tiles = (nb + (int64_t)kThreads * kUnrollB - 1) / ((int64_t)kThreads * kUnrollB);
#define MO_STA(B, F) nvfp4_static_kernel<Tin, B, kIeee, F><<<resident_grid((const void*)nvfp4_static_kernel<Tin, B, kIeee, F>, tiles), kThreads, 0, s>>>(xp, yp, nb, scale, ga, quantize, fp8_max)
      if (from_amax) {
        if (block == 16) MO_STA(16, true);
        else if (block == 32) MO_STA(32, true);
        else MO_STA(64, true);
      } else {
        if (block == 16) MO_STA(16, false);
        else if (block == 32) MO_STA(32, false);
        else MO_STA(64, false);
      }
#undef MO_STA
      return check_launch("mo_qdq_nvfp4_static");
    }
  }
  if (from_amax) return MO_ERR_ALIGN;  // the caller falls back to mo_fp4_scales + the generic kernel
  const int grid = persistent_grid((nb * block + kThreads - 1) / kThreads, 8);
  nvfp4_static_generic_kernel<Tin, Tout, kIeee><<<grid, kThreads, 0, s>>>(xp, yp, nb, block, scale);
  return check_launch("mo_qdq_nvfp4_static(generic)");
}

template <typename Tin, bool kIeee>
static int static_out_dispatch(const void* x, void* y, int64_t nb, int block, int out_dtype, const float* scale,
                               cudaStream_t s) {
  switch (out_dtype) {
    case MO_F32: return launch_static<Tin, float, kIeee>(x, y, nb, block, scale, s);
    case MO_F16: return launch_static<Tin, __half, kIeee>(x, y, nb, block, scale, s);
    case MO_BF16: return launch_static<Tin, __nv_bfloat16, kIeee>(x, y, nb, block, scale, s);
  }
  set_error("mo_qdq_nvfp4_static: unsupported out dtype %d", out_dtype);
  return MO_ERR_DTYPE;
}

template <typename Tin>
static int cast_out_dispatch(const void* x, void* y, int64_t n, int out_dtype, cudaStream_t s) {
  const int grid = persistent_grid((n + kThreads - 1) / kThreads, 8);
  const Tin* xp = static_cast<const Tin*>(x);
  switch (out_dtype) {
    case MO_F32: fp4_cast_kernel<Tin, float><<<grid, kThreads, 0, s>>>(xp, static_cast<float*>(y), n); break;
    case MO_F16: fp4_cast_kernel<Tin, __half><<<grid, kThreads, 0, s>>>(xp, static_cast<__half*>(y), n); break;
    case MO_BF16: fp4_cast_kernel<Tin, __nv_bfloat16><<<grid, kThreads, 0, s>>>(xp, static_cast<__nv_bfloat16*>(y), n); break;
    default: set_error("mo_fp4_cast: unsupported out dtype %d", out_dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_fp4_cast");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_qdq_nvfp4_dynamic(const void* x, void* y, int64_t rows, int64_t cols, int dtype,
                                    const float* global_amax, int block, int div_mode, void* stream) {
  if (rows == 0 || cols == 0) return MO_OK;
  if (!x || !y || !global_amax) { set_error("mo_qdq_nvfp4_dynamic: NULL pointer"); return MO_ERR_NULL; }
  if (rows < 0 || cols < 0) { set_error("mo_qdq_nvfp4_dynamic: negative size"); return MO_ERR_SHAPE; }
  if (block != 16 && block != 32 && block != 64) {
    set_error("mo_qdq_nvfp4_dynamic: block=%d not in {16,32,64}", block);
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool ieee = div_mode == MO_DIV_IEEE;
  switch (dtype) {
    case MO_F32: return ieee ? launch_dynamic<float, true>(x, y, rows, cols, global_amax, block, s)
                             : launch_dynamic<float, false>(x, y, rows, cols, global_amax, block, s);
    case MO_F16: return ieee ? launch_dynamic<__half, true>(x, y, rows, cols, global_amax, block, s)
                             : launch_dynamic<__half, false>(x, y, rows, cols, global_amax, block, s);
    case MO_BF16: return ieee ? launch_dynamic<__nv_bfloat16, true>(x, y, rows, cols, global_amax, block, s)
                              : launch_dynamic<__nv_bfloat16, false>(x, y, rows, cols, global_amax, block, s);
  }
  set_error("mo_qdq_nvfp4_dynamic: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_qdq_nvfp4_static(const void* x, void* y, int64_t n_blocks, int block, int in_dtype, int out_dtype,
                                   const float* scale, int div_mode, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !y || !scale) { set_error("mo_qdq_nvfp4_static: NULL pointer"); return MO_ERR_NULL; }
  if (block < 1 || n_blocks < 0) { set_error("mo_qdq_nvfp4_static: bad block/n_blocks"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool ieee = div_mode == MO_DIV_IEEE;
  switch (in_dtype) {
    case MO_F32: return ieee ? static_out_dispatch<float, true>(x, y, n_blocks, block, out_dtype, scale, s)
                             : static_out_dispatch<float, false>(x, y, n_blocks, block, out_dtype, scale, s);
    case MO_F16: return ieee ? static_out_dispatch<__half, true>(x, y, n_blocks, block, out_dtype, scale, s)
                             : static_out_dispatch<__half, false>(x, y, n_blocks, block, out_dtype, scale, s);
    case MO_BF16: return ieee ? static_out_dispatch<__nv_bfloat16, true>(x, y, n_blocks, block, out_dtype, scale, s)
                              : static_out_dispatch<__nv_bfloat16, false>(x, y, n_blocks, block, out_dtype, scale, s);
  }
  set_error("mo_qdq_nvfp4_static: unsupported in dtype %d", in_dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_fp4_cast(const void* x, void* y, int64_t n, int in_dtype, int out_dtype, void* stream) {
  if (n == 0) return MO_OK;
  if (!x || !y) { set_error("mo_fp4_cast: NULL pointer"); return MO_ERR_NULL; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (in_dtype) {
    case MO_F32: return cast_out_dispatch<float>(x, y, n, out_dtype, s);
    case MO_F16: return cast_out_dispatch<__half>(x, y, n, out_dtype, s);
    case MO_BF16: return cast_out_dispatch<__nv_bfloat16>(x, y, n, out_dtype, s);
  }
  set_error("mo_fp4_cast: unsupported in dtype %d", in_dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_fp4_scales(const float* amax, float* scale, int64_t n_blocks, const float* global_amax,
                             int quantize_block_scales, float fp8_max_for_normalization, int div_mode,
                             uint32_t* scratch, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!amax || !scale) { set_error("mo_fp4_scales: NULL pointer"); return MO_ERR_NULL; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = persistent_grid((n_blocks + kThreads - 1) / kThreads, 4);
  if (quantize_block_scales && global_amax == nullptr) {
    // global_amax = reduce_amax(amax): needs a 4-byte device scratch word
    if (!scratch) { set_error("mo_fp4_scales: scratch is required when global_amax is NULL"); return MO_ERR_NULL; }
    cudaMemsetAsync(scratch, 0, 4, s);
    absmax_f32_kernel<<<grid, kThreads, 0, s>>>(amax, n_blocks, scratch);
    global_amax = reinterpret_cast<const float*>(scratch);
  }
  if (div_mode == MO_DIV_IEEE)
    fp4_scales_kernel<true><<<grid, kThreads, 0, s>>>(amax, scale, n_blocks, global_amax, quantize_block_scales,
                                                      fp8_max_for_normalization);
  else
    fp4_scales_kernel<false><<<grid, kThreads, 0, s>>>(amax, scale, n_blocks, global_amax, quantize_block_scales,
                                                       fp8_max_for_normalization);
  return check_launch("mo_fp4_scales");
}

// Fused form: per-block amax in, scale derived in the kernel (one launch instead of mo_fp4_scales + mo_qdq_nvfp4_static).
// Returns MO_ERR_ALIGN (without launching) when the vector path does not apply: the caller uses the two-call form.
template <typename T>
static int static_amax_dispatch(const void* x, void* y, int64_t nb, int block, const float* amax, const float* ga, int q,
                                float fp8_max, bool ieee, cudaStream_t s) {
  return ieee ? launch_static<T, T, true>(x, y, nb, block, amax, s, true, ga, q, fp8_max)
              : launch_static<T, T, false>(x, y, nb, block, amax, s, true, ga, q, fp8_max);
}

extern "C" int mo_qdq_nvfp4_static_amax(const void* x, void* y, int64_t n_blocks, int block, int dtype, const float* amax,
                                        const float* global_amax, int quantize_block_scales,
                                        float fp8_max_for_normalization, int div_mode, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !y || !amax || (quantize_block_scales && !global_amax)) {
    set_error("mo_qdq_nvfp4_static_amax: NULL pointer");
    return MO_ERR_NULL;
  }
  if (block != 16 && block != 32 && block != 64) return MO_ERR_ALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool ieee = div_mode == MO_DIV_IEEE;
  int rc;
  switch (dtype) {
    case MO_F32: rc = static_amax_dispatch<float>(x, y, n_blocks, block, amax, global_amax, quantize_block_scales, fp8_max_for_normalization, ieee, s); break;
    case MO_F16: rc = static_amax_dispatch<__half>(x, y, n_blocks, block, amax, global_amax, quantize_block_scales, fp8_max_for_normalization, ieee, s); break;
    case MO_BF16: rc = static_amax_dispatch<__nv_bfloat16>(x, y, n_blocks, block, amax, global_amax, quantize_block_scales, fp8_max_for_normalization, ieee, s); break;
    default: set_error("mo_qdq_nvfp4_static_amax: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
  if (rc == MO_ERR_ALIGN) set_error("mo_qdq_nvfp4_static_amax: vector path not applicable (alignment / block)");
  return rc;
}
