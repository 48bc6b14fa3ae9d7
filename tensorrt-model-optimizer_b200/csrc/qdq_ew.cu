// Elementwise fake quantize-dequantize: INT-N and FP8 E4M3, per-tensor / per-axis / per-block.
//
// Replaces fake_tensor_quant_kernel / fake_tensor_quant_with_axis_cuda_kernel
// (reference gemm/tensor_quant_gpu.cu:58-140) and fake_e4m3fy_kernel / ..._with_axis_cuda_kernel
// (gemm/tensor_quant_gpu_fp8.cu:35-107), including the host-side torch ops that build the FP8
// scale pair.  HBM-bound: 4 algorithmic bytes per bf16 element (2 read + 2 written).
//
// Shape of the kernel (B200): persistent grid (a multiple of 148 SMs), 256 threads, every thread
// moves kUnroll independent 32-byte vectors per iteration with LDG.E.256 / STG.E.256 so that
// >= 64 KB per SM is in flight; no shared memory (no reuse); math in fp32 in exactly the reference's
// operation order (x*scale -> round -> clamp -> q/scale; never x*(1/scale) for INT).
#include "common.cuh"

namespace mo {

constexpr int kUnroll = 2;

// How amax is indexed.
enum AxisMode { kPerTensor = 0, kPerChunk = 1, kPerElem = 2 };

struct AxisArgs {
  const float* amax;
  int64_t axis_size;
  int64_t inner;
  int mode;
  // channel of element idx
  __device__ __forceinline__ int64_t channel(int64_t idx) const {
    if (idx < (int64_t(1) << 32) && inner < (int64_t(1) << 32) && axis_size < (int64_t(1) << 32)) {
      return (uint32_t)((uint32_t)idx / (uint32_t)inner) % (uint32_t)axis_size;
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
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0) {
    if (ax.mode == kPerChunk) set(__ldg(ax.amax + ax.channel(idx0)));
    if (ax.mode == kPerElem) {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = one(f[i], idx0 + i);
      return;
    }
#pragma unroll
    for (int i = 0; i < N; i += 2) pair(f[i], f[i + 1]);
  }
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

template <bool kUnsigned>
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
  __device__ __forceinline__ float q1(float x) {
    if (kUnsigned && x < 0.f) flags |= MO_FLAG_NEGATIVE_IN_UNSIGNED;
    float v = __fmul_rn(x, scale);
    // clamp first (bounds are integers, so clamp and rint commute), NaN kept like the reference's
    // `output > max ? max : output` selects (:52-53)
    v = max_nan(min_nan(v, maxb), minb);
    // rint for |v| < 2^22: add/subtract 1.5*2^23 (round-half-even in the fp32 adder)
    v = __fsub_rn(__fadd_rn(v, 12582912.0f), 12582912.0f);
    float o;
    if (dfast) {
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
  template <int N>
  __device__ __forceinline__ void vec(float* f, int64_t idx0) {
    if (ax.mode == kPerChunk) set(__ldg(ax.amax + ax.channel(idx0)));
    if (ax.mode == kPerElem) {
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] = one(f[i], idx0 + i);
      return;
    }
#pragma unroll
    for (int i = 0; i < N; ++i) f[i] = q1(f[i]);
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
template <bool U>
__device__ __forceinline__ void finish_op(IntOp<U>& op) { op.finish(); }

// ------------------------------------------------------------------------------------------------
// The streaming kernel.  kVec = true: 32-byte vectors (x, y 32-byte aligned); tail and the
// unaligned case run the scalar path.
// ------------------------------------------------------------------------------------------------
template <typename T, typename Op>
__global__ void __launch_bounds__(kThreads) qdq_stream_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                              int64_t n, Op op, bool vectorized) {
  constexpr int E = Elem<T>::kPerVec;
  op.init();
  if (vectorized) {
    const int64_t n_vec = n / E;
    const int64_t tile_vecs = (int64_t)kThreads * kUnroll;
    const int64_t n_tiles = (n_vec + tile_vecs - 1) / tile_vecs;
    const V32* xv = reinterpret_cast<const V32*>(x);
    V32* yv = reinterpret_cast<V32*>(y);
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int64_t base = tile * tile_vecs + threadIdx.x;
      V32 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t c = base + (int64_t)u * kThreads;
        if (c < n_vec) v[u] = ld_stream(xv + c);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t c = base + (int64_t)u * kThreads;
        if (c < n_vec) {
          float f[E];
          Elem<T>::unpack(v[u], f);
          op.template vec<E>(f, c * E);
          Elem<T>::pack(f, v[u]);
          st_stream(yv + c, v[u]);
        }
      }
    }
    // tail (< E elements)
    if (blockIdx.x == 0) {
      const int64_t i = n_vec * E + threadIdx.x;
      if (i < n) Elem<T>::store(y + i, op.one(Elem<T>::load(x + i), i));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
      Elem<T>::store(y + i, op.one(Elem<T>::load(x + i), i));
  }
  finish_op(op);
}

template <typename T, typename Op>
static int launch_stream(const void* x, void* y, int64_t n, Op op, cudaStream_t stream, const char* what) {
  constexpr int E = Elem<T>::kPerVec;
  const bool vec = aligned32(x) && aligned32(y) && n >= E;
  int64_t tiles = vec ? (n / E + (int64_t)kThreads * kUnroll - 1) / ((int64_t)kThreads * kUnroll)
                      : (n + kThreads - 1) / kThreads;
  const int grid = persistent_grid(tiles, 8);
  qdq_stream_kernel<T, Op><<<grid, kThreads, 0, stream>>>(static_cast<const T*>(x), static_cast<T*>(y), n, op, vec);
  return check_launch(what);
}

static int make_axis(AxisArgs& ax, const float* amax, int64_t n, int64_t axis_size, int64_t inner, int per_vec) {
  if (axis_size < 1 || inner < 1) { set_error("axis_size and inner must be >= 1"); return MO_ERR_SHAPE; }
  if (axis_size == 1) inner = n;
  if (n % (axis_size * inner) != 0) {
    set_error("n=%lld is not a multiple of axis_size*inner=%lld", (long long)n, (long long)(axis_size * inner));
    return MO_ERR_SHAPE;
  }
  ax.amax = amax;
  ax.axis_size = axis_size;
  ax.inner = inner;
  ax.mode = axis_size == 1 ? kPerTensor : (inner % per_vec == 0 ? kPerChunk : kPerElem);
  return MO_OK;
}

}  // namespace mo

using namespace mo;

extern "C" int mo_qdq_fp8(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                          int64_t inner, int numerics, void* stream) {
  if (n == 0) return MO_OK;
  if (!x || !y) { set_error("mo_qdq_fp8: x/y is NULL"); return MO_ERR_NULL; }
  Fp8Op op;
  op.plain = amax == nullptr;
  op.eager = numerics == MO_NUMERICS_CPU;
  if (op.plain) { axis_size = 1; inner = n; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  switch (dtype) {
    case MO_F32:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 8))) return rc;
                  return launch_stream<float, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<f32>");
    case MO_F16:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__half, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<f16>");
    case MO_BF16: if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__nv_bfloat16, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<bf16>");
  }
  set_error("mo_qdq_fp8: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

template <bool U>
static int qdq_int_impl(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                        int64_t inner, int num_bits, int narrow_range, uint32_t* status, cudaStream_t s) {
  IntOp<U> op;
  const float bound = (float)((1 << (num_bits - 1 + (U ? 1 : 0))) - 1);  // tensor_quant_gpu.cu:38-41
  op.maxb = bound;
  op.minb = -(bound + (narrow_range ? 0.f : 1.f));                       // :69
  op.status = status;
  int rc;
  switch (dtype) {
    case MO_F32:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 8))) return rc;
                  return launch_stream<float, IntOp<U>>(x, y, n, op, s, "mo_qdq_int<f32>");
    case MO_F16:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__half, IntOp<U>>(x, y, n, op, s, "mo_qdq_int<f16>");
    case MO_BF16: if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__nv_bfloat16, IntOp<U>>(x, y, n, op, s, "mo_qdq_int<bf16>");
  }
  set_error("mo_qdq_int: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_qdq_int(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                          int64_t inner, int num_bits, int is_unsigned, int narrow_range, uint32_t* status,
                          void* stream) {
  if (n == 0) return MO_OK;
  if (!x || !y || !amax) { set_error("mo_qdq_int: x/y/amax is NULL"); return MO_ERR_NULL; }
  if (num_bits < 1 || num_bits + (is_unsigned ? 1 : 0) > 22) {
    set_error("mo_qdq_int: num_bits=%d out of range [1, 22 - unsigned]", num_bits);
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return is_unsigned ? qdq_int_impl<true>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s)
                     : qdq_int_impl<false>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s);
}
