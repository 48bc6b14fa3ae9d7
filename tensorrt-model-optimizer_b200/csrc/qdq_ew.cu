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
#include <type_traits>

#include "qdq_ops.cuh"

namespace mo {

constexpr int kUnroll = 2;

// ------------------------------------------------------------------------------------------------
// The streaming kernel.  kVec = true: 32-byte vectors (x, y 32-byte aligned); tail and the
// unaligned case run the scalar path.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float prescaled(const T* x, int64_t i, const T* pre, int64_t cols) {
  float v = Elem<T>::load(x + i);
  if (pre != nullptr) {
    T t;
    Elem<T>::store(&t, __fmul_rn(v, Elem<T>::load(pre + i % cols)));
    v = Elem<T>::load(&t);
  }
  return v;
}

template <typename T, typename Op, bool kPre>
__global__ void __launch_bounds__(kThreads) qdq_stream_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                              int64_t n, Op op, bool vectorized,
                                                              const T* __restrict__ pre_arg, int64_t cols) {
  const T* pre = kPre ? pre_arg : nullptr;  // compile-time NULL: the plain kernels carry no prescale code
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
      float am[kUnroll];  // per-channel / per-block amax fetched WITH the data, not after it arrived
      const bool chunk = op.ax.mode == kPerChunk;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t c = base + (int64_t)u * kThreads;
        am[u] = 0.f;
        if (c < n_vec) {
          v[u] = ld_stream(xv + c);
          if (chunk) am[u] = __ldg(op.ax.amax + op.ax.channel(c * E));
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t c = base + (int64_t)u * kThreads;
        if (c < n_vec) {
          float f[E];
          Elem<T>::unpack(v[u], f);
          if constexpr (kPre) {  // inputs * pre_quant_scale, a T tensor in the reference (tensor_quantizer.py:1143)
            // the scale vector travels in the tensor's dtype (its values ARE of that dtype): one 32-byte load per vector
            float sc[E];
            Elem<T>::unpack(ld_cached(pre + (c * E) % cols), sc);
#pragma unroll
            for (int i = 0; i < E; ++i) f[i] = __fmul_rn(f[i], sc[i]);
            Elem<T>::pack(f, v[u]);
            Elem<T>::unpack(v[u], f);
          }
          op.template vec<E>(f, c * E, am[u], chunk);
          Elem<T>::pack(f, v[u]);
          st_stream(yv + c, v[u]);
        }
      }
    }
    // tail (< E elements)
    if (blockIdx.x == 0) {
      const int64_t i = n_vec * E + threadIdx.x;
      if (i < n) Elem<T>::store(y + i, op.one(prescaled<T>(x, i, pre, cols), i));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
      Elem<T>::store(y + i, op.one(prescaled<T>(x, i, pre, cols), i));
  }
  finish_op(op);
}

// pre_quant_scale + fake-quant, row-block tiling: a thread owns ONE column vector of the scale (16 values in registers,
// read once per tile) and walks kPreRows rows of it, two rows in flight.  The streaming kernel above re-read the 16 scale
// values through L1 for every vector (as many bytes as the data itself for bf16).
constexpr int kPreRows = 8;

template <typename T, typename Op>
__global__ void __launch_bounds__(kThreads) qdq_pre_rows_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows,
                                                                int64_t vpr, Op op, const T* __restrict__ pre) {
  constexpr int E = Elem<T>::kPerVec;
  op.init();
  const V32* xv = reinterpret_cast<const V32*>(x);
  V32* yv = reinterpret_cast<V32*>(y);
  const int64_t col_tiles = (vpr + kThreads - 1) / kThreads, row_tiles = (rows + kPreRows - 1) / kPreRows;
  const bool chunk = op.ax.mode == kPerChunk;
  for (int64_t tile = blockIdx.x; tile < col_tiles * row_tiles; tile += gridDim.x) {
    const int64_t ct = tile % col_tiles, rt = tile / col_tiles;
    const int64_t j = ct * kThreads + threadIdx.x;
    if (j >= vpr) continue;
    float sc[E];
    Elem<T>::unpack(ld_cached(pre + j * E), sc);
    const int64_t r0 = rt * kPreRows, r1 = min(rows, r0 + (int64_t)kPreRows);
    for (int64_t r = r0; r < r1; r += 2) {
      V32 v[2];
      float am[2] = {0.f, 0.f};
      int64_t c[2] = {r * vpr + j, (r + 1) * vpr + j};
      const bool ok[2] = {true, r + 1 < r1};
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (ok[u]) {
          v[u] = ld_stream(xv + c[u]);
          if (chunk) am[u] = __ldg(op.ax.amax + op.ax.channel(c[u] * E));
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (!ok[u]) continue;
        float f[E];
        Elem<T>::unpack(v[u], f);
#pragma unroll
        for (int i = 0; i < E; ++i) f[i] = __fmul_rn(f[i], sc[i]);
        Elem<T>::pack(f, v[u]);  // inputs * pre_quant_scale is a T tensor in the reference (tensor_quantizer.py:1143)
        Elem<T>::unpack(v[u], f);
        op.template vec<E>(f, c[u] * E, am[u], chunk);
        Elem<T>::pack(f, v[u]);
        st_stream(yv + c[u], v[u]);
      }
    }
  }
  finish_op(op);
}

template <typename T, typename Op>
static int launch_stream(const void* x, void* y, int64_t n, Op op, cudaStream_t stream, const char* what,
                         const void* pre_v = nullptr, int64_t cols = 1) {
  constexpr int E = Elem<T>::kPerVec;
  const T* pre = static_cast<const T*>(pre_v);
  const bool vec = aligned32(x) && aligned32(y) && n >= E && (pre == nullptr || (cols % E == 0 && aligned32(pre)));
  int64_t tiles = vec ? (n / E + (int64_t)kThreads * kUnroll - 1) / ((int64_t)kThreads * kUnroll)
                      : (n + kThreads - 1) / kThreads;
  // FP8 only: the INT operator's longer per-element chain needs the occupancy the 80-register row-block kernel gives up
  // (measured: FP8 86.6 -> 93.4 % of the copy peak, INT8 79.8 -> 72.5 %)
  if (std::is_same<Op, Fp8Op>::value && pre != nullptr && vec && n % cols == 0) {
    const int64_t rows = n / cols, vpr = cols / E;
    const int64_t t2 = ((vpr + kThreads - 1) / kThreads) * ((rows + kPreRows - 1) / kPreRows);
    qdq_pre_rows_kernel<T, Op><<<resident_grid((const void*)qdq_pre_rows_kernel<T, Op>, t2), kThreads, 0, stream>>>(
        static_cast<const T*>(x), static_cast<T*>(y), rows, vpr, op, pre);
    return check_launch(what);
  }
  if (pre != nullptr)
    qdq_stream_kernel<T, Op, true><<<resident_grid((const void*)qdq_stream_kernel<T, Op, true>, tiles), kThreads, 0, stream>>>(static_cast<const T*>(x), static_cast<T*>(y), n, op, vec, pre, cols);
  else
    qdq_stream_kernel<T, Op, false><<<resident_grid((const void*)qdq_stream_kernel<T, Op, false>, tiles), kThreads, 0, stream>>>(static_cast<const T*>(x), static_cast<T*>(y), n, op, vec, nullptr, 1);
  return check_launch(what);
}

// ------------------------------------------------------------------------------------------------
// INT-N with the amax of every `block` consecutive elements taken from the data (TensorQuantizer with block_sizes
// and no calibrated amax: _get_amax -> reduce_amax on the [n/block, block] view, tensor_quantizer.py:736-751), and an
// optional per-column multiply in front (pre_quant_scale, :1143).  One pass: read x once, write y once.
//   p = T(x * pre[col])   amax = max|p| (a T value)   y = T(int_qdq(p, amax))
// Vector kernel: `L = block / E` adjacent lanes own one block (L a power of two <= 32).
// ------------------------------------------------------------------------------------------------
template <typename T, bool kUnsigned>
__global__ void __launch_bounds__(kThreads) qdq_int_blockdyn_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                    int64_t n_vec, int lanes, const T* __restrict__ pre,
                                                                    int64_t cols, IntOp<kUnsigned> op,
                                                                    float* __restrict__ amax_out) {
  constexpr int E = Elem<T>::kPerVec;
  op.flags = 0;
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
      const bool live = c < n_vec;
      float f[E];
      uint32_t m = 0;
      if (live) {
        Elem<T>::unpack(v[u], f);
        if (pre != nullptr) {
          float sc[E];
          Elem<T>::unpack(ld_cached(pre + (c * E) % cols), sc);
#pragma unroll
          for (int i = 0; i < E; ++i) f[i] = __fmul_rn(f[i], sc[i]);
          Elem<T>::pack(f, v[u]);    // the product is a T tensor in the reference
          Elem<T>::unpack(v[u], f);
        }
#pragma unroll
        for (int i = 0; i < E; ++i) m = max(m, __float_as_uint(f[i]) & 0x7fffffffu);
      }
      for (int o = 1; o < lanes; o <<= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (live) {
        const float am = __uint_as_float(m);
        op.set(am);
        if (op.dfast) {
#pragma unroll
          for (int i = 0; i < E; ++i) f[i] = op.template q1t<true>(f[i]);
        } else {
#pragma unroll
          for (int i = 0; i < E; ++i) f[i] = op.template q1t<false>(f[i]);
        }
        Elem<T>::pack(f, v[u]);
        st_stream(yv + c, v[u]);
        if (amax_out != nullptr && (c & (lanes - 1)) == 0) amax_out[c / lanes] = am;
      }
    }
  }
}

// Any block size / alignment: one warp per block, two sweeps over the (L1-resident) block.
template <typename T, bool kUnsigned>
__global__ void __launch_bounds__(kThreads) qdq_int_blockdyn_generic_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                            int64_t n_blocks, int block,
                                                                            const T* __restrict__ pre, int64_t cols,
                                                                            IntOp<kUnsigned> op,
                                                                            float* __restrict__ amax_out) {
  op.flags = 0;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t b = warp; b < n_blocks; b += n_warps) {
    const int64_t e0 = b * block;
    uint32_t m = 0;
    for (int i = lane; i < block; i += 32) {
      float p = Elem<T>::load(x + e0 + i);
      if (pre != nullptr) {
        T t;
        Elem<T>::store(&t, __fmul_rn(p, Elem<T>::load(pre + (e0 + i) % cols)));
        p = Elem<T>::load(&t);
      }
      m = max(m, __float_as_uint(p) & 0x7fffffffu);
    }
    m = warp_max_u32(m);
    const float am = __uint_as_float(m);
    op.set(am);
    for (int i = lane; i < block; i += 32) {
      float p = Elem<T>::load(x + e0 + i);
      if (pre != nullptr) {
        T t;
        Elem<T>::store(&t, __fmul_rn(p, Elem<T>::load(pre + (e0 + i) % cols)));
        p = Elem<T>::load(&t);
      }
      Elem<T>::store(y + e0 + i, op.q1(p));
    }
    if (amax_out != nullptr && lane == 0) amax_out[b] = am;
  }
}

template <typename T, bool U>
static int blockdyn_launch(const void* x, void* y, int64_t n_blocks, int block, const void* pre_v, int64_t cols,
                           int num_bits, int narrow_range, float* amax_out, cudaStream_t s) {
  constexpr int E = Elem<T>::kPerVec;
  IntOp<U> op;
  const T* pre = static_cast<const T*>(pre_v);
  const float bound = (float)((1 << (num_bits - 1 + (U ? 1 : 0))) - 1);
  op.maxb = bound;
  op.minb = -(bound + (narrow_range ? 0.f : 1.f));
  op.status = nullptr;
  op.ax.mode = kPerElem;  // scale is set per block by the kernel
  op.ax.amax = nullptr; op.ax.axis_size = 1; op.ax.inner = 1; op.ax.small = false;
  const int lanes = block / E;
  const bool vec = aligned32(x) && aligned32(y) && block % E == 0 && lanes >= 1 && lanes <= 32 &&
                   (lanes & (lanes - 1)) == 0 && (pre == nullptr || (cols % E == 0 && aligned32(pre)));
  if (vec) {
    const int64_t n_vec = n_blocks * lanes;
    const int64_t tiles = (n_vec + (int64_t)kThreads * kUnroll - 1) / ((int64_t)kThreads * kUnroll);
    qdq_int_blockdyn_kernel<T, U><<<resident_grid((const void*)qdq_int_blockdyn_kernel<T, U>, tiles), kThreads, 0, s>>>(
        static_cast<const T*>(x), static_cast<T*>(y), n_vec, lanes, pre, cols, op, amax_out);
  } else {
    qdq_int_blockdyn_generic_kernel<T, U><<<persistent_grid((n_blocks * 32 + kThreads - 1) / kThreads, 8), kThreads, 0, s>>>(
        static_cast<const T*>(x), static_cast<T*>(y), n_blocks, block, pre, cols, op, amax_out);
  }
  return check_launch("mo_qdq_int_block_dynamic");
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
  ax.small = n < (int64_t(1) << 32) && inner < (int64_t(1) << 32) && axis_size < (int64_t(1) << 32);
  if (ax.small) { ax.div_inner.init((uint32_t)inner); ax.div_axis.init((uint32_t)axis_size); }
  return MO_OK;
}

}  // namespace mo

using namespace mo;

static int check_pre(const void* pre, int64_t cols, int64_t n, const char* what) {
  if (pre && (cols < 1 || n % cols != 0)) { set_error("%s: cols=%lld must divide n=%lld", what, (long long)cols, (long long)n); return MO_ERR_SHAPE; }
  return MO_OK;
}

extern "C" int mo_qdq_fp8_pqs(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                              int64_t inner, int numerics, const void* prescale, int64_t cols, void* stream) {
  if (n == 0) return MO_OK;
  if (!x || !y) { set_error("mo_qdq_fp8: x/y is NULL"); return MO_ERR_NULL; }
  if (int rc0 = check_pre(prescale, cols, n, "mo_qdq_fp8_pqs")) return rc0;
  Fp8Op op;
  op.plain = amax == nullptr;
  op.eager = numerics == MO_NUMERICS_CPU;
  if (op.plain) { axis_size = 1; inner = n; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  switch (dtype) {
    case MO_F32:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 8))) return rc;
                  return launch_stream<float, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<f32>", prescale, cols);
    case MO_F16:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__half, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<f16>", prescale, cols);
    case MO_BF16: if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__nv_bfloat16, Fp8Op>(x, y, n, op, s, "mo_qdq_fp8<bf16>", prescale, cols);
  }
  set_error("mo_qdq_fp8: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_qdq_fp8(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                          int64_t inner, int numerics, void* stream) {
  return mo_qdq_fp8_pqs(x, y, n, dtype, amax, axis_size, inner, numerics, nullptr, 1, stream);
}

template <bool U, bool PT>
static int qdq_int_impl2(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                         int64_t inner, int num_bits, int narrow_range, uint32_t* status, cudaStream_t s,
                         const void* pre, int64_t cols) {
  IntOp<U, PT> op;
  const float bound = (float)((1 << (num_bits - 1 + (U ? 1 : 0))) - 1);  // tensor_quant_gpu.cu:38-41
  op.maxb = bound;
  op.minb = -(bound + (narrow_range ? 0.f : 1.f));                       // :69
  op.status = status;
  int rc;
  switch (dtype) {
    case MO_F32:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 8))) return rc;
                  return launch_stream<float, IntOp<U, PT>>(x, y, n, op, s, "mo_qdq_int<f32>", pre, cols);
    case MO_F16:  if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__half, IntOp<U, PT>>(x, y, n, op, s, "mo_qdq_int<f16>", pre, cols);
    case MO_BF16: if ((rc = make_axis(op.ax, amax, n, axis_size, inner, 16))) return rc;
                  return launch_stream<__nv_bfloat16, IntOp<U, PT>>(x, y, n, op, s, "mo_qdq_int<bf16>", pre, cols);
  }
  set_error("mo_qdq_int: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

// per-tensor amax and no pre_quant_scale: the packed-arithmetic instantiation (with a prescale it measured the same)
template <bool U>
static int qdq_int_impl(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                        int64_t inner, int num_bits, int narrow_range, uint32_t* status, cudaStream_t s,
                        const void* pre, int64_t cols) {
  if (axis_size == 1 && pre == nullptr)
    return qdq_int_impl2<U, true>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s, pre, cols);
  return qdq_int_impl2<U, false>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s, pre, cols);
}

extern "C" int mo_qdq_int_pqs(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                              int64_t inner, int num_bits, int is_unsigned, int narrow_range, uint32_t* status,
                              const void* prescale, int64_t cols, void* stream) {
  if (n == 0) return MO_OK;
  if (int rc0 = check_pre(prescale, cols, n, "mo_qdq_int_pqs")) return rc0;
  if (!x || !y || !amax) { set_error("mo_qdq_int: x/y/amax is NULL"); return MO_ERR_NULL; }
  if (num_bits < 1 || num_bits + (is_unsigned ? 1 : 0) > 22) {
    set_error("mo_qdq_int: num_bits=%d out of range [1, 22 - unsigned]", num_bits);
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return is_unsigned ? qdq_int_impl<true>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s, prescale, cols)
                     : qdq_int_impl<false>(x, y, n, dtype, amax, axis_size, inner, num_bits, narrow_range, status, s, prescale, cols);
}

extern "C" int mo_qdq_int(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
                          int64_t inner, int num_bits, int is_unsigned, int narrow_range, uint32_t* status,
                          void* stream) {
  return mo_qdq_int_pqs(x, y, n, dtype, amax, axis_size, inner, num_bits, is_unsigned, narrow_range, status, nullptr, 1,
                        stream);
}

extern "C" int mo_qdq_int_block_dynamic(const void* x, void* y, int64_t n_blocks, int block, int dtype,
                                        const void* prescale, int64_t cols, int num_bits, int is_unsigned,
                                        int narrow_range, float* amax_out, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !y) { set_error("mo_qdq_int_block_dynamic: x/y is NULL"); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 1) { set_error("mo_qdq_int_block_dynamic: bad n_blocks / block"); return MO_ERR_SHAPE; }
  if (prescale && (cols < 1 || cols % block != 0)) {
    set_error("mo_qdq_int_block_dynamic: cols=%lld must be a positive multiple of block=%d", (long long)cols, block);
    return MO_ERR_SHAPE;
  }
  if (num_bits < 1 || num_bits + (is_unsigned ? 1 : 0) > 22) { set_error("mo_qdq_int_block_dynamic: bad num_bits"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define MO_BD(T) (is_unsigned ? blockdyn_launch<T, true>(x, y, n_blocks, block, prescale, cols, num_bits, narrow_range, amax_out, s) \
                              : blockdyn_launch<T, false>(x, y, n_blocks, block, prescale, cols, num_bits, narrow_range, amax_out, s))
  switch (dtype) {
    case MO_F32: return MO_BD(float);
    case MO_F16: return MO_BD(__half);
    case MO_BF16: return MO_BD(__nv_bfloat16);
  }
#undef MO_BD
  set_error("mo_qdq_int_block_dynamic: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}
