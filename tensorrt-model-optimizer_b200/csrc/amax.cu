// Absolute-max reductions for calibration (MaxCalibrator.collect / reduce_amax / reduce_block_amax).
//
// Replaces the eager torch sequence max + min + 2x abs + maximum (reference
// quantization/utils/core_utils.py:146-183: reads the tensor twice) and the running
// torch.max merge plus three host-syncing asserts of MaxCalibrator.collect (calib/max.py:52-86).
// HBM-bound: one read, 2 algorithmic bytes per bf16 element.
//
// The reduction works on the fp32 BIT PATTERN of |x| as an unsigned integer: for non-negative
// floats the unsigned order is the numeric order, +inf sorts above every finite value and every NaN
// pattern above +inf, so NaN/inf propagate exactly like torch.max and the merge into the running
// amax is a single atomicMax per CTA (per channel) on the calibrator's own state word.
//
// Three kernels:
//   amax_flat_kernel   per-tensor: persistent grid, LDG.E.256, packed 16-bit integer max (VIMNMX),
//                      redux.sync + one atomicMax per CTA.
//   amax_rows_kernel   view [rows, len] reduced along the contiguous `len`: a group of G lanes
//                      (G = 1..32, or a whole CTA) per row; out[row % axis_size].  Covers per-channel
//                      weights (axis 0), block-16/128 amax and every "middle axis" view.
//   amax_cols_kernel   view [rows, cols] reduced over rows (axis = last dim, e.g. SmoothQuant's
//                      per-input-channel activation amax): each thread owns one 32-byte column strip.
#include "common.cuh"

namespace mo {

constexpr int kUnrollA = 4;

// |x| bit patterns of one 32-byte vector folded into a running max `m` kept in the element's own
// packed layout (2 x u16 for 16-bit types, 1 x u32 for fp32).
template <typename T>
__device__ __forceinline__ void fold_abs_max(const V32& v, uint32_t& m) {
  if constexpr (sizeof(T) == 2) {
#pragma unroll
    for (int i = 0; i < 8; ++i) m = __vmaxu2(m, v.w[i] & 0x7fff7fffu);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) m = max(m, v.w[i] & 0x7fffffffu);
  }
}

// Packed running max -> fp32 bit pattern of the max |x|.
template <typename T>
__device__ __forceinline__ uint32_t packed_to_f32_bits(uint32_t m) {
  if constexpr (sizeof(T) == 4) return m;
  const uint32_t h = max(m & 0xffffu, m >> 16);
  if constexpr (Elem<T>::kDType == MO_BF16) return h << 16;  // bf16 -> fp32 is a shift, NaN stays NaN
  // fp16: widen the value; NaN must stay NaN and keep sorting above +inf
  if (h > 0x7c00u) return 0x7fc00000u;
  return __float_as_uint(__half2float(__ushort_as_half((unsigned short)h)));
}

template <typename T>
__device__ __forceinline__ uint32_t abs_bits_scalar(const T* p) {
  return __float_as_uint(Elem<T>::load(p)) & 0x7fffffffu;  // fp16/bf16 -> fp32 keeps NaN-ness
}

// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) amax_flat_kernel(const T* __restrict__ x, int64_t n, bool vectorized,
                                                             uint32_t* __restrict__ out) {
  constexpr int E = Elem<T>::kPerVec;
  uint32_t best = 0;
  if (vectorized) {
    const int64_t n_vec = n / E;
    const V32* xv = reinterpret_cast<const V32*>(x);
    const int64_t tile_vecs = (int64_t)kThreads * kUnrollA;
    const int64_t n_tiles = (n_vec + tile_vecs - 1) / tile_vecs;
    uint32_t m = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int64_t base = tile * tile_vecs + threadIdx.x;
      V32 v[kUnrollA];
#pragma unroll
      for (int u = 0; u < kUnrollA; ++u) {
        const int64_t c = base + (int64_t)u * kThreads;
        if (c < n_vec) v[u] = ld_stream(xv + c);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[u].w[i] = 0;
        }
      }
#pragma unroll
      for (int u = 0; u < kUnrollA; ++u) fold_abs_max<T>(v[u], m);
    }
    best = packed_to_f32_bits<T>(m);
    if (blockIdx.x == 0) {
      const int64_t i = n_vec * E + threadIdx.x;
      if (i < n) best = max(best, abs_bits_scalar(x + i));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
      best = max(best, abs_bits_scalar(x + i));
  }
  best = block_max_u32(best);
  if (threadIdx.x == 0 && best != 0) atomicMax(out, best);
}

// ------------------------------------------------------------------------------------------------
// rows kernel: G lanes cooperate on one row of `len` elements; G is a power of two <= 32.
// Vector path requires len % E == 0 and 32-byte aligned base.
template <typename T, int G>
__global__ void __launch_bounds__(kThreads) amax_rows_kernel(const T* __restrict__ x, int64_t rows, int64_t len,
                                                             int64_t axis_size, bool vectorized,
                                                             uint32_t* __restrict__ out) {
  constexpr int E = Elem<T>::kPerVec;
  const int lane_in_group = threadIdx.x % G;
  const int64_t groups_per_cta = kThreads / G;
  for (int64_t row = (int64_t)blockIdx.x * groups_per_cta + threadIdx.x / G; row < rows;
       row += (int64_t)gridDim.x * groups_per_cta) {
    const T* xr = x + row * len;
    uint32_t best;
    if (vectorized) {
      const V32* xv = reinterpret_cast<const V32*>(xr);
      const int64_t n_vec = len / E;
      uint32_t m = 0;
      int64_t c = lane_in_group;
      for (; c + (int64_t)G * (kUnrollA - 1) < n_vec; c += (int64_t)G * kUnrollA) {
        V32 v[kUnrollA];
#pragma unroll
        for (int u = 0; u < kUnrollA; ++u) v[u] = ld_stream(xv + c + (int64_t)u * G);
#pragma unroll
        for (int u = 0; u < kUnrollA; ++u) fold_abs_max<T>(v[u], m);
      }
      for (; c < n_vec; c += G) fold_abs_max<T>(ld_stream(xv + c), m);
      best = packed_to_f32_bits<T>(m);
    } else {
      best = 0;
      for (int64_t i = lane_in_group; i < len; i += G) best = max(best, abs_bits_scalar(xr + i));
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane_in_group == 0 && best != 0) atomicMax(out + row % axis_size, best);
  }
}

// Short rows of exactly G vectors (block-128 / block-256 amax, G >= 4): a group of G lanes takes FOUR rows per step -
// four independent 32-byte loads per lane - and the four partial maxima are reduced with a butterfly that halves the
// number of values a lane carries at every exchange (2 + 1 shuffles, then log2(G) - 2 plain ones) instead of
// 4 x log2(G).  It leaves every row's result in a distinct lane, and a warp's 4 * 32 / G rows are consecutive, so the
// merge into `out` is one coalesced atomicMax per warp step rather than one scattered atomic per row.
template <typename T, int G>
__global__ void __launch_bounds__(kThreads) amax_blocks_kernel(const T* __restrict__ x, int64_t rows, int64_t axis_size,
                                                               uint32_t* __restrict__ out) {
  static_assert(G >= 4 && G <= 32, "G");
  constexpr int GPW = 32 / G;            // groups per warp
  constexpr int RPW = 4 * GPW;           // rows per warp step
  const int lane = threadIdx.x & 31, lig = lane % G, g = lane / G;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * kThreads) >> 5;
  const V32* xv = reinterpret_cast<const V32*>(x);
  const bool hi_a = (lig & (G / 2)) != 0, hi_b = (lig & (G / 4)) != 0;
  for (int64_t base = warp * RPW; base < rows; base += n_warps * RPW) {
    V32 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t row = base + u * GPW + g;  // for a fixed u the warp reads GPW consecutive rows: 1 KB contiguous
      if (row < rows) v[u] = ld_stream(xv + row * G + lig);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[u].w[i] = 0;
      }
    }
    uint32_t b[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint32_t m = 0;
      fold_abs_max<T>(v[u], m);
      b[u] = packed_to_f32_bits<T>(m);
    }
    // exchange A: lanes with bit G/2 clear keep rows {0,1}, the others rows {2,3}
    uint32_t k0 = hi_a ? b[2] : b[0], k1 = hi_a ? b[3] : b[1];
    const uint32_t s0 = hi_a ? b[0] : b[2], s1 = hi_a ? b[1] : b[3];
    k0 = max(k0, __shfl_xor_sync(0xffffffffu, s0, G / 2));
    k1 = max(k1, __shfl_xor_sync(0xffffffffu, s1, G / 2));
    // exchange B: bit G/4 clear keeps the first of the two
    uint32_t r = hi_b ? k1 : k0;
    r = max(r, __shfl_xor_sync(0xffffffffu, hi_b ? k0 : k1, G / 4));
#pragma unroll
    for (int o = G / 8; o > 0; o >>= 1) r = max(r, __shfl_xor_sync(0xffffffffu, r, o));
    const int u = (hi_a ? 2 : 0) + (hi_b ? 1 : 0);
    const int64_t row = base + u * GPW + g;
    if ((lig & (G / 4 - 1)) == 0 && row < rows && r != 0) atomicMax(out + row % axis_size, r);
  }
}

// Long rows: a whole CTA per row (grid-strided over rows x splits).
template <typename T>
__global__ void __launch_bounds__(kThreads) amax_longrows_kernel(const T* __restrict__ x, int64_t rows, int64_t len,
                                                                 int64_t axis_size, int splits, bool vectorized,
                                                                 uint32_t* __restrict__ out) {
  constexpr int E = Elem<T>::kPerVec;
  const int64_t units = rows * splits;
  for (int64_t unit = blockIdx.x; unit < units; unit += gridDim.x) {
    const int64_t row = unit / splits;
    const int split = (int)(unit % splits);
    const T* xr = x + row * len;
    uint32_t best = 0;
    if (vectorized) {
      const V32* xv = reinterpret_cast<const V32*>(xr);
      const int64_t n_vec = len / E;
      const int64_t per = (n_vec + splits - 1) / splits;
      const int64_t lo = split * per, hi = min(n_vec, lo + per);
      uint32_t m = 0;
      int64_t c = lo + threadIdx.x;
      for (; c + (int64_t)kThreads * (kUnrollA - 1) < hi; c += (int64_t)kThreads * kUnrollA) {
        V32 v[kUnrollA];
#pragma unroll
        for (int u = 0; u < kUnrollA; ++u) v[u] = ld_stream(xv + c + (int64_t)u * kThreads);
#pragma unroll
        for (int u = 0; u < kUnrollA; ++u) fold_abs_max<T>(v[u], m);
      }
      for (; c < hi; c += kThreads) fold_abs_max<T>(ld_stream(xv + c), m);
      best = packed_to_f32_bits<T>(m);
    } else {
      const int64_t per = (len + splits - 1) / splits;
      const int64_t lo = split * per, hi = min(len, lo + per);
      for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) best = max(best, abs_bits_scalar(xr + i));
    }
    best = block_max_u32(best);
    if (threadIdx.x == 0 && best != 0) atomicMax(out + row % axis_size, best);
    __syncthreads();  // block_max_u32's shared scratch is reused by the next unit
  }
}

// ------------------------------------------------------------------------------------------------
// cols kernel: x is [rows, cols], out[c] = max_r |x[r, c]|.  grid = (col strips of kThreads vectors,
// row slabs).  Vector path: cols % E == 0, aligned.
template <typename T>
__global__ void __launch_bounds__(kThreads) amax_cols_kernel(const T* __restrict__ x, int64_t rows, int64_t cols,
                                                             int64_t rows_per_slab, bool vectorized,
                                                             uint32_t* __restrict__ out) {
  constexpr int E = Elem<T>::kPerVec;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = min(rows, r0 + rows_per_slab);
  if (vectorized) {
    const int64_t vec_cols = cols / E;
    const int64_t vc = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (vc >= vec_cols) return;
    uint32_t m[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) m[i] = 0;
    const V32* xv = reinterpret_cast<const V32*>(x);
    int64_t r = r0;
    for (; r + kUnrollA <= r1; r += kUnrollA) {
      V32 v[kUnrollA];
#pragma unroll
      for (int u = 0; u < kUnrollA; ++u) v[u] = ld_stream(xv + (r + u) * vec_cols + vc);
#pragma unroll
      for (int u = 0; u < kUnrollA; ++u)
#pragma unroll
        for (int i = 0; i < 8; ++i)
          m[i] = sizeof(T) == 2 ? __vmaxu2(m[i], v[u].w[i] & 0x7fff7fffu) : max(m[i], v[u].w[i] & 0x7fffffffu);
    }
    for (; r < r1; ++r) {
      const V32 v = ld_stream(xv + r * vec_cols + vc);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        m[i] = sizeof(T) == 2 ? __vmaxu2(m[i], v.w[i] & 0x7fff7fffu) : max(m[i], v.w[i] & 0x7fffffffu);
    }
    uint32_t* o = out + vc * E;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if constexpr (sizeof(T) == 4) {
        if (m[i]) atomicMax(o + i, m[i]);
      } else {
        const uint32_t lo = packed_to_f32_bits<T>(m[i] & 0xffffu), hi = packed_to_f32_bits<T>(m[i] >> 16);
        if (lo) atomicMax(o + 2 * i, lo);
        if (hi) atomicMax(o + 2 * i + 1, hi);
      }
    }
  } else {
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (c >= cols) return;
    uint32_t best = 0;
    for (int64_t r = r0; r < r1; ++r) best = max(best, abs_bits_scalar(x + r * cols + c));
    if (best) atomicMax(out + c, best);
  }
}

// cols kernel, vector path: the CTA is W column lanes x (kThreads / W) row lanes, every column lane owning K vectors
// (columns cl + k*W), so ONE CTA covers whole rows whenever cols <= 256*4 vectors and its reads are one contiguous
// slab of memory, exactly like the per-tensor kernel's (a strip-per-CTA layout, which reads 8 KB pieces at a row
// pitch, stayed at 67-71 % of the copy peak).  Row lanes are merged through shared memory so that only W threads per
// CTA touch `out`.
template <typename T, int W, int K>
__global__ void __launch_bounds__(kThreads, 3) amax_cols_tiled_kernel(const T* __restrict__ x, int64_t rows, int64_t cols,
                                                                   int64_t rows_per_slab, uint32_t* __restrict__ out) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int RL = kThreads / W;
  constexpr int U = K >= 4 ? 1 : 4 / K;   // rows in flight per lane: U*K independent 32-byte loads
  __shared__ uint32_t red[RL > 1 ? RL - 1 : 1][RL > 1 ? K * 8 : 1][RL > 1 ? W : 1];
  const int cl = threadIdx.x % W, rl = threadIdx.x / W;
  const int64_t vec_cols = cols / E;
  const int64_t vc0 = (int64_t)blockIdx.x * (W * K) + cl;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = min(rows, r0 + rows_per_slab);
  uint32_t m[K][8];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) m[k][i] = 0;
  const V32* xv = reinterpret_cast<const V32*>(x);
  auto fold = [&](const V32& v, int k) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      m[k][i] = sizeof(T) == 2 ? __vmaxu2(m[k][i], v.w[i] & 0x7fff7fffu) : max(m[k][i], v.w[i] & 0x7fffffffu);
  };
  for (int64_t r = r0 + rl; r < r1; r += (int64_t)RL * U) {
    V32 v[U][K];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const int64_t rr = r + (int64_t)u * RL, vc = vc0 + (int64_t)k * W;
        if (rr < r1 && vc < vec_cols) v[u][k] = ld_stream(xv + rr * vec_cols + vc);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[u][k].w[i] = 0;
        }
      }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < K; ++k) fold(v[u][k], k);
  }
  if constexpr (RL > 1) {
    if (rl > 0) {
#pragma unroll
      for (int k = 0; k < K; ++k)
#pragma unroll
        for (int i = 0; i < 8; ++i) red[rl - 1][k * 8 + i][cl] = m[k][i];
    }
    __syncthreads();
    if (rl == 0) {
#pragma unroll
    for (int q = 0; q < RL - 1; ++q)
#pragma unroll
      for (int k = 0; k < K; ++k)
#pragma unroll
        for (int i = 0; i < 8; ++i)
          m[k][i] = sizeof(T) == 2 ? __vmaxu2(m[k][i], red[q][k * 8 + i][cl]) : max(m[k][i], red[q][k * 8 + i][cl]);
    }
  }
  // Merge into `out` with consecutive lanes on consecutive columns (one 128-byte atomic request per warp instead of
  // 32 scattered ones): the W x 16 (or 8) per-thread results of one k go through shared memory, transposed.
  constexpr int EW = sizeof(T) == 2 ? 16 : 8;          // columns per vector
  __shared__ uint32_t tr[EW][W + 1];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    __syncthreads();
    if (rl == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if constexpr (sizeof(T) == 4) tr[i][cl] = m[k][i];
        else {
          tr[2 * i][cl] = packed_to_f32_bits<T>(m[k][i] & 0xffffu);
          tr[2 * i + 1][cl] = packed_to_f32_bits<T>(m[k][i] >> 16);
        }
      }
    }
    __syncthreads();
    const int64_t col0 = ((int64_t)blockIdx.x * (W * K) + (int64_t)k * W) * E;   // first column of this k
    for (int c = threadIdx.x; c < W * EW; c += kThreads) {
      const uint32_t v = tr[c % EW][c / EW];
      if (v != 0 && col0 + c < cols) atomicMax(out + col0 + c, v);
    }
  }
}

template <typename T>
static int amax_dispatch(const void* xv, int64_t outer, int64_t axis_size, int64_t inner, uint32_t* out,
                         cudaStream_t s) {
  constexpr int E = Elem<T>::kPerVec;
  const T* x = static_cast<const T*>(xv);
  const int64_t n = outer * axis_size * inner;
  const bool al = aligned32(xv);
  if (axis_size == 1) {
    const bool vec = al && n >= E;
    const int64_t tiles = vec ? (n / E + (int64_t)kThreads * kUnrollA - 1) / ((int64_t)kThreads * kUnrollA)
                              : (n + kThreads - 1) / kThreads;
    amax_flat_kernel<T><<<resident_grid((const void*)amax_flat_kernel<T>, tiles), kThreads, 0, s>>>(x, n, vec, out);
    return check_launch("mo_amax(flat)");
  }
  if (inner == 1) {
    // keep the last axis: [outer, axis_size]
    const int64_t rows = outer, cols = axis_size;
    const bool vec = al && cols % E == 0;
    if (vec) {
      const int64_t vcols = cols / E;
      // narrow rows: W = the smallest power of two >= vcols (32..256), K = 1; wider: W = 256, K = 1, 2 or 4 (1024 vectors
      // per strip; longer rows take several strips of 32 KB)
      int W = 256, K = 1;
      if (vcols <= 128) { W = 32; while (W < vcols) W <<= 1; }
      else { while (K < 4 && (int64_t)256 * K < vcols) K <<= 1; }
      const int RL = kThreads / W, U = K >= 4 ? 1 : 4 / K;
      const int64_t strips = (vcols + (int64_t)W * K - 1) / ((int64_t)W * K);
      int64_t slabs = (3 * kNumSMs + strips - 1) / strips;           // one resident wave; the per-CTA merge is not free
      const int64_t min_rows = (int64_t)RL * U * 8;
      if (slabs > (rows + min_rows - 1) / min_rows) slabs = (rows + min_rows - 1) / min_rows;
      if (slabs < 1) slabs = 1;
      if (slabs > 65535) slabs = 65535;
      const int64_t rps = (rows + slabs - 1) / slabs;
      dim3 grid((unsigned)strips, (unsigned)((rows + rps - 1) / rps));
#define MO_COLS(WW, KK) amax_cols_tiled_kernel<T, WW, KK><<<grid, kThreads, 0, s>>>(x, rows, cols, rps, out)
      if (W == 32) MO_COLS(32, 1);
      else if (W == 64) MO_COLS(64, 1);
      else if (W == 128) MO_COLS(128, 1);
      else if (K == 1) MO_COLS(256, 1);
      else if (K == 2) MO_COLS(256, 2);
      else MO_COLS(256, 4);
#undef MO_COLS
      return check_launch("mo_amax(cols)");
    }
    const int64_t strips = ((vec ? cols / E : cols) + kThreads - 1) / kThreads;
    // enough row slabs to fill the machine ~4x, at least 32 rows per slab
    int64_t slabs = (4 * kNumSMs + strips - 1) / strips;
    if (slabs > (rows + 31) / 32) slabs = (rows + 31) / 32;
    if (slabs < 1) slabs = 1;
    if (slabs > 65535) slabs = 65535;
    const int64_t rps = (rows + slabs - 1) / slabs;
    dim3 grid((unsigned)strips, (unsigned)((rows + rps - 1) / rps));
    amax_cols_kernel<T><<<grid, kThreads, 0, s>>>(x, rows, cols, rps, vec, out);
    return check_launch("mo_amax(cols)");
  }
  // rows of `inner` contiguous elements, row r belongs to channel r % axis_size
  const int64_t rows = outer * axis_size, len = inner;
  const bool vec = al && len % E == 0;
  const int64_t work = vec ? len / E : len;  // per-row units one lane handles per step
  if (work >= 32 * 64) {
    // long rows: CTA per row, split rows further when there are too few to fill the machine
    int splits = 1;
    while (rows * splits < 2 * kNumSMs && work / (splits * 2) >= kThreads * 2) splits *= 2;
    const int grid = persistent_grid(rows * splits, 8);
    amax_longrows_kernel<T><<<grid, kThreads, 0, s>>>(x, rows, len, axis_size, splits, vec, out);
    return check_launch("mo_amax(longrows)");
  }
  if (vec && work >= 4 && work <= 32 && (work & (work - 1)) == 0) {
    const int64_t tiles = (rows * work + kThreads - 1) / kThreads / 4 + 1;
#define MO_BLK(GG) \
  case GG: amax_blocks_kernel<T, GG><<<resident_grid((const void*)amax_blocks_kernel<T, GG>, tiles), kThreads, 0, s>>>( \
      x, rows, axis_size, out); break;
    switch ((int)work) { MO_BLK(4) MO_BLK(8) MO_BLK(16) MO_BLK(32) }
#undef MO_BLK
    return check_launch("mo_amax(blocks)");
  }
  int G = 1;
  while (G < 32 && G * 2 <= work) G *= 2;  // one vector per lane at least
  const int64_t groups_per_cta = kThreads / G;
  const int grid = persistent_grid((rows + groups_per_cta - 1) / groups_per_cta, 8);
#define MO_ROWS(GG) \
  case GG: amax_rows_kernel<T, GG><<<grid, kThreads, 0, s>>>(x, rows, len, axis_size, vec, out); break;
  switch (G) {
    MO_ROWS(1) MO_ROWS(2) MO_ROWS(4) MO_ROWS(8) MO_ROWS(16) MO_ROWS(32)
  }
#undef MO_ROWS
  return check_launch("mo_amax(rows)");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_amax(const void* x, int64_t outer, int64_t axis_size, int64_t inner, int dtype, uint32_t* amax_bits,
                       void* stream) {
  if (outer < 0 || axis_size < 1 || inner < 0) { set_error("mo_amax: bad shape"); return MO_ERR_SHAPE; }
  if (outer == 0 || inner == 0) return MO_OK;
  if (!x || !amax_bits) { set_error("mo_amax: NULL pointer"); return MO_ERR_NULL; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (axis_size == 1) { inner = outer * inner; outer = 1; }
  switch (dtype) {
    case MO_F32: return amax_dispatch<float>(x, outer, axis_size, inner, amax_bits, s);
    case MO_F16: return amax_dispatch<__half>(x, outer, axis_size, inner, amax_bits, s);
    case MO_BF16: return amax_dispatch<__nv_bfloat16>(x, outer, axis_size, inner, amax_bits, s);
  }
  set_error("mo_amax: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}
