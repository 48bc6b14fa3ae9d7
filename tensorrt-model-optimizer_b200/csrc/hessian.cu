// Local-Hessian weight calibration: the per-cin-block Hessian H_n = sum_t x_t,n^T x_t,n of a
// linear's input activations, and the NVFP4 FP8-scale sweep that minimises dw^T H dw per block.
//
// Replaces `_LocalHessianAccumulator.accumulate` (model_calib.py:848-862: float copy, transpose, reshape, batched fp32
// matmul - four passes over the activation) and the Triton `_fp8_scale_sweep_hessian_kernel`
// (gemm/nvfp4_fp8_sweep.py:173-290: a [32, 16] x [16, 16] tl.dot per candidate).
//
// block_hessian_kernel: a CTA owns 256 input columns (16 blocks of 16) and a slice of the tokens; thread (n, i) keeps
// row i of H_n in 16 registers; a 32-token tile is staged in shared memory as fp32 and every thread does 16 FMAs per
// token from one conflict-free scalar read and four broadcast 16-byte reads.  The activation is read from HBM exactly
// once (2 B per element); partial Hessians of the token slices are summed in a fixed order (deterministic).
//
// hessian_sweep_kernel: a CTA owns one cin-block and 256 output rows; H_n (1 KB) and the 126 candidate scales live in shared
// memory, each thread keeps its 16 weights in registers and evaluates dw^T H dw for every candidate in fp32 FMAs
// (the reference: allow_tf32 = False), strict `<` keeps the first minimum.
#include <math.h>

#include "common.cuh"

namespace mo {

constexpr int kHB = 16;          // Hessian block (NVFP4 block size)
constexpr int kHCols = 256;      // input columns per CTA = kThreads
constexpr int kHTok = 32;        // tokens per shared-memory tile
constexpr int kNumCand = 126;    // positive finite E4M3 values (gemm/_fp8_scale_candidates.py:26-31)

template <typename T>
__global__ void __launch_bounds__(kThreads) block_hessian_kernel(const T* __restrict__ x, int64_t tokens, int64_t cin,
                                                                 int64_t tok_per_split, float* __restrict__ partial) {
  __shared__ __align__(16) float tile[kHTok][kHCols];
  const int64_t col0 = (int64_t)blockIdx.x * kHCols;
  const int64_t t0 = (int64_t)blockIdx.y * tok_per_split;
  const int64_t t1 = min(tokens, t0 + tok_per_split);
  const int tid = threadIdx.x;
  const int nb = tid >> 4;  // block within the CTA's 16
  const bool col_ok = col0 + tid < cin;
  float acc[kHB];
#pragma unroll
  for (int j = 0; j < kHB; ++j) acc[j] = 0.f;
  for (int64_t tb = t0; tb < t1; tb += kHTok) {
    // stage: thread `tid` loads column col0 + tid of 32 tokens?  No - coalesce along the row: 256 threads cover one token
    // row of 256 columns per step (512 contiguous bytes for bf16), 32 steps per tile
#pragma unroll 8
    for (int r = 0; r < kHTok; ++r) {
      const int64_t t = tb + r;
      tile[r][tid] = (t < t1 && col_ok) ? Elem<T>::load(x + t * cin + col0 + tid) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < kHTok; ++r) {
      const float a = tile[r][tid];
      const float4* row = reinterpret_cast<const float4*>(&tile[r][nb * kHB]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 v = row[q];
        acc[4 * q + 0] = __fmaf_rn(a, v.x, acc[4 * q + 0]);
        acc[4 * q + 1] = __fmaf_rn(a, v.y, acc[4 * q + 1]);
        acc[4 * q + 2] = __fmaf_rn(a, v.z, acc[4 * q + 2]);
        acc[4 * q + 3] = __fmaf_rn(a, v.w, acc[4 * q + 3]);
      }
    }
    __syncthreads();
  }
  if (col_ok) {
    // partial[split][cin_block][i][j]; thread (nb, i = tid & 15)
    float* out = partial + ((int64_t)blockIdx.y * (cin / kHB) + col0 / kHB + nb) * (kHB * kHB) + (tid & 15) * kHB;
#pragma unroll
    for (int j = 0; j < kHB; j += 4) *reinterpret_cast<float4*>(out + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
  }
}

// H[i] += sum_s partial[s][i], s ascending (fixed order)
__global__ void __launch_bounds__(kThreads) hessian_reduce_kernel(const float* __restrict__ partial, int splits, int64_t n,
                                                                  float* __restrict__ h) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s = __fadd_rn(s, partial[(int64_t)k * n + i]);
    h[i] = __fadd_rn(h[i], s);
  }
}

template <typename T, bool kIeee>
__global__ void __launch_bounds__(kThreads) hessian_sweep_kernel(const T* __restrict__ w, int64_t cout, int64_t n_cin,
                                                                 const float* __restrict__ hessian,
                                                                 const float* __restrict__ global_amax,
                                                                 float* __restrict__ best_amax) {
  __shared__ __align__(16) float s_h[kHB * kHB];
  __shared__ float s_scale[kNumCand], s_amax[kNumCand], s_rcp[kNumCand], s_pre[kNumCand];
  const int64_t n = blockIdx.x;  // cin block
  const float ga = __ldg(global_amax);
  const Fp4ScaleOf<kIeee> scale_of(global_amax, 1, 448.0f);
  for (int k = threadIdx.x; k < kNumCand; k += kThreads) {
    uint32_t h2;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)(k + 1)));
    const float v = __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
    const float c = kIeee ? __fdiv_rn(v, 448.0f) : __fmul_rn(v, 1.0f / 448.0f);  // fp8_scale_candidates()
    const float am = __fmul_rn(c, ga);                                           // candidate_amaxes (:263)
    s_amax[k] = am;
    const float sc = scale_of(am);                                               // compute_fp4_scales (:264-266)
    s_scale[k] = sc == 0.0f ? 1.0f : sc;                                         // scale_safe (:219)
    const DivFull df(s_scale[k]);
    s_rcp[k] = df.r;
    s_pre[k] = df.pre;
  }
  for (int i = threadIdx.x; i < kHB * kHB; i += kThreads) s_h[i] = __ldg(hessian + n * (kHB * kHB) + i);
  __syncthreads();
  const int64_t row = (int64_t)blockIdx.y * kThreads + threadIdx.x;
  if (row >= cout) return;
  const int64_t b = row * n_cin + n;
  float wv[kHB], wa[kHB], sg[kHB];
  {
    constexpr int E = Elem<T>::kPerVec;
    constexpr int NV = kHB / E;
    const V32* xv = reinterpret_cast<const V32*>(w);
#pragma unroll
    for (int k = 0; k < NV; ++k) Elem<T>::unpack(ld_stream(xv + b * NV + k), wv + k * E);
  }
#pragma unroll
  for (int i = 0; i < kHB; ++i) {
    wa[i] = fabsf(wv[i]);
    sg[i] = wv[i] >= 0.f ? 1.0f : -1.0f;  // tl.where(w >= 0, 1, -1) (:196)
  }
  float best_loss = __int_as_float(0x7f800000);
  int best_k = 0;
  for (int k = 0; k < kNumCand; ++k) {
    const float sc = s_scale[k], rc = s_rcp[k], pre = s_pre[k];
    float dw[kHB];
#pragma unroll
    for (int i = 0; i < kHB; i += 2) {
      float a, c;
      if constexpr (kIeee) { a = __fdiv_rn(wa[i], sc); c = __fdiv_rn(wa[i + 1], sc); }
      else if (pre == 1.0f) { a = __fmul_rn(rc, wa[i]); c = __fmul_rn(rc, wa[i + 1]); }
      else { a = __fmul_rn(rc, __fmul_rn(wa[i], pre)); c = __fmul_rn(rc, __fmul_rn(wa[i + 1], pre)); }
      e2m1_roundtrip2(a, c);
      dw[i] = __fmul_rn(sg[i], __fsub_rn(wa[i], __fmul_rn(a, sc)));          // w - quant(w), signed
      dw[i + 1] = __fmul_rn(sg[i + 1], __fsub_rn(wa[i + 1], __fmul_rn(c, sc)));
    }
    // loss = sum_j (sum_i dw_i H_ij) dw_j
    float hdw[kHB];
#pragma unroll
    for (int j = 0; j < kHB; ++j) hdw[j] = 0.f;
#pragma unroll
    for (int i = 0; i < kHB; ++i) {
      const float4* hr = reinterpret_cast<const float4*>(&s_h[i * kHB]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 h = hr[q];
        hdw[4 * q + 0] = __fmaf_rn(dw[i], h.x, hdw[4 * q + 0]);
        hdw[4 * q + 1] = __fmaf_rn(dw[i], h.y, hdw[4 * q + 1]);
        hdw[4 * q + 2] = __fmaf_rn(dw[i], h.z, hdw[4 * q + 2]);
        hdw[4 * q + 3] = __fmaf_rn(dw[i], h.w, hdw[4 * q + 3]);
      }
    }
    float loss = 0.f;
#pragma unroll
    for (int j = 0; j < kHB; ++j) loss = __fmaf_rn(hdw[j], dw[j], loss);
    if (loss < best_loss) { best_loss = loss; best_k = k; }  // strict <: the first minimum wins (:229-231)
  }
  best_amax[b] = s_amax[best_k];
}

}  // namespace mo

using namespace mo;

extern "C" int64_t mo_block_hessian_scratch_bytes(int64_t tokens, int64_t cin) {
  if (tokens <= 0 || cin <= 0) return 0;
  const int64_t col_ctas = (cin + kHCols - 1) / kHCols;
  int64_t splits = (2 * kNumSMs + col_ctas - 1) / col_ctas;
  const int64_t max_splits = (tokens + kHTok - 1) / kHTok;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  return splits * (cin / kHB) * kHB * kHB * (int64_t)sizeof(float);
}

extern "C" int mo_block_hessian(const void* x, int64_t tokens, int64_t cin, int dtype, int block, float* hessian,
                                void* scratch, void* stream) {
  if (tokens == 0 || cin == 0) return MO_OK;
  if (!x || !hessian || !scratch) { set_error("mo_block_hessian: NULL pointer"); return MO_ERR_NULL; }
  if (block != kHB || cin % kHB != 0 || tokens < 0) {
    set_error("mo_block_hessian: block must be 16 and cin a multiple of 16 (got block=%d, cin=%lld)", block, (long long)cin);
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t col_ctas = (cin + kHCols - 1) / kHCols;
  const int64_t splits = mo_block_hessian_scratch_bytes(tokens, cin) / ((cin / kHB) * kHB * kHB * (int64_t)sizeof(float));
  int64_t tps = (tokens + splits - 1) / splits;
  tps = (tps + kHTok - 1) / kHTok * kHTok;
  const dim3 grid((unsigned)col_ctas, (unsigned)splits);
  float* partial = static_cast<float*>(scratch);
  // a split whose token range is empty must still write zeros: clear the scratch first (it is small)
  cudaMemsetAsync(partial, 0, (size_t)splits * (cin / kHB) * kHB * kHB * sizeof(float), s);
  switch (dtype) {
    case MO_F32: block_hessian_kernel<float><<<grid, kThreads, 0, s>>>(static_cast<const float*>(x), tokens, cin, tps, partial); break;
    case MO_F16: block_hessian_kernel<__half><<<grid, kThreads, 0, s>>>(static_cast<const __half*>(x), tokens, cin, tps, partial); break;
    case MO_BF16: block_hessian_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(static_cast<const __nv_bfloat16*>(x), tokens, cin, tps, partial); break;
    default: set_error("mo_block_hessian: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
  const int64_t n = cin * kHB;
  hessian_reduce_kernel<<<persistent_grid((n + kThreads - 1) / kThreads, 4), kThreads, 0, s>>>(partial, (int)splits, n, hessian);
  return check_launch("mo_block_hessian");
}

extern "C" int mo_nvfp4_fp8_scale_sweep_hessian(const void* w, int64_t cout, int64_t n_cin_blocks, int block, int dtype,
                                                const float* hessian, const float* global_amax, float* best_amax,
                                                int numerics, void* stream) {
  if (cout == 0 || n_cin_blocks == 0) return MO_OK;
  if (!w || !hessian || !global_amax || !best_amax) { set_error("mo_nvfp4_fp8_scale_sweep_hessian: NULL pointer"); return MO_ERR_NULL; }
  if (block != kHB || cout < 0 || n_cin_blocks < 0 || !aligned32(w)) {
    set_error("mo_nvfp4_fp8_scale_sweep_hessian: block must be 16 and w 32-byte aligned");
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const dim3 grid((unsigned)n_cin_blocks, (unsigned)((cout + kThreads - 1) / kThreads));
  const bool ieee = numerics == MO_DIV_IEEE;
#define MO_HS(T)                                                                                                       \
  if (ieee) hessian_sweep_kernel<T, true><<<grid, kThreads, 0, s>>>(static_cast<const T*>(w), cout, n_cin_blocks, hessian, global_amax, best_amax); \
  else hessian_sweep_kernel<T, false><<<grid, kThreads, 0, s>>>(static_cast<const T*>(w), cout, n_cin_blocks, hessian, global_amax, best_amax)
  switch (dtype) {
    case MO_F32: MO_HS(float); break;
    case MO_F16: MO_HS(__half); break;
    case MO_BF16: MO_HS(__nv_bfloat16); break;
    default: set_error("mo_nvfp4_fp8_scale_sweep_hessian: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
#undef MO_HS
  return check_launch("mo_nvfp4_fp8_scale_sweep_hessian");
}
