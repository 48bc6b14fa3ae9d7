// Local-Hessian weight calibration: the per-cin-block Hessian H_n = sum_t x_t,n^T x_t,n of a
// linear's input activations, and the NVFP4 FP8-scale sweep that minimises dw^T H dw per block.
//
// Replaces `_LocalHessianAccumulator.accumulate` (model_calib.py:848-862: float copy, transpose, reshape, batched fp32
// matmul - four passes over the activation) and the Triton `_fp8_scale_sweep_hessian_kernel`
// (gemm/nvfp4_fp8_sweep.py:173-290: a [32, 16] x [16, 16] tl.dot per candidate).
//
// block_hessian_kernel: a CTA owns 1024 input columns (64 blocks of 16) and a slice of the tokens; four threads share a
// block, each keeping an 8x8 quarter of H_n in registers; an 8-token tile is staged in shared memory with 32-byte vector
// copies.  The activation is read from HBM exactly once (2 B per element); partial Hessians of the token slices are
// summed in a fixed order (deterministic).
//
// hessian_sweep_kernel: a CTA owns one cin-block and 256 output rows; H_n (1 KB) and the 126 candidate scales live in shared
// memory, each thread keeps its 16 weights in registers and evaluates dw^T H dw for every candidate in fp32 FMAs
// (the reference: allow_tf32 = False), strict `<` keeps the first minimum.
#include <math.h>

#include "common.cuh"

namespace mo {

constexpr int kHB = 16;          // Hessian block (NVFP4 block size)
constexpr int kHCols = 1024;     // input columns per CTA: 64 blocks of 16, four threads per block
constexpr int kHTok = 8;         // tokens per shared-memory tile (8 x 1024 elements)
constexpr int kNumCand = 126;    // positive finite E4M3 values (gemm/_fp8_scale_candidates.py:26-31)

// Thread (n, qi, qj) of a CTA keeps the 8x8 quarter [8qi, 8qi+8) x [8qj, 8qj+8) of H_n in 64 registers: per token it reads
// 8 + 8 values (two 16-byte shared-memory loads for bf16) and does 64 FMAs, so the FMA pipe, not the shared-memory port,
// is the limit (the first version - one row of H_n per thread, 5 loads per 16 FMAs - ran at 10 % of the FMA rate).
template <typename T>
__global__ void __launch_bounds__(kThreads) block_hessian_kernel(const T* __restrict__ x, int64_t tokens, int64_t cin,
                                                                 int64_t tok_per_split, float* __restrict__ partial) {
  constexpr int E = Elem<T>::kPerVec;                 // elements per 32-byte vector
  constexpr int VPR = kHCols / E;                     // vectors per tile row
  constexpr int LOADS = kHTok * VPR / kThreads;       // vectors per thread per tile
  __shared__ __align__(32) T tile[kHTok][kHCols];     // raw elements: converted when read
  const int64_t col0 = (int64_t)blockIdx.x * kHCols;
  const int64_t t0 = (int64_t)blockIdx.y * tok_per_split;
  const int64_t t1 = min(tokens, t0 + tok_per_split);
  const int tid = threadIdx.x;
  const int nb = tid >> 2, qi = (tid >> 1) & 1, qj = tid & 1;
  const bool blk_ok = col0 + (int64_t)nb * kHB < cin;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  // the next tile's vectors are fetched into registers while this tile is multiplied: with two CTAs per SM the load
  // latency (about as long as a tile's FMAs) was exposed once per tile
  V32 pre[LOADS];
  auto fetch = [&](int64_t tb) {
#pragma unroll
    for (int l = 0; l < LOADS; ++l) {
      const int v = tid + l * kThreads, r = v / VPR, cv = v % VPR;
      const int64_t t = tb + r, c = col0 + (int64_t)cv * E;
      if (t < t1 && c < cin) pre[l] = ld_stream(reinterpret_cast<const V32*>(x + t * cin + c));
      else {
#pragma unroll
        for (int k = 0; k < 8; ++k) pre[l].w[k] = 0u;
      }
    }
  };
  if (t0 < t1) fetch(t0);
  for (int64_t tb = t0; tb < t1; tb += kHTok) {
#pragma unroll
    for (int l = 0; l < LOADS; ++l) {
      const int v = tid + l * kThreads, r = v / VPR, cv = v % VPR;
      *reinterpret_cast<V32*>(&tile[r][cv * E]) = pre[l];
    }
    __syncthreads();
    if (tb + kHTok < t1) fetch(tb + kHTok);
#pragma unroll 2
    for (int r = 0; r < kHTok; ++r) {
      float a[8], b[8];
      if constexpr (E == 16) {  // 8 elements = 16 bytes
        const uint4 ua = *reinterpret_cast<const uint4*>(&tile[r][nb * kHB + 8 * qi]);
        const uint4 ub = *reinterpret_cast<const uint4*>(&tile[r][nb * kHB + 8 * qj]);
        const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            a[2 * k] = __uint_as_float(wa[k] << 16); a[2 * k + 1] = __uint_as_float(wa[k] & 0xffff0000u);
            b[2 * k] = __uint_as_float(wb[k] << 16); b[2 * k + 1] = __uint_as_float(wb[k] & 0xffff0000u);
          } else {
            const float2 fa = __half22float2(*reinterpret_cast<const __half2*>(&wa[k]));
            const float2 fb = __half22float2(*reinterpret_cast<const __half2*>(&wb[k]));
            a[2 * k] = fa.x; a[2 * k + 1] = fa.y; b[2 * k] = fb.x; b[2 * k + 1] = fb.y;
          }
        }
      } else {
        const float4* pa = reinterpret_cast<const float4*>(&tile[r][nb * kHB + 8 * qi]);
        const float4* pb = reinterpret_cast<const float4*>(&tile[r][nb * kHB + 8 * qj]);
        const float4 a0 = pa[0], a1 = pa[1], b0 = pb[0], b1 = pb[1];
        a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
      }
      // 16-bit inputs: two accumulators per packed-fp32 instruction (each half is an fma.rn: same bits, half the issue
      // slots; 0.086 -> 0.077 ms at [16384, 4096] bf16).  fp32 inputs are nearer the HBM floor and measured slower
      // packed (0.089 -> 0.096 ms), so they keep the scalar FMAs.
      if constexpr (E != 16) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = __fmaf_rn(a[i], b[j], acc[i][j]);
      } else
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; j += 2)
          asm("{ .reg .b64 A, B, C;\n"
              "  mov.b64 A, {%2, %2}; mov.b64 B, {%3, %4}; mov.b64 C, {%0, %1};\n"
              "  fma.rn.f32x2 C, A, B, C; mov.b64 {%0, %1}, C; }"
              : "+f"(acc[i][j]), "+f"(acc[i][j + 1]) : "f"(a[i]), "f"(b[j]), "f"(b[j + 1]));
    }
    __syncthreads();
  }
  if (blk_ok) {
    // partial[split][cin_block][i][j]
    float* out = partial + ((int64_t)blockIdx.y * (cin / kHB) + col0 / kHB + nb) * (kHB * kHB);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float* row = out + (8 * qi + i) * kHB + 8 * qj;
      *reinterpret_cast<float4*>(row) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      *reinterpret_cast<float4*>(row + 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
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
        // packed fp32: two columns per instruction, each half an fma.rn (same bits, half the issue slots)
        asm("{ .reg .b64 A, B, C, D, E;\n"
            "  mov.b64 A, {%4, %4}; mov.b64 B, {%5, %6}; mov.b64 C, {%0, %1};\n"
            "  mov.b64 D, {%7, %8}; mov.b64 E, {%2, %3};\n"
            "  fma.rn.f32x2 C, A, B, C; fma.rn.f32x2 E, A, D, E;\n"
            "  mov.b64 {%0, %1}, C; mov.b64 {%2, %3}, E; }"
            : "+f"(hdw[4 * q + 0]), "+f"(hdw[4 * q + 1]), "+f"(hdw[4 * q + 2]), "+f"(hdw[4 * q + 3])
            : "f"(dw[i]), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w));
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
  if (block != kHB || cin % kHB != 0 || tokens < 0 || !aligned32(x)) {
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

// ---------------------------------------------------------------------------------------------------------------------
// GPTQ column loop of one block, fused with the NVFP4 quantizer (replaces Triton _gptq_scalar_kernel,
// gemm/gptq_fused_kernel.py:43-136; the reference runs one Triton program per row with masked [BLOCK] vector ops per
// column).  One warp owns a row: lane l keeps columns l, l + 32, ... in registers; per column the owner's value is
// broadcast by shuffle, every lane derives q and err = (w - q) / H^-1[c][c], and the lanes holding later columns apply
// w_j -= err * H^-1[c][j].  Rows of the block's inverse-Hessian factor are read through L1 (all warps of a CTA walk
// the same row at the same time).
//   scale = e4m3(min(amax / (6 * gs), 448)) * gs          (common/nvfp4_quant.py:104-125)
//   q     = sign(w) * e2m1(|w| / scale_safe) * scale_safe  (:66-101), zero scale -> 0
// Numerics: GPU = Triton's `/` (div.full.f32) and fused multiply-subtract; CPU = IEEE division, separate multiply and
// subtract (the torch restatement, oracle/hessian.py).
// ---------------------------------------------------------------------------------------------------------------------
namespace mo {

template <int K, bool kIeee>
__global__ void __launch_bounds__(kThreads) gptq_block_nvfp4_kernel(const float* __restrict__ w, float* __restrict__ qw,
                                                                    float* __restrict__ err, const float* __restrict__ amax,
                                                                    float gs, const float* __restrict__ hinv, int64_t rows,
                                                                    int n_amax_blocks, int quant_block, int block_start) {
  constexpr int B = 32 * K;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;  // whole warp
  float wr[K];
#pragma unroll
  for (int k = 0; k < K; ++k) wr[k] = w[row * B + k * 32 + lane];
  const float six_gs = __fmul_rn(6.0f, gs);
  const float* arow = amax + row * n_amax_blocks;
#pragma unroll
  for (int kk = 0; kk < K; ++kk) {
    for (int l = 0; l < 32; ++l) {
      const int col = kk * 32 + l;
      const float ws = __shfl_sync(0xffffffffu, wr[kk], l);
      const float bam = __ldg(arow + (block_start + col) / quant_block);
      // fp8_quantize_scale
      float t = kIeee ? __fdiv_rn(bam, six_gs) : div_full(bam, six_gs);
      t = fminf(t, 448.0f);
      const float scale = __fmul_rn(e4m3_roundtrip1(t), gs);
      // nvfp4_scalar_quant
      const bool zero = scale == 0.0f;
      const float safe = (zero || isnan(scale) || isinf(scale)) ? 1.0f : scale;
      float a = kIeee ? __fdiv_rn(fabsf(ws), safe) : div_full(fabsf(ws), safe), z = 0.f;
      e2m1_roundtrip2(a, z);
      const float resc = __fmul_rn(a, safe);
      const float q = zero ? 0.0f : (ws >= 0.0f ? resc : -resc);
      const float d = __ldg(hinv + col * B + col);
      const float diff = __fsub_rn(ws, q);
      const float e = kIeee ? __fdiv_rn(diff, d) : div_full(diff, d);
      if (lane == l) {
        qw[row * B + col] = q;
        err[row * B + col] = e;
      }
#pragma unroll
      for (int k2 = 0; k2 < K; ++k2) {
        const int j = k2 * 32 + lane;
        if (j > col) {
          const float h = __ldg(hinv + col * B + j);
          wr[k2] = kIeee ? __fsub_rn(wr[k2], __fmul_rn(e, h)) : __fmaf_rn(-e, h, wr[k2]);
        }
      }
    }
  }
}

}  // namespace mo

extern "C" int mo_gptq_block_nvfp4(const float* w, float* qw, float* err, const float* amax, float global_scale,
                                   const float* hinv, int64_t rows, int block, int n_amax_blocks, int quant_block,
                                   int block_start, int numerics, void* stream) {
  if (rows == 0) return MO_OK;
  if (!w || !qw || !err || !amax || !hinv) { set_error("mo_gptq_block_nvfp4: NULL pointer"); return MO_ERR_NULL; }
  if (rows < 0 || quant_block < 1 || n_amax_blocks < 1 || block_start < 0 ||
      (block != 32 && block != 64 && block != 128 && block != 256)) {
    set_error("mo_gptq_block_nvfp4: block must be 32, 64, 128 or 256 (got %d)", block);
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = (int)((rows + kThreads / 32 - 1) / (kThreads / 32));
  const bool ieee = numerics == MO_DIV_IEEE;
#define MO_GQ(KK)                                                                                                         \
  if (ieee) mo::gptq_block_nvfp4_kernel<KK, true><<<grid, kThreads, 0, s>>>(w, qw, err, amax, global_scale, hinv, rows, n_amax_blocks, quant_block, block_start); \
  else mo::gptq_block_nvfp4_kernel<KK, false><<<grid, kThreads, 0, s>>>(w, qw, err, amax, global_scale, hinv, rows, n_amax_blocks, quant_block, block_start)
  switch (block) {
    case 32: MO_GQ(1); break;
    case 64: MO_GQ(2); break;
    case 128: MO_GQ(4); break;
    default: MO_GQ(8); break;
  }
#undef MO_GQ
  return check_launch("mo_gptq_block_nvfp4");
}
