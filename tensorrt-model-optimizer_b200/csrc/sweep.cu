// Scale searches that call fake-quant repeatedly, fused so the tensor is read ONCE for every candidate:
//
//  mo_nvfp4_fp8_scale_sweep   per 16-element block try the 126 positive finite E4M3 values c/448 as the block
//                             scale (scale = c * global_amax / 6) and keep the one with the least squared error.
//                             Replaces Triton _fp8_scale_sweep_kernel (reference gemm/nvfp4_fp8_sweep.py:57-162;
//                             its comment quotes 5.08 ms on B300 for the autotuned config) and the 126-pass
//                             Python sweep of NVFP4MSECalibrator (calib/mse.py:175-311).
//  mo_mse_sweep               MseCalibrator (calib/mse.py:31-172): for K amax multipliers accumulate
//                             sum((x - fq(x, amax*c_k))^2) per tensor / per channel; INT-N or FP8 E4M3.
//                             The reference runs K full fake-quant + mse_loss + reduce passes (K = 39).
//
// Both are ALU-bound (126 x ~8 and 39 x ~10 thread-instructions per element), not HBM-bound.
#include "common.cuh"

namespace mo {

constexpr int kNumFp8Cand = 126;

// ------------------------------------------------------------------------------------------------
// NVFP4 FP8-scale sweep.  One thread = one block of BLOCK elements kept in registers as |w|.
// ------------------------------------------------------------------------------------------------
template <typename T, int BLOCK, bool kIeee>
__global__ void __launch_bounds__(kThreads) nvfp4_fp8_sweep_kernel(const T* __restrict__ x, int64_t n_blocks,
                                                                   const float* __restrict__ global_amax,
                                                                   float* __restrict__ best_amax) {
  constexpr int E = Elem<T>::kPerVec;
  constexpr int NV = BLOCK / E;
  __shared__ float s_scale[kNumFp8Cand];  // scale_safe per candidate
  __shared__ float s_cand[kNumFp8Cand];   // c_k = e4m3(k + 1) / 448
  __shared__ float s_rcp[kNumFp8Cand], s_pre[kNumFp8Cand];  // div.full.f32's divisor-only part (DivFull), per candidate
  const float ga = __ldg(global_amax);
  for (int k = threadIdx.x; k < kNumFp8Cand; k += kThreads) {
    uint32_t h2;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)(k + 1)));
    const float v = __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
    // fp8_scale_candidates(): `fp8_values / 448.0` is a torch op: CUDA multiplies by the fp32 reciprocal, CPU divides
    const float c = kIeee ? __fdiv_rn(v, 448.0f) : __fmul_rn(v, 1.0f / 448.0f);
    s_cand[k] = c;
    const float sc = div_mode<kIeee>(__fmul_rn(c, ga), 6.0f);  // scale = c * global_amax / 6.0
    s_scale[k] = sc == 0.0f ? 1.0f : sc;
    const DivFull df(s_scale[k]);
    s_rcp[k] = df.r;
    s_pre[k] = df.pre;
  }
  __syncthreads();
  const V32* xv = reinterpret_cast<const V32*>(x);
  for (int64_t b = (int64_t)blockIdx.x * kThreads + threadIdx.x; b < n_blocks; b += (int64_t)gridDim.x * kThreads) {
    float w[BLOCK];
#pragma unroll
    for (int k = 0; k < NV; ++k) Elem<T>::unpack(ld_stream(xv + b * NV + k), w + k * E);
#pragma unroll
    for (int i = 0; i < BLOCK; ++i) w[i] = fabsf(w[i]);
    float best_loss = __int_as_float(0x7f800000);
    int best_k = 0;
#pragma unroll 2
    for (int k = 0; k < kNumFp8Cand; ++k) {
      const float sc = s_scale[k];
      const float rc = s_rcp[k], pre = s_pre[k];
      float part[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < BLOCK; i += 2) {
        float a, c;
        if constexpr (kIeee) { a = __fdiv_rn(w[i], sc); c = __fdiv_rn(w[i + 1], sc); }
        else if (pre == 1.0f) { a = __fmul_rn(rc, w[i]); c = __fmul_rn(rc, w[i + 1]); }
        else { a = __fmul_rn(rc, __fmul_rn(w[i], pre)); c = __fmul_rn(rc, __fmul_rn(w[i + 1], pre)); }
        e2m1_roundtrip2(a, c);  // fp4_round_magnitude
        const float d0 = __fsub_rn(w[i], __fmul_rn(a, sc));
        const float d1 = __fsub_rn(w[i + 1], __fmul_rn(c, sc));
        part[(i / 2) & 3] = __fadd_rn(part[(i / 2) & 3], __fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)));
      }
      const float loss = __fadd_rn(__fadd_rn(part[0], part[1]), __fadd_rn(part[2], part[3]));
      if (loss < best_loss) { best_loss = loss; best_k = k; }  // strict <: the first minimum wins
    }
    best_amax[b] = __fmul_rn(ga, s_cand[best_k]);  // best_amax = global_amax * c[best]
  }
}

// generic (any block size / unaligned): one thread per block, scalar loads, block re-read per candidate from L1
template <typename T, bool kIeee>
__global__ void __launch_bounds__(kThreads) nvfp4_fp8_sweep_generic_kernel(const T* __restrict__ x, int64_t n_blocks,
                                                                           int block,
                                                                           const float* __restrict__ global_amax,
                                                                           float* __restrict__ best_amax) {
  const float ga = __ldg(global_amax);
  for (int64_t b = (int64_t)blockIdx.x * kThreads + threadIdx.x; b < n_blocks; b += (int64_t)gridDim.x * kThreads) {
    const T* xp = x + b * block;
    float best_loss = __int_as_float(0x7f800000), best_c = 0.f;
    bool first = true;
    for (int k = 0; k < kNumFp8Cand; ++k) {
      uint32_t h2;
      asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)(k + 1)));
      const float v = __half22float2(*reinterpret_cast<__half2*>(&h2)).x;
      const float c = kIeee ? __fdiv_rn(v, 448.0f) : __fmul_rn(v, 1.0f / 448.0f);
      float sc = div_mode<kIeee>(__fmul_rn(c, ga), 6.0f);
      if (sc == 0.0f) sc = 1.0f;
      float loss = 0.f;
      for (int i = 0; i < block; ++i) {
        const float w = fabsf(Elem<T>::load(xp + i));
        float a = div_mode<kIeee>(w, sc), z = 0.f;
        e2m1_roundtrip2(a, z);
        const float d = __fsub_rn(w, __fmul_rn(a, sc));
        loss = __fadd_rn(loss, __fmul_rn(d, d));
      }
      if (first || loss < best_loss) { best_loss = loss; best_c = c; first = false; }
    }
    best_amax[b] = __fmul_rn(ga, best_c);
  }
}

// ------------------------------------------------------------------------------------------------
// MSE multiplier sweep.  View [rows, len], channel = row % axis_size; one WARP per row.
// losses[k * axis_size + ch] += sum_row (x - fq(x, amax[ch] * mult[k]))^2      (fp64 accumulation)
// ------------------------------------------------------------------------------------------------

struct MseFmt {
  int fp8;          // 1: E4M3 (CUDA-extension scale arithmetic), 0: INT-N
  float maxb, minb; // INT bounds
};

__device__ __forceinline__ float mse_fq_int(float x, float scale, const ExactDiv& d, float minb, float maxb, bool zero) {
  float v = __fmul_rn(x, scale);
  v = fmaxf(fminf(v, maxb), minb);
  v = __fsub_rn(__fadd_rn(v, 12582912.0f), 12582912.0f);  // rint
  return zero ? 0.0f : d(v);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) mse_sweep_kernel(const T* __restrict__ x, int64_t rows, int64_t len,
                                                             int64_t axis_size, const float* __restrict__ cand,
                                                             int K, MseFmt fmt, int numerics_cpu,
                                                             double* __restrict__ losses) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * kThreads) >> 5;
  for (int64_t row = warp; row < rows; row += n_warps) {
    const int64_t ch = row % axis_size;
    const T* xr = x + row * len;
    for (int k = 0; k < K; ++k) {
      const float am = __ldg(cand + (int64_t)k * axis_size + ch);  // initial_amax * candidate, formed by the caller
      float acc = 0.f;
      double dacc = 0.0;
      if (fmt.fp8) {
        const float safe = am <= kEps ? 1.0f : am;
        const float sc = numerics_cpu ? __fmul_rn(__fdiv_rn(1.0f, safe), 448.0f) : __fdiv_rn(448.0f, safe);
        const float inv = __fdiv_rn(1.0f, sc);
        int cnt = 0;
        for (int64_t i = lane; i < len; i += 32) {
          const float v = Elem<T>::load(xr + i);
          const float q = __fmul_rn(e4m3_roundtrip1(__fmul_rn(v, sc)), inv);
          const float d = __fsub_rn(v, q);
          acc = __fmaf_rn(d, d, acc);
          if (++cnt == 64) { dacc += (double)acc; acc = 0.f; cnt = 0; }
        }
      } else {
        const bool zero = am < kEps;
        const float sc = __fdiv_rn(fmt.maxb, am);
        const ExactDiv dv(sc);
        int cnt = 0;
        for (int64_t i = lane; i < len; i += 32) {
          const float v = Elem<T>::load(xr + i);
          const float q = mse_fq_int(v, sc, dv, fmt.minb, fmt.maxb, zero);
          const float d = __fsub_rn(v, q);
          acc = __fmaf_rn(d, d, acc);
          if (++cnt == 64) { dacc += (double)acc; acc = 0.f; cnt = 0; }
        }
      }
      dacc += (double)acc;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dacc += __shfl_xor_sync(0xffffffffu, dacc, o);
      if (lane == 0) atomicAdd(&losses[(int64_t)k * axis_size + ch], dacc);
    }
  }
}

template <typename T, bool kIeee>
static int sweep_launch(const void* xv, int64_t n_blocks, int block, const float* ga, float* best, cudaStream_t s) {
  constexpr int E = Elem<T>::kPerVec;
  const T* x = static_cast<const T*>(xv);
  const int grid = persistent_grid((n_blocks + kThreads - 1) / kThreads, 4);
  if (aligned32(xv) && block == 16 && 16 % E == 0)
    nvfp4_fp8_sweep_kernel<T, 16, kIeee><<<grid, kThreads, 0, s>>>(x, n_blocks, ga, best);
  else
    nvfp4_fp8_sweep_generic_kernel<T, kIeee><<<grid, kThreads, 0, s>>>(x, n_blocks, block, ga, best);
  return check_launch("mo_nvfp4_fp8_scale_sweep");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_nvfp4_fp8_scale_sweep(const void* x, int64_t n_blocks, int block, int dtype,
                                        const float* global_amax, float* best_amax, int numerics, void* stream) {
  if (n_blocks == 0) return MO_OK;
  if (!x || !global_amax || !best_amax) { set_error("mo_nvfp4_fp8_scale_sweep: NULL pointer"); return MO_ERR_NULL; }
  if (n_blocks < 0 || block < 1) { set_error("mo_nvfp4_fp8_scale_sweep: bad block / n_blocks"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool ieee = numerics == MO_NUMERICS_CPU;
  switch (dtype) {
    case MO_F32: return ieee ? sweep_launch<float, true>(x, n_blocks, block, global_amax, best_amax, s)
                             : sweep_launch<float, false>(x, n_blocks, block, global_amax, best_amax, s);
    case MO_F16: return ieee ? sweep_launch<__half, true>(x, n_blocks, block, global_amax, best_amax, s)
                             : sweep_launch<__half, false>(x, n_blocks, block, global_amax, best_amax, s);
    case MO_BF16: return ieee ? sweep_launch<__nv_bfloat16, true>(x, n_blocks, block, global_amax, best_amax, s)
                              : sweep_launch<__nv_bfloat16, false>(x, n_blocks, block, global_amax, best_amax, s);
  }
  set_error("mo_nvfp4_fp8_scale_sweep: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}

extern "C" int mo_mse_sweep(const void* x, int64_t outer, int64_t axis_size, int64_t inner, int dtype,
                            const float* cand_amax, int n_cand, int num_bits, int is_unsigned,
                            int narrow_range, int fp8, int numerics, double* losses, void* stream) {
  if (outer < 0 || axis_size < 1 || inner < 0) { set_error("mo_mse_sweep: bad shape"); return MO_ERR_SHAPE; }
  if (outer == 0 || inner == 0) return MO_OK;
  if (!x || !cand_amax || !losses) { set_error("mo_mse_sweep: NULL pointer"); return MO_ERR_NULL; }
  if (n_cand < 1) { set_error("mo_mse_sweep: n_cand < 1"); return MO_ERR_ARG; }
  if (!fp8 && (num_bits < 1 || num_bits + (is_unsigned ? 1 : 0) > 22)) { set_error("mo_mse_sweep: bad num_bits"); return MO_ERR_ARG; }
  MseFmt fmt;
  fmt.fp8 = fp8;
  fmt.maxb = (float)((1 << (num_bits - 1 + (is_unsigned ? 1 : 0))) - 1);
  fmt.minb = -(fmt.maxb + (narrow_range ? 0.f : 1.f));
  int64_t rows = outer * axis_size, len = inner;
  if (axis_size == 1) {  // per-tensor: chop into rows of 4096 elements (+ one ragged row) for parallelism
    const int64_t n = outer * inner;
    len = n >= 4096 && n % 4096 == 0 ? 4096 : (n % 1024 == 0 ? 1024 : n);
    rows = n / len;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = persistent_grid((rows * 32 + kThreads - 1) / kThreads, 8);
  const int cpu = numerics == MO_NUMERICS_CPU;
  switch (dtype) {
    case MO_F32: mse_sweep_kernel<float><<<grid, kThreads, 0, s>>>(static_cast<const float*>(x), rows, len, axis_size, cand_amax, n_cand, fmt, cpu, losses); break;
    case MO_F16: mse_sweep_kernel<__half><<<grid, kThreads, 0, s>>>(static_cast<const __half*>(x), rows, len, axis_size, cand_amax, n_cand, fmt, cpu, losses); break;
    case MO_BF16: mse_sweep_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>(static_cast<const __nv_bfloat16*>(x), rows, len, axis_size, cand_amax, n_cand, fmt, cpu, losses); break;
    default: set_error("mo_mse_sweep: unsupported dtype %d", dtype); return MO_ERR_DTYPE;
  }
  return check_launch("mo_mse_sweep");
}
