// Real-quant FP8 linear: out[M,N] = (a[M,K] . b[N,K]^T) * scale_a * scale_b (+ bias), E4M3 operands, fp32 accumulation in
// TMEM, bf16 / f16 output.
//
// Replaces the reference's FP8 per-tensor GEMM backend (backends/fp8_per_tensor_gemm.py:36-50: torch._scaled_mm on the
// E4M3 input and weight with per-tensor scale_a / scale_b, use_fast_accum) - the GEMM a RealQuantLinear with
// FP8_DEFAULT_CFG runs after mtq.compress.
//
// Same shape as the bf16 kernel (csrc/gemm.cu): persistent grid, warp 0 = TMA producer, warp 1 = tcgen05.mma issuer and
// TMEM owner, warps 2..5 = epilogue; CTA tile 128 x 256, K step 128 (one 128-byte swizzle row of E4M3), 4-stage ring,
// two 256-column fp32 accumulators so the epilogue of tile i overlaps the MMAs of tile i+1.  The MMA is
// tcgen05.mma.kind::f8f6f4 (K = 32 per instruction: twice the math per shared-memory byte of the bf16 kernel).
// K must be a multiple of 16 (TMA row stride) and N a multiple of 8; M is arbitrary.
#include "tc_utils.cuh"

namespace mo {

namespace f8 {
constexpr int BM = 128, BN = 256, BK = 128;  // BK in elements = bytes
constexpr int UMMA_K = 32;
constexpr int kStages = 4;
constexpr int kThreadsG = 192;
constexpr uint32_t kABytes = BM * BK, kBBytes = BN * BK, kStageBytes = kABytes + kBBytes;
constexpr int kGroupM = 16;

struct Smem {
  uint8_t a[kStages][kABytes];
  uint8_t b[kStages][kBBytes];
  uint64_t full[kStages], empty[kStages], tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void tile_coord(int tile, int tiles_m, int tiles_n, int& m_blk, int& n_blk) {
  const int per_group = kGroupM * tiles_n;
  const int g = tile / per_group, r = tile - g * per_group;
  const int gm = min(kGroupM, tiles_m - g * kGroupM);
  n_blk = r / gm;
  m_blk = g * kGroupM + (r - n_blk * gm);
}
}  // namespace f8

template <bool kHalf>
__global__ void __launch_bounds__(f8::kThreadsG, 1)
linear_fp8_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const float* __restrict__ scale_a, const float* __restrict__ scale_b, const void* __restrict__ bias_v,
                  void* __restrict__ out_v, int M, int N, int K) {
  using namespace f8;
  using T = typename std::conditional<kHalf, __half, __nv_bfloat16>::type;
  extern __shared__ uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN, num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < kStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s.tmem_full[i], 1); mbar_init(&s.tmem_empty[i], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coord(tile, tiles_m, tiles_n, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.empty[stage], phase ^ 1);
          mbar_expect_tx(&s.full[stage], kStageBytes);
          tma_load_2d(s.a[stage], &map_a, &s.full[stage], kb * BK, m_blk * BM);
          tma_load_2d(s.b[stage], &map_b, &s.full[stage], kb * BK, n_blk * BN);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      // instruction descriptor (kind::f8f6f4): D = F32 at [4,6), A format [7,10) and B format [10,13) = 0 (E4M3), both
      // K-major, N >> 3 at [17,23), M >> 4 at [24,29)
      const uint32_t idesc = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&s.tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.full[stage], phase);
          tc_fence_after();
          const uint64_t da = umma_desc(smem_u32(s.a[stage])), db = umma_desc(smem_u32(s.b[stage]));
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)  // +32 bytes per step inside the 128-byte swizzle row
            umma_f8(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          umma_commit(&s.empty[stage]);
          if (kb == num_kb - 1) umma_commit(&s.tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue: warps 2..5 own TMEM lanes 32*(warp%4) .. +31 =====
    const int quarter = warp & 3;
    const T* bias = static_cast<const T*>(bias_v);
    T* out = static_cast<T*>(out_v);
    const float alpha = __fmul_rn(__ldg(scale_a), __ldg(scale_b));
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coord(tile, tiles_m, tiles_n, m_blk, n_blk);
      mbar_wait(&s.tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = m_blk * BM + quarter * 32 + lane;
      const int n0 = n_blk * BN;
      const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + (uint32_t)c0, r);
        if (row < M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const int col = n0 + c0 + j;
            if (col < N) {
              float f[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = __fmul_rn(__uint_as_float(r[j + i]), alpha);
              if (bias != nullptr) {
                const uint4 bv = __ldg(reinterpret_cast<const uint4*>(bias + col));
                const T* bp = reinterpret_cast<const T*>(&bv);
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = __fadd_rn(f[i], Elem<T>::load(bp + i));
              }
              uint4 o;
              T* op = reinterpret_cast<T*>(&o);
#pragma unroll
              for (int i = 0; i < 8; ++i) Elem<T>::store(op + i, f[i]);
              *reinterpret_cast<uint4*>(out + (int64_t)row * N + col) = o;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// [rows, K] row-major E4M3 matrix, box = [box_rows, 128 bytes] with the 128-byte swizzle.
static bool make_map_u8(CUtensorMap* m, const void* base, int64_t rows, int64_t K, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)K};
  const cuuint32_t box[2] = {(cuuint32_t)f8::BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return false; }
  return true;
}

template <bool kHalf>
static int linear_fp8_launch(const CUtensorMap& ma, const CUtensorMap& mb, const float* sa, const float* sb, const void* bias,
                             void* out, int M, int N, int K, cudaStream_t s) {
  auto kern = linear_fp8_kernel<kHalf>;
  const int smem = (int)sizeof(f8::Smem) + 1024;
  static bool configured = false;
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mo_linear_fp8: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    configured = true;
  }
  const int tiles = ((M + f8::BM - 1) / f8::BM) * ((N + f8::BN - 1) / f8::BN);
  const int grid = tiles < kNumSMs ? tiles : kNumSMs;
  kern<<<grid, f8::kThreadsG, smem, s>>>(ma, mb, sa, sb, bias, out, M, N, K);
  return check_launch("mo_linear_fp8");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_linear_fp8(const uint8_t* a, const uint8_t* b, const float* scale_a, const float* scale_b, const void* bias,
                             void* out, int64_t M, int64_t N, int64_t K, int out_dtype, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_fp8: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (out_dtype != MO_BF16 && out_dtype != MO_F16) { set_error("mo_linear_fp8: bf16 / f16 output only (dtype %d)", out_dtype); return MO_ERR_DTYPE; }
  if (!a || !b || !scale_a || !scale_b || !out) { set_error("mo_linear_fp8: NULL pointer"); return MO_ERR_NULL; }
  if (K < 16 || K % 16 != 0 || N % 8 != 0) { set_error("mo_linear_fp8: K must be a multiple of 16 and N of 8 (K=%lld N=%lld)", (long long)K, (long long)N); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_fp8: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(a) || mis(b) || mis(bias) || mis(out)) { set_error("mo_linear_fp8: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  CUtensorMap ma, mb;
  if (!make_map_u8(&ma, a, M, K, f8::BM) || !make_map_u8(&mb, b, N, K, f8::BN)) return MO_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return out_dtype == MO_F16 ? linear_fp8_launch<true>(ma, mb, scale_a, scale_b, bias, out, (int)M, (int)N, (int)K, s)
                             : linear_fp8_launch<false>(ma, mb, scale_a, scale_b, bias, out, (int)M, (int)N, (int)K, s);
}
