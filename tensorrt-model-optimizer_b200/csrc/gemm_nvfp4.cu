// NVFP4 W4A4 linear on the block-scaled tensor-core path:
//   out[m,n] = T( alpha * sum_k (a[m,k] * sfa[m,k/16]) * (b[n,k] * sfb[n,k/16]) + bias[n] ),  alpha = gs_a * gs_b
// a, b: packed E2M1 codes (two per byte, even element in the low nibble) as NVFP4QTensor.quantize / mo_nvfp4_pack
// write them (reference qtensor/nvfp4_tensor.py:253-342); sfa, sfb: their E4M3 block scales re-ordered by
// mo_nvfp4_sf_swizzle into the 128x4 tile order tcgen05 reads; gs_*: the fp32 global scales.  This is the role of the
// reference's real-quant NVFP4 GEMM backend (backends/nvfp4_gemm.py:28-125, which needs TensorRT-LLM) and the
// "kind::mxf4nvf4" path of SURVEY 7 step 9: the products of the DEQUANTIZED values are formed exactly by
// tcgen05.mma.kind::mxf4nvf4.block_scale and accumulated in fp32 in TMEM.
//
// Kernel shape: as csrc/gemm.cu (TMA producer warp, single-thread MMA issuer, 4 epilogue warps, persistent grid), with
//   * CTA tile 128 x 256, K step 256 elements (= 128 bytes of packed codes, one swizzle row), MMA K = 64;
//   * per stage: A 16 KB + B 32 KB by cp.async.bulk.tensor, scale tiles 2 KB + 4 KB by cp.async.bulk;
//   * scale factors go shared -> TMEM with tcgen05.cp.32x128b.warpx4 (4 columns per 128 rows x 64 K), double-buffered;
//   * one 256-column accumulator (TMEM: 256 accumulator + 96 scale-factor columns).
// Requirements: K % 256 == 0, N % 8 == 0; scale arrays padded as mo_nvfp4_sf_swizzle pads them.
#include <stdlib.h>

#include <type_traits>

#include "tc_utils.cuh"

namespace mo {

namespace f4 {
constexpr int BM = 128, BN = 256, BK = 256;          // BK in elements
constexpr int kStages = 4;
constexpr int kThreadsGemm = 192;
constexpr uint32_t kABytes = BM * BK / 2, kBBytes = BN * BK / 2;      // 16 KB, 32 KB
constexpr uint32_t kSfaBytes = (BM / 128) * (BK / 64) * 512;           // 2 KB
constexpr uint32_t kSfbBytes = (BN / 128) * (BK / 64) * 512;           // 4 KB
constexpr uint32_t kStageBytes = kABytes + kBBytes + kSfaBytes + kSfbBytes;
constexpr int kGroupM = 16;
constexpr uint32_t kAccCols = 256, kSfCols = 48;   // scale columns per stage: SFA 4 tiles x 4 + SFB 8 tiles x 4

struct Smem {
  uint8_t a[kStages][kABytes];
  uint8_t b[kStages][kBBytes];
  uint8_t sfa[kStages][kSfaBytes];
  uint8_t sfb[kStages][kSfbBytes];
  uint64_t full[kStages], empty[kStages], sf_ready[kStages], tmem_full, tmem_empty;
  uint32_t tmem_base;
};
}  // namespace f4

__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 32 rows x 16 bytes, rows contiguous (no swizzle): 8-row groups are 128 bytes apart.
__device__ __forceinline__ uint64_t sf_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(128u >> 4) << 32;   // stride byte offset
  d |= (uint64_t)1 << 46;             // descriptor version (sm_100)
  return d;                           // layout type 0 = no swizzle, leading byte offset 0 (one 16-byte column)
}
__device__ __forceinline__ void utccp_32x128b(uint32_t tmem_dst, uint64_t desc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(desc) : "memory");
}
__device__ __forceinline__ void umma_nvfp4(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t sfa_tmem, uint32_t sfb_tmem, uint32_t accumulate) {
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
      " tcgen05.mma.cta_group::1.kind::mxf4nvf4.block_scale.block16 [%0], %1, %2, %3, [%5], [%6], p;\n}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa_tmem), "r"(sfb_tmem)
      : "memory");
}

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,"
      "%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}

template <bool kHalf>
__global__ void __launch_bounds__(f4::kThreadsGemm, 1)
linear_nvfp4_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const __grid_constant__ CUtensorMap map_sfa, const __grid_constant__ CUtensorMap map_sfb,
                    const float* __restrict__ gs_a, const float* __restrict__ gs_b, const void* __restrict__ bias_v,
                    void* __restrict__ out_v, int M, int N, int K, int dbg) {
  using namespace f4;
  using T = typename std::conditional<kHalf, __half, __nv_bfloat16>::type;
  extern __shared__ uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN, num_tiles = tiles_m * tiles_n;
  const int num_kb = K / BK;
  const int sf_tiles_k = K / 64;  // 512-byte scale tiles per 128-row block

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_sfa) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_sfb) : "memory");
    for (int i = 0; i < kStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); mbar_init(&s.sf_ready[i], 4); }
    mbar_init(&s.tmem_full, 1);
    mbar_init(&s.tmem_empty, 4);
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
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int per_group = kGroupM * tiles_n;
        const int g = tile / per_group, r = tile - g * per_group;
        const int gm = min(kGroupM, tiles_m - g * kGroupM);
        const int n_blk = r / gm, m_blk = g * kGroupM + (r - n_blk * gm);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.empty[stage], phase ^ 1);
          mbar_expect_tx(&s.full[stage], kStageBytes);
          tma_load_2d(s.a[stage], &map_a, &s.full[stage], kb * (BK / 2), m_blk * BM);
          tma_load_2d(s.b[stage], &map_b, &s.full[stage], kb * (BK / 2), n_blk * BN);
          // scale tiles: rows of the [tiles, 128 x u32] view, BK/64 = 4 consecutive tiles per 128-row block
          tma_load_2d(s.sfa[stage], &map_sfa, &s.full[stage], 0, m_blk * sf_tiles_k + kb * (BK / 64));
#pragma unroll
          for (int rb = 0; rb < BN / 128; ++rb)
            tma_load_2d(s.sfb[stage] + rb * (kSfbBytes / (BN / 128)), &map_sfb, &s.full[stage], 0,
                        (n_blk * (BN / 128) + rb) * sf_tiles_k + kb * (BK / 64));
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // block-scaled instruction descriptor: A/B = E2M1 (1), scale format UE4M3 (0), K-major, N>>3 @17, M>>4 @24
      const uint32_t idesc = (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0, acc_phase = 0, it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&s.tmem_empty, acc_phase ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          mbar_wait(&s.full[stage], phase);
          tc_fence_after();
          const uint32_t sf_base = tmem + kAccCols + (uint32_t)stage * kSfCols;
          if (dbg & 2) {  // debug: scale factors through tcgen05.cp from the MMA thread (12 copies per K step)
            const uint32_t sfa_s = smem_u32(s.sfa[stage]), sfb_s = smem_u32(s.sfb[stage]);
#pragma unroll
            for (int j = 0; j < BK / 64; ++j) {
              utccp_32x128b(sf_base + 4 * j, sf_desc(sfa_s + j * 512));
#pragma unroll
              for (int rb = 0; rb < BN / 128; ++rb)
                utccp_32x128b(sf_base + 16 + 8 * j + 4 * rb, sf_desc(sfb_s + rb * (BK / 64) * 512 + j * 512));
            }
          } else {
            mbar_wait(&s.sf_ready[stage], phase);  // the epilogue warps have written this stage's scale columns
            tc_fence_after();
          }
          const uint64_t da = umma_desc(smem_u32(s.a[stage])), db = umma_desc(smem_u32(s.b[stage]));
#pragma unroll
          for (int j = 0; j < BK / 64; ++j)  // 64 elements = 32 bytes per MMA: +2 in the (>>4) address field
            umma_nvfp4(tmem, da + (uint64_t)(2 * j), db + (uint64_t)(2 * j), idesc, sf_base + 4 * j,
                       sf_base + 16 + 8 * j, (kb | j) != 0);
          umma_commit(&s.empty[stage]);
          if (kb == num_kb - 1) umma_commit(&s.tmem_full);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        acc_phase ^= 1;
      }
    }
  } else {
    const int quarter = warp & 3;
    const T* bias = static_cast<const T*>(bias_v);
    T* out = static_cast<T*>(out_v);
    const float alpha = __fmul_rn(__ldg(gs_a), __ldg(gs_b));
    uint32_t acc_phase = 0;
    int sf_stage = 0;
    uint32_t sf_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      // Scale factors: shared memory -> registers -> TMEM (tcgen05.st), one stage ahead of the MMAs.  Every warp
      // writes the same 48 columns into its own 32-lane quarter (the replication tcgen05.cp.warpx4 would do):
      // lane l holds, per 128-row x 64-K tile, the 16 bytes [row l + 32c][k 0..3], c = 0..3 = four TMEM columns.
      // A stage's TMEM columns are free again when its smem slot was refilled (full[stage] completes after the MMAs
      // that read them have committed empty[stage]).
      if (!(dbg & 2)) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.full[sf_stage], sf_phase);
          const uint4* pa = reinterpret_cast<const uint4*>(s.sfa[sf_stage]) + lane;
          const uint4* pb = reinterpret_cast<const uint4*>(s.sfb[sf_stage]) + lane;
          uint32_t ra[16], rb_[32];
#pragma unroll
          for (int j = 0; j < BK / 64; ++j) {
            const uint4 v = pa[j * 32];
            ra[4 * j] = v.x; ra[4 * j + 1] = v.y; ra[4 * j + 2] = v.z; ra[4 * j + 3] = v.w;
#pragma unroll
            for (int rb = 0; rb < BN / 128; ++rb) {
              const uint4 u = pb[(rb * (BK / 64) + j) * 32];
              rb_[8 * j + 4 * rb] = u.x; rb_[8 * j + 4 * rb + 1] = u.y; rb_[8 * j + 4 * rb + 2] = u.z; rb_[8 * j + 4 * rb + 3] = u.w;
            }
          }
          const uint32_t sf_t = tmem + ((uint32_t)(quarter * 32) << 16) + kAccCols + (uint32_t)sf_stage * kSfCols;
          tmem_st16(sf_t, ra);
          tmem_st32(sf_t + 16, rb_);
          asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&s.sf_ready[sf_stage]);
          if (++sf_stage == kStages) { sf_stage = 0; sf_phase ^= 1; }
        }
      }
      const int per_group = kGroupM * tiles_n;
      const int g = tile / per_group, r = tile - g * per_group;
      const int gm = min(kGroupM, tiles_m - g * kGroupM);
      const int n_blk = r / gm, m_blk = g * kGroupM + (r - n_blk * gm);
      mbar_wait(&s.tmem_full, acc_phase);
      tc_fence_after();
      const int row = m_blk * BM + quarter * 32 + lane;
      const int n0 = n_blk * BN;
      const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t rr[32];
        tmem_ld32(taddr + (uint32_t)c0, rr);
        if (row < M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const int col = n0 + c0 + j;
            if (col < N) {
              float f[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = __fmul_rn(__uint_as_float(rr[j + i]), alpha);
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
      if (lane == 0) mbar_arrive(&s.tmem_empty);
      acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// E4M3 block scales [rows, kblocks] (row-major) -> 128 x 4 tiles: tile (rb, kq) at (rb * ceil(kblocks/4) + kq) * 512,
// element (r, k) of the tile at (r % 32) * 16 + (r / 32) * 4 + k.  Rows are padded to `row_pad` (128 or 256) and
// k-blocks to a multiple of 4 with zeros.
__global__ void __launch_bounds__(kThreads) sf_swizzle_kernel(const uint8_t* __restrict__ sf, int64_t rows,
                                                              int64_t kblocks, int64_t rows_padded,
                                                              uint8_t* __restrict__ out) {
  const int64_t kq_n = (kblocks + 3) / 4;
  const int64_t total = rows_padded * kq_n;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int64_t row = i / kq_n, kq = i - row * kq_n;
    uint32_t v = 0;
    if (row < rows) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t kb = kq * 4 + k;
        if (kb < kblocks) v |= (uint32_t)sf[row * kblocks + kb] << (8 * k);
      }
    }
    const int64_t rb = row >> 7, r = row & 127;
    *reinterpret_cast<uint32_t*>(out + (rb * kq_n + kq) * 512 + (r & 31) * 16 + (r >> 5) * 4) = v;
  }
}

static bool make_map_u8(CUtensorMap* m, const void* base, int64_t rows, int64_t row_bytes, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
  const cuuint32_t box[2] = {128u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return false; }
  return true;
}

// scale buffer as [n_tiles, 128] u32 (512-byte tiles), box = 4 tiles, no swizzle
static bool make_map_sf(CUtensorMap* m, const void* base, int64_t n_tiles) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {128u, (cuuint64_t)n_tiles};
  const cuuint64_t strides[1] = {512u};
  const cuuint32_t box[2] = {128u, (cuuint32_t)(f4::BK / 64)};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (scales) failed (CUresult %d)", (int)r); return false; }
  return true;
}


// ------------------------------------------------------------------------------------------------------------------------
// CTA-pair form (tcgen05 cta_group::2, 256 x 256 pair tile): each CTA of the 2-CTA cluster holds its own 128 rows of A and
// HALF of the 256 B rows, so per K step a CTA fills and reads 32 KB of codes instead of 48 KB for its 128 x 256 of output -
// shared-memory bandwidth (TMA fill + MMA operand reads share 128 B/clk per SM) is what bounds the single-CTA kernel above.
// Every CTA still needs the scale factors of ALL 256 B rows (its tensor core forms its 128 rows against the whole N) and of
// its own 128 A rows, in its OWN tensor memory.
// Warp roles per CTA (320 threads): 0 TMA producer, 1 MMA issuer (leader CTA only) + TMEM allocation, 2..5 epilogue,
// 6..9 scale-factor transport.  Protocol:
//   * producer (both CTAs), per ring stage: scale tiles by plain TMA onto the CTA's own `sf_full[stage]`; codes by
//     cp.async.bulk.tensor...cta_group::2 whose complete_tx lands on the LEADER's `full[stage]`;
//   * scale warps (both CTAs): shared memory -> registers -> tcgen05.st into one of three 48-column slots (one K step: four
//     MMAs' worth of SFA 4 x 4 + SFB 4 x 8 columns), each warp its own lane quarter.  Slot k % 3 is free again when the MMAs
//     of K step k - 3 have completed, which is what `empty[stage of k - 3]` already says (tcgen05.commit, multicast to both
//     CTAs): no commit of its own - each extra commit per K step measured ~150 clocks of tensor-pipe bubble.  Written slots
//     are announced on the LEADER's `sf_ready[slot]` (8 arrivals);
//   * MMA thread (leader): full[stage] + sf_ready[slot] -> four M = 256 MMAs -> commit empty[stage]; tmem_full[acc] after a
//     tile's last K step (all multicast);
//   * epilogue (both CTAs): TMEM -> registers -> * alpha + bias -> bf16 -> swizzled staging chunk -> TMA store, as the
//     bf16 pair kernel (csrc/gemm_2cta.cu).
// Tensor memory (512 columns): the scale slots take 144, which leaves no room for two disjoint 256-column accumulators.
// kOverlap: accumulator 0 = columns [0, 256), accumulator 1 = [96, 352), tiles alternate, and the epilogue drains the
// shared columns [96, 256) FIRST (accumulator 0 from its top column down, accumulator 1 from its bottom up) and releases
// them (`r_free`) after 5 of its 8 column groups: the next tile's MMAs start then instead of after the whole drain.
// !kOverlap: one accumulator, `tmem_empty` after the whole drain.
// Tried here and measured slower (tools/gemm_nvfp4_lab.py, profiles/r02_gemm_nvfp4_lab.txt): scale factors by
// tcgen05.cp.cta_group::2 from the MMA thread (each copy occupies the tensor pipe for ~40 clocks, 12 per K step, in order
// with the MMAs: 3.0 against 3.3 PFLOP/s); cluster-scope release / acquire on the scale barriers (~1000 clocks per wait:
// 1.8 PFLOP/s); half-K-step slots with their own commits.
namespace f4p {
constexpr int BM = 128, BN = 256, BK = 256, BNH = 128;
constexpr int kStages = 5;
constexpr int kThreads = 320;
constexpr uint32_t kABytes = BM * BK / 2, kBBytes = BNH * BK / 2;      // 16 KB each
constexpr uint32_t kSfaBytes = 2048, kSfbBytes = 4096;                 // 4 + 2 x 4 tiles of 512 bytes per K step
constexpr uint32_t kCodeBytes = kABytes + kBBytes, kSfBytes = kSfaBytes + kSfbBytes;
constexpr int kSfSlots = 3;
constexpr uint32_t kSfSlotCols = 48, kAcc1Col = 96, kSfCol = 352;
constexpr int kRelGroups = 5;   // 32-column groups of an accumulator inside the shared columns [96, 256)
constexpr int kEpiCols = 64;
constexpr uint32_t kEpiBytes = 32 * kEpiCols * 2;
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;

struct Smem {
  uint8_t a[kStages][kABytes];
  uint8_t b[kStages][kBBytes];
  uint8_t epi[4][2][kEpiBytes];
  uint8_t sfa[kStages][kSfaBytes];
  uint8_t sfb[kStages][kSfbBytes];
  uint64_t full[kStages], empty[kStages], sf_full[kStages];
  uint64_t sf_ready[kSfSlots];
  uint64_t tmem_full[2], tmem_empty, r_free;
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_nvfp4_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                uint32_t sfa_tmem, uint32_t sfb_tmem, uint32_t accumulate) {
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
      " tcgen05.mma.cta_group::2.kind::mxf4nvf4.block_scale.block16 [%0], %1, %2, %3, [%5], [%6], p;\n}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa_tmem), "r"(sfb_tmem)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}
// arrive on CTA `cta`'s copy of `bar`; the arriving warp's tcgen05.st / tcgen05.ld are ordered before the leader's MMAs by
// tcgen05.fence::before_thread_sync on this side and ::after_thread_sync on the waiter's, as in csrc/gemm_2cta.cu
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {
  asm volatile("{\n .reg .b32 ra;\n mapa.shared::cluster.u32 ra, %0, %1;\n mbarrier.arrive.shared::cluster.b64 _, [ra];\n}"
               ::"r"(smem_u32(bar)), "r"(cta)
               : "memory");
}
__device__ __forceinline__ void tile_coord(int tile, int tiles_m, int tiles_n, int group_m, int& m_blk, int& n_blk) {
  const int per_group = group_m * tiles_n;
  const int g = tile / per_group, r = tile - g * per_group;
  const int gm = min(group_m, tiles_m - g * group_m);
  n_blk = r / gm;
  m_blk = g * group_m + (r - n_blk * gm);
}
}  // namespace f4p

template <bool kHalf, bool kOverlap>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(f4p::kThreads, 1)
linear_nvfp4_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                         const __grid_constant__ CUtensorMap map_sfa, const __grid_constant__ CUtensorMap map_sfb,
                         const __grid_constant__ CUtensorMap map_out, const float* __restrict__ gs_a,
                         const float* __restrict__ gs_b, const void* __restrict__ bias_v, int M, int N, int K, int group_m) {
  using namespace f4p;
  using T = typename std::conditional<kHalf, __half, __nv_bfloat16>::type;
  extern __shared__ uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_rank();
  const bool leader = rank == 0;
  const int tiles_m = (M + 2 * BM - 1) / (2 * BM), tiles_n = (N + BN - 1) / BN, num_tiles = tiles_m * tiles_n;
  const int num_kb = K / BK;
  const int sf_tiles_k = K / 64;  // 512-byte scale tiles per 128-row block
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_sfa) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_sfb) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_out) : "memory");
    for (int i = 0; i < kStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); mbar_init(&s.sf_full[i], 1); }
    for (int i = 0; i < kSfSlots; ++i) mbar_init(&s.sf_ready[i], 8);
    mbar_init(&s.tmem_full[0], 1);
    mbar_init(&s.tmem_full[1], 1);
    mbar_init(&s.tmem_empty, 8);
    mbar_init(&s.r_free, 8);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer (both CTAs) =====
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int m_blk, n_blk;
        tile_coord(tile, tiles_m, tiles_n, group_m, m_blk, n_blk);
        const int m0 = m_blk * 2 * BM + (int)rank * BM, n0 = n_blk * BN + (int)rank * BNH;
        const int sfa_row = (m_blk * 2 + (int)rank) * sf_tiles_k, sfb_row = n_blk * 2 * sf_tiles_k;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.empty[stage], phase ^ 1);
          mbar_expect_tx(&s.sf_full[stage], kSfBytes);
          tma_load_2d(s.sfa[stage], &map_sfa, &s.sf_full[stage], 0, sfa_row + kb * (BK / 64));
          tma_load_2d(s.sfb[stage], &map_sfb, &s.sf_full[stage], 0, sfb_row + kb * (BK / 64));
          tma_load_2d(s.sfb[stage] + kSfbBytes / 2, &map_sfb, &s.sf_full[stage], 0, sfb_row + sf_tiles_k + kb * (BK / 64));
          if (leader) mbar_expect_tx(&s.full[stage], 2 * kCodeBytes);  // the codes of both CTAs land on this barrier
          tma_load_2d_pair(s.a[stage], &map_a, &s.full[stage], kb * (BK / 2), m0);
          tma_load_2d_pair(s.b[stage], &map_b, &s.full[stage], kb * (BK / 2), n0);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {  // ===== MMA issuer (leader CTA only) =====
      // block-scaled instruction descriptor: A / B = E2M1 (1), scale format UE4M3 (0), K-major, N >> 3 @17, M >> 4 @24 (M = 256)
      const uint32_t idesc = (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
      int stage = 0, slot = 0;
      uint32_t phase = 0, slot_phase = 0, unit = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs, ++unit) {
        const uint32_t acc = kOverlap ? (unit & 1u) : 0u;
        const uint32_t d_tmem = tmem + acc * kAcc1Col;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.full[stage], phase);
          mbar_wait(&s.sf_ready[slot], slot_phase);
          if (kb == 0) {
            if constexpr (kOverlap) {
              if (unit > 0) mbar_wait(&s.r_free, (unit - 1u) & 1u);  // the previous tile's shared columns are drained
            } else {
              mbar_wait(&s.tmem_empty, (unit & 1u) ^ 1u);
            }
          }
          tc_fence_after();
          const uint64_t da = umma_desc(smem_u32(s.a[stage])), db = umma_desc(smem_u32(s.b[stage]));
          const uint32_t sf_t = tmem + kSfCol + (uint32_t)slot * kSfSlotCols;
#pragma unroll
          for (int j = 0; j < BK / 64; ++j)  // 64 elements = 32 bytes per MMA: +2 in the (>>4) address field
            umma_nvfp4_pair(d_tmem, da + (uint64_t)(2 * j), db + (uint64_t)(2 * j), idesc, sf_t + 4 * j, sf_t + 16 + 8 * j,
                            (kb | j) != 0);
          umma_commit_pair(&s.empty[stage]);
          if (kb == num_kb - 1) umma_commit_pair(&s.tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
          if (++slot == kSfSlots) { slot = 0; slot_phase ^= 1; }
        }
      }
    }
  } else if (warp >= 6) {
    // ===== scale-factor transport (both CTAs): lane l holds, per 128-row x 64-K tile, the 16 bytes [row l + 32c][k 0..3],
    // c = 0..3 = four TMEM columns; every warp writes the same columns into its own 32-lane quarter =====
    const int quarter = warp & 3;
    int stage = 0, slot = 0, free_stage = 0, step = 0;
    uint32_t phase = 0, free_phase = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      for (int kb = 0; kb < num_kb; ++kb, ++step) {
        mbar_wait(&s.sf_full[stage], phase);
        const uint4* pa = reinterpret_cast<const uint4*>(s.sfa[stage]) + lane;
        const uint4* pb = reinterpret_cast<const uint4*>(s.sfb[stage]) + lane;
        uint32_t ra[16], rb_[32];
#pragma unroll
        for (int j = 0; j < BK / 64; ++j) {
          const uint4 v = pa[j * 32];
          ra[4 * j] = v.x; ra[4 * j + 1] = v.y; ra[4 * j + 2] = v.z; ra[4 * j + 3] = v.w;
#pragma unroll
          for (int rb = 0; rb < 2; ++rb) {
            const uint4 u = pb[(rb * (BK / 64) + j) * 32];
            rb_[8 * j + 4 * rb] = u.x; rb_[8 * j + 4 * rb + 1] = u.y; rb_[8 * j + 4 * rb + 2] = u.z; rb_[8 * j + 4 * rb + 3] = u.w;
          }
        }
        if (step >= kSfSlots) {  // the MMAs of K step `step - kSfSlots` (the slot's previous user) have completed
          mbar_wait(&s.empty[free_stage], free_phase);
          if (++free_stage == kStages) { free_stage = 0; free_phase ^= 1; }
        }
        tc_fence_after();
        const uint32_t sf_t = tmem + ((uint32_t)(quarter * 32) << 16) + kSfCol + (uint32_t)slot * kSfSlotCols;
        tmem_st16(sf_t, ra);
        tmem_st32(sf_t + 16, rb_);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&s.sf_ready[slot], 0);
        if (++slot == kSfSlots) slot = 0;
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue (both CTAs): warps 2..5 own TMEM lanes 32 * (warp % 4) .. +31 of their CTA's 128 rows =====
    const int quarter = warp & 3;
    const T* bias = static_cast<const T*>(bias_v);
    const float alpha = __fmul_rn(__ldg(gs_a), __ldg(gs_b));
    uint32_t unit = 0, chunks = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs, ++unit) {
      int m_blk, n_blk;
      tile_coord(tile, tiles_m, tiles_n, group_m, m_blk, n_blk);
      const int n0 = n_blk * BN;
      const uint32_t acc = kOverlap ? (unit & 1u) : 0u;
      mbar_wait(&s.tmem_full[acc], kOverlap ? ((unit >> 1) & 1u) : (unit & 1u));
      tc_fence_after();
      const int row0 = m_blk * 2 * BM + (int)rank * BM + quarter * 32;
      const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + acc * kAcc1Col;
      const bool down = kOverlap && acc == 0;  // accumulator 0 shares its TOP columns: drain from column 255 down
      uint8_t* buf = nullptr;
      bool live = false;
#pragma unroll 1
      for (int gi = 0; gi < BN / 32; ++gi) {
        const int g = down ? BN / 32 - 1 - gi : gi;       // 32-column group of the accumulator
        const int c0 = (g >> 1) * kEpiCols, h = g & 1;    // its staged chunk (64 columns) and half
        if ((gi & 1) == 0) {                              // a new chunk
          buf = s.epi[quarter][chunks & 1u];
          ++chunks;
          // the TMA store that last read this buffer (two chunks ago) must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          live = n0 + c0 < N && row0 < M;  // a chunk wholly outside the tensor is neither staged nor stored
        }
        if (live) {
          uint32_t r[32];
          tmem_ld32(taddr + (uint32_t)(c0 + 32 * h), r);
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const int col = n0 + c0 + 32 * h + j;
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = __fmul_rn(__uint_as_float(r[j + i]), alpha);
            if (bias != nullptr && col < N) {
              const uint4 bv = __ldg(reinterpret_cast<const uint4*>(bias + col));
              const T* bp = reinterpret_cast<const T*>(&bv);
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = __fadd_rn(f[i], Elem<T>::load(bp + i));
            }
            uint4 o;
            T* op = reinterpret_cast<T*>(&o);
#pragma unroll
            for (int i = 0; i < 8; ++i) Elem<T>::store(op + i, f[i]);
            const int chunk = 4 * h + (j >> 3);  // 16-byte chunk of the 128-byte row; SWIZZLE_128B: chunk ^= row % 8
            *reinterpret_cast<uint4*>(buf + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = o;
          }
        }
        if (kOverlap && gi == kRelGroups - 1) {  // this warp has read the columns the two accumulators share
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cta(&s.r_free, 0);
        }
        if (gi & 1) {                                     // the chunk is complete
          if (live) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the TMA
          __syncwarp();
          if (lane == 0) {
            if (live)
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                           ::"l"(&map_out), "r"(smem_u32(buf)), "r"(n0 + c0), "r"(row0)
                           : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");  // one group per chunk keeps the wait count exact
          }
        }
      }
      if constexpr (!kOverlap) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&s.tmem_empty, 0);  // the leader's barrier collects both CTAs' 4 warps
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // every store has landed before exit
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer may still be reading / arriving: nobody frees TMEM or exits early
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// output [M, N] row-major, stored in boxes of 32 rows x 64 columns (one warp's staged chunk), 128-byte swizzle
static bool make_map_out_f4(CUtensorMap* m, void* base, int64_t M, int64_t N, bool half) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M};
  const cuuint64_t strides[1] = {(cuuint64_t)N * 2};
  const cuuint32_t box[2] = {(cuuint32_t)f4p::kEpiCols, 32u};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (output) failed (CUresult %d)", (int)r); return false; }
  return true;
}

template <bool kHalf, bool kOverlap>
static int linear_nvfp4_pair_launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa,
                                    const CUtensorMap& msb, const CUtensorMap& mo, const float* a_gs, const float* b_gs,
                                    const void* bias, int M, int N, int K, int group_m, cudaStream_t s) {
  auto kern = linear_nvfp4_pair_kernel<kHalf, kOverlap>;
  const int smem = (int)sizeof(f4p::Smem) + 1024;
  static bool configured = false;
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mo_linear_nvfp4_pair: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    configured = true;
  }
  const int tiles = ((M + 2 * f4p::BM - 1) / (2 * f4p::BM)) * ((N + f4p::BN - 1) / f4p::BN);
  const int pairs = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  kern<<<2 * pairs, f4p::kThreads, smem, s>>>(ma, mb, msa, msb, mo, a_gs, b_gs, bias, M, N, K, group_m);
  return check_launch("mo_linear_nvfp4_pair");
}

}  // namespace mo

using namespace mo;

extern "C" int64_t mo_nvfp4_sf_swizzle_bytes(int64_t rows, int64_t kblocks, int row_pad) {
  if (rows < 0 || kblocks < 0 || (row_pad != 128 && row_pad != 256)) return -1;
  const int64_t rp = (rows + row_pad - 1) / row_pad * row_pad;
  return rp / 128 * ((kblocks + 3) / 4) * 512;
}

extern "C" int mo_nvfp4_sf_swizzle(const void* sf, int64_t rows, int64_t kblocks, int row_pad, void* out, void* stream) {
  if (rows < 0 || kblocks < 0 || (row_pad != 128 && row_pad != 256)) { set_error("mo_nvfp4_sf_swizzle: bad arguments"); return MO_ERR_ARG; }
  if (rows == 0 || kblocks == 0) return MO_OK;
  if (!sf || !out) { set_error("mo_nvfp4_sf_swizzle: NULL pointer"); return MO_ERR_NULL; }
  const int64_t rp = (rows + row_pad - 1) / row_pad * row_pad;
  const int64_t total = rp * ((kblocks + 3) / 4);
  sf_swizzle_kernel<<<persistent_grid((total + kThreads - 1) / kThreads, 8), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(sf), rows, kblocks, rp, static_cast<uint8_t*>(out));
  return check_launch("mo_nvfp4_sf_swizzle");
}

extern "C" int mo_linear_nvfp4(const void* a_codes, const void* a_sf, const float* a_gs, const void* b_codes,
                               const void* b_sf, const float* b_gs, const void* bias, void* out, int64_t M, int64_t N,
                               int64_t K, int out_dtype, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_nvfp4: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (out_dtype != MO_BF16 && out_dtype != MO_F16) { set_error("mo_linear_nvfp4: bf16 / f16 output only"); return MO_ERR_DTYPE; }
  if (!a_codes || !a_sf || !a_gs || !b_codes || !b_sf || !b_gs || !out) { set_error("mo_linear_nvfp4: NULL pointer"); return MO_ERR_NULL; }
  if (K < f4::BK || K % f4::BK != 0 || N % 8 != 0) { set_error("mo_linear_nvfp4: K must be a multiple of %d and N of 8", f4::BK); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_nvfp4: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(a_codes) || mis(b_codes) || mis(a_sf) || mis(b_sf) || mis(bias) || mis(out)) { set_error("mo_linear_nvfp4: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  CUtensorMap ma, mb;
  if (!make_map_u8(&ma, a_codes, M, K / 2, f4::BM) || !make_map_u8(&mb, b_codes, N, K / 2, f4::BN)) return MO_ERR_ARG;
  CUtensorMap msa, msb;
  if (!make_map_sf(&msa, a_sf, (M + 127) / 128 * (K / 64)) || !make_map_sf(&msb, b_sf, (N + 255) / 256 * 2 * (K / 64)))
    return MO_ERR_ARG;
  const int smem = (int)sizeof(f4::Smem) + 1024;
  const bool half = out_dtype == MO_F16;
  auto kern = half ? linear_nvfp4_kernel<true> : linear_nvfp4_kernel<false>;
  static bool configured[2] = {false, false};
  if (!configured[half]) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mo_linear_nvfp4: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    configured[half] = true;
  }
  const int tiles = (int)(((M + f4::BM - 1) / f4::BM) * ((N + f4::BN - 1) / f4::BN));
  const int grid = tiles < kNumSMs ? tiles : kNumSMs;
  kern<<<grid, f4::kThreadsGemm, smem, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, msa, msb, a_gs, b_gs, bias, out, (int)M, (int)N, (int)K, getenv("MO_DBG_NVFP4") ? atoi(getenv("MO_DBG_NVFP4")) : 0);
  return check_launch("mo_linear_nvfp4");
}

// The same product on CTA pairs (256 x 256 pair tile, see linear_nvfp4_pair_kernel): same operands, same results.
extern "C" int mo_linear_nvfp4_pair(const void* a_codes, const void* a_sf, const float* a_gs, const void* b_codes,
                                    const void* b_sf, const float* b_gs, const void* bias, void* out, int64_t M, int64_t N,
                                    int64_t K, int out_dtype, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_nvfp4_pair: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (out_dtype != MO_BF16 && out_dtype != MO_F16) { set_error("mo_linear_nvfp4_pair: bf16 / f16 output only"); return MO_ERR_DTYPE; }
  if (!a_codes || !a_sf || !a_gs || !b_codes || !b_sf || !b_gs || !out) { set_error("mo_linear_nvfp4_pair: NULL pointer"); return MO_ERR_NULL; }
  if (K < f4p::BK || K % f4p::BK != 0 || N % 8 != 0) { set_error("mo_linear_nvfp4_pair: K must be a multiple of %d and N of 8", f4p::BK); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_nvfp4_pair: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(a_codes) || mis(b_codes) || mis(a_sf) || mis(b_sf) || mis(bias) || mis(out)) { set_error("mo_linear_nvfp4_pair: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  CUtensorMap ma, mb, msa, msb, mo_;
  if (!make_map_u8(&ma, a_codes, M, K / 2, f4p::BM) || !make_map_u8(&mb, b_codes, N, K / 2, f4p::BNH)) return MO_ERR_ARG;
  if (!make_map_sf(&msa, a_sf, (M + 127) / 128 * (K / 64)) || !make_map_sf(&msb, b_sf, (N + 255) / 256 * 2 * (K / 64)))
    return MO_ERR_ARG;
  const bool half = out_dtype == MO_F16;
  if (!make_map_out_f4(&mo_, out, M, N, half)) return MO_ERR_ARG;
  // lab switches (tools/gemm_nvfp4_lab.py): MO_LAB_F4P = 1 single accumulator; MO_LAB_F4P_GROUPM = M-blocks per pass over B
  const int lab = getenv("MO_LAB_F4P") ? atoi(getenv("MO_LAB_F4P")) : 0;
  const int lab_gm = getenv("MO_LAB_F4P_GROUPM") ? atoi(getenv("MO_LAB_F4P_GROUPM")) : 0;
  const int group_m = lab_gm > 0 ? lab_gm : (N <= 4096 ? 4 : 16);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int m = (int)M, n = (int)N, k = (int)K;
  if (lab == 1)
    return half ? linear_nvfp4_pair_launch<true, false>(ma, mb, msa, msb, mo_, a_gs, b_gs, bias, m, n, k, group_m, s)
                : linear_nvfp4_pair_launch<false, false>(ma, mb, msa, msb, mo_, a_gs, b_gs, bias, m, n, k, group_m, s);
  return half ? linear_nvfp4_pair_launch<true, true>(ma, mb, msa, msb, mo_, a_gs, b_gs, bias, m, n, k, group_m, s)
              : linear_nvfp4_pair_launch<false, true>(ma, mb, msa, msb, mo_, a_gs, b_gs, bias, m, n, k, group_m, s);
}
