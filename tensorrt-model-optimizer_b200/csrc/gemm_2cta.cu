// Quant-aware linear on CTA PAIRS (tcgen05 cta_group::2), bf16 / f16 operands (kind::f16; mo_linear_pair) and E4M3 operands
// with per-tensor scales (kind::f8f6f4; mo_linear_fp8_pair, the math of csrc/gemm_fp8.cu): the math of csrc/gemm.cu's plain and loss
// modes with the 256 x 256 tile of a pair of SMs.  Each CTA of the 2-CTA cluster holds its own 128 rows of A and HALF of the
// B tile, so per K step a CTA fills and reads 32 KB of shared memory instead of 48 KB for its 128 x 256 of output - the
// shared-memory bandwidth per SM is what bounds the single-CTA kernel - and the ring is 5 stages deep.
// Epilogue (plain mode): TMEM -> registers -> bf16 -> a 128-byte-swizzled staging tile in shared memory -> TMA store
// (cp.async.bulk.tensor ... global.shared::cta, SASS UTMASTG): each warp stages 32 rows x 64 columns at a time in one of
// two 4 KB buffers, so the global writes are whole 128-byte rows issued by the copy engine instead of 16-byte stores
// scattered over 32 rows per instruction, the epilogue warps never wait for a store, and ragged M / N are clipped by
// the tensor map.  The loss mode reads `ref` and writes nothing, as before.
// Protocol (per stage of the ring):
//   * warp 0 of BOTH CTAs is a TMA producer: it waits on its own `empty[stage]`, then loads its A rows and its B half with
//     cp.async.bulk.tensor...cta_group::2 whose complete_tx lands on the LEADER's `full[stage]` (peer bit of the barrier
//     address cleared); the leader's producer posts the expect_tx for the bytes of both CTAs;
//   * warp 1 lane 0 of the LEADER waits on `full[stage]` and issues tcgen05.mma.cta_group::2 (M = 256: TMEM lanes 0..127
//     of each CTA), then tcgen05.commit...multicast::cluster arrives on `empty[stage]` of both CTAs; after an accumulator's
//     last K step it commits to `tmem_full[acc]` of both CTAs;
//   * warps 2..5 of both CTAs drain their own 128 rows of the accumulator and arrive on the LEADER's `tmem_empty[acc]`
//     (8 arrivals: the peer's through mapa + mbarrier.arrive.shared::cluster).
#include <type_traits>

#include "tc_utils.cuh"

namespace mo {

namespace f16p {
// kMH = 128-row M blocks per CTA.  1: pair tile 256 x 256 (each CTA: 128 rows of A + 128 of the 256 B rows per stage, the
// two TMEM accumulators alternate between tiles).  2: pair tile 512 x 256 - each CTA holds 256 rows of A and issues two
// M = 256 MMAs per K step against the same B tile, one per accumulator: per flop 0.75x the bytes through L2 and shared
// memory (112 KB per 1024 tensor clocks per SM instead of 64 KB per 512), which is what bounds the 256 x 256 form (ncu
// side by side with cuBLAS's nvjet 256x256 2cta kernel, profiles/r02_gemm_vs_cublas.txt: 124 M vs 76 M L2 read sectors,
// tensor pipe 82.5 % vs 88.2 % active).  Both accumulators belong to the same tile, so its epilogue is only
// overlapped with the first K steps of the next tile's other half.
#ifndef MO_LAB_MH
#define MO_LAB_MH 0  /* 0 = chosen per call by the launcher; 1 / 2 force one form (tools/gemm_lab.py) */
#endif
constexpr int BM = 128, BN = 256;              // MMA M per CTA, pair-tile N
constexpr int kRowBytes = 128;                // K step: one 128-byte swizzle row = 64 16-bit elements or 128 E4M3 codes
constexpr int BNH = BN / 2;
constexpr int kMmaBytes = 32;                 // K bytes per tcgen05.mma (16 x 16-bit, 32 x 8-bit)
#ifndef MO_LAB_GROUPM
#define MO_LAB_GROUPM 0  /* 0 = chosen per shape by the launcher */
#endif
#ifndef MO_LAB_L2PROMO
#define MO_LAB_L2PROMO CU_TENSOR_MAP_L2_PROMOTION_L2_256B
#endif
#ifndef MO_LAB_PROLOGUE
#define MO_LAB_PROLOGUE 4
#endif
constexpr int kPrologue = MO_LAB_PROLOGUE;    // 512 x 256 form: K steps per tile issued accumulator-major (see the MMA issuer)
constexpr int kEpiCols = 64;                  // columns per staged chunk: 64 bf16 = one 128-byte swizzle row
constexpr uint32_t kEpiBytes = 32 * kEpiCols * 2;  // one warp's chunk: 32 rows x 128 bytes
constexpr int kThreadsG = 192;
constexpr uint32_t kBBytes = BNH * kRowBytes;
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;   // clears the CTA-rank bit of a shared::cluster address (leader's copy)

template <int kMH>
struct Cfg {
  static constexpr int BMC = BM * kMH;                 // A rows per CTA and tile
#ifdef MO_LAB_STAGES
  static constexpr int kStages = MO_LAB_STAGES;
#else
  static constexpr int kStages = kMH == 2 ? 4 : 5;     // 48 KB / 32 KB stages: 192 KB / 160 KB of the ring
#endif
  static constexpr uint32_t kABytes = BMC * kRowBytes, kStageBytes = kABytes + kBBytes;
};

template <int kMH>
struct Smem {
  uint8_t a[Cfg<kMH>::kStages][Cfg<kMH>::kABytes];
  uint8_t b[Cfg<kMH>::kStages][kBBytes];
  uint8_t epi[4][2][kEpiBytes];  // per epilogue warp: two staging chunks (1024-byte aligned: the swizzle atom)
  uint64_t full[Cfg<kMH>::kStages], empty[Cfg<kMH>::kStages], tmem_full[2], tmem_empty[2];
  uint64_t epi_bar[4][2];        // loss mode: a `ref` chunk has landed in epi[warp][buffer]
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_rank16() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all16() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair16(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
template <bool kFp8>
__device__ __forceinline__ void umma_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  if constexpr (kFp8)
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  else
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair16(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta16(uint64_t* bar, uint32_t cta) {
  asm volatile("{\n .reg .b32 ra;\n mapa.shared::cluster.u32 ra, %0, %1;\n mbarrier.arrive.shared::cluster.b64 _, [ra];\n}"
               ::"r"(smem_u32(bar)), "r"(cta)
               : "memory");
}
// Tiles are walked in groups of `group_m` M-blocks x all N-blocks (M fastest), so that the pairs running at the same
// time share A rows and B rows in L2.
__device__ __forceinline__ void tile_coord(int tile, int tiles_m, int tiles_n, int group_m, int& m_blk, int& n_blk) {
  const int per_group = group_m * tiles_n;
  const int g = tile / per_group, r = tile - g * per_group;
  const int gm = min(group_m, tiles_m - g * group_m);
  n_blk = r / gm;
  m_blk = g * group_m + (r - n_blk * gm);
}
}  // namespace f16p

// kFp8: E4M3 operands (K in codes, 128 per row) and out = T(acc * (*scale_a * *scale_b) + bias); else bf16 / f16 operands.
template <bool kHalf, bool kLoss, int kMH, bool kFp8>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(f16p::kThreadsG, 1)
linear_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_out,
                   const void* __restrict__ bias_v, void* __restrict__ out_v, const void* __restrict__ ref_v,
                   double* __restrict__ loss_acc, double loss_scale, const float* __restrict__ scale_a,
                   const float* __restrict__ scale_b, int M, int N, int K, int group_m) {
  using namespace f16p;
  using T = typename std::conditional<kHalf, __half, __nv_bfloat16>::type;
  constexpr int BMC = Cfg<kMH>::BMC, kStages = Cfg<kMH>::kStages;
  constexpr int BK = kFp8 ? kRowBytes : kRowBytes / 2;  // K elements per stage
  constexpr uint32_t kStageBytes = Cfg<kMH>::kStageBytes;
  extern __shared__ uint8_t smem_raw[];
  Smem<kMH>& s = *reinterpret_cast<Smem<kMH>*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_rank16();
  const bool leader = rank == 0;
  const int tiles_m = (M + 2 * BMC - 1) / (2 * BMC), tiles_n = (N + BN - 1) / BN, num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + BK - 1) / BK;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_out) : "memory");
    for (int i = 0; i < kStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s.tmem_full[i], 1); mbar_init(&s.tmem_empty[i], 8); }
    for (int i = 0; i < 8; ++i) mbar_init(&s.epi_bar[i >> 1][i & 1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the same warp of both CTAs allocates the pair's TMEM (same address in both)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all16();  // both CTAs' barriers are initialised before any remote arrive / complete_tx
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer (both CTAs) =====
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int m_blk, n_blk;
        tile_coord(tile, tiles_m, tiles_n, group_m, m_blk, n_blk);
        const int m0 = m_blk * 2 * BMC + (int)rank * BMC, n0 = n_blk * BN + (int)rank * BNH;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.empty[stage], phase ^ 1);
#if defined(MO_GEMM_EXP) && (MO_GEMM_EXP & 1)  // lab: no operand traffic at all (the MMAs read stale shared memory)
          if (leader) mbar_expect_tx(&s.full[stage], 0);
#else
          if (leader) mbar_expect_tx(&s.full[stage], 2 * kStageBytes);  // bytes of both CTAs land on this barrier
          tma_load_2d_pair16(s.a[stage], &map_a, &s.full[stage], kb * BK, m0);
          tma_load_2d_pair16(s.b[stage], &map_b, &s.full[stage], kb * BK, n0);
#endif
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {  // ===== MMA issuer (leader CTA only) =====
      // kind::f16 descriptor: D = F32, A / B = BF16 (1) or F16 (0) K-major, N >> 3 at [17,23), M >> 4 at [24,29), M = 256
      // (kind::f8f6f4: A / B format 0 = E4M3)
      const uint32_t fmt = (kFp8 || kHalf) ? 0u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0, unit = 0;  // unit = accumulator use counter: slot = unit % 2, its barrier phase = (unit / 2) % 2
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int kb0 = 0;
        if constexpr (kMH == 2 && kPrologue > 0) {
          // Both accumulators belong to this tile and the previous tile's epilogue drains them one after the other, so the
          // second one is free ~2000 clocks after the first.  The tile's first kPrologue K steps are therefore issued
          // accumulator-major: all of them for half 0, then for half 1 (their stages are released on the second pass) -
          // the second accumulator's drain hides behind the first half's MMAs instead of stalling the pipe.
          const int pro = min(kPrologue, num_kb);
#pragma unroll 1
          for (int h = 0; h < 2; ++h) {
            const uint32_t u = unit + (uint32_t)h, slot = u & 1u;
            mbar_wait(&s.tmem_empty[slot], ((u >> 1) & 1u) ^ 1u);
            tc_fence_after();
            const uint32_t d_tmem = tmem + slot * (uint32_t)BN;
            int st = stage;
            uint32_t ph = phase;
            for (int kb = 0; kb < pro; ++kb) {
              mbar_wait(&s.full[st], ph);
              tc_fence_after();
              const uint64_t dah = umma_desc(smem_u32(s.a[st])) + (uint64_t)((h * BM * kRowBytes) >> 4);
              const uint64_t db = umma_desc(smem_u32(s.b[st]));
#pragma unroll
              for (int k = 0; k < kRowBytes / kMmaBytes; ++k)
                umma_pair<kFp8>(d_tmem, dah + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
              if (kb == num_kb - 1) umma_commit_pair16(&s.tmem_full[slot]);
              if (h == 1) umma_commit_pair16(&s.empty[st]);
              if (++st == kStages) { st = 0; ph ^= 1; }
            }
            if (h == 1) { stage = st; phase = ph; }
          }
          kb0 = pro;
        }
        for (int kb = kb0; kb < num_kb; ++kb) {
          mbar_wait(&s.full[stage], phase);
          tc_fence_after();
          const uint64_t da = umma_desc(smem_u32(s.a[stage])), db = umma_desc(smem_u32(s.b[stage]));
#pragma unroll
          for (int h = 0; h < kMH; ++h) {
            const uint32_t u = unit + (uint32_t)h, slot = u & 1u;
            if (kb == 0) {  // the epilogue has drained this accumulator's previous use
              mbar_wait(&s.tmem_empty[slot], ((u >> 1) & 1u) ^ 1u);
              tc_fence_after();
            }
            const uint32_t d_tmem = tmem + slot * (uint32_t)BN;
            const uint64_t dah = da + (uint64_t)((h * BM * kRowBytes) >> 4);  // the CTA's h-th 128-row block of A
#pragma unroll
            for (int k = 0; k < kRowBytes / kMmaBytes; ++k)
              umma_pair<kFp8>(d_tmem, dah + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
            if (kb == num_kb - 1) umma_commit_pair16(&s.tmem_full[slot]);
          }
          umma_commit_pair16(&s.empty[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        unit += kMH;
      }
    }
  } else {
    // ===== epilogue (both CTAs): warps 2..5 own TMEM lanes 32*(warp%4) .. +31 of their CTA's accumulator half =====
    const int quarter = warp & 3;
    const T* bias = static_cast<const T*>(bias_v);
    [[maybe_unused]] T* out = static_cast<T*>(out_v);
    [[maybe_unused]] const T* ref = static_cast<const T*>(ref_v);
    float alpha = 1.0f;
    if constexpr (kFp8) alpha = __fmul_rn(__ldg(scale_a), __ldg(scale_b));
    uint32_t unit = 0;  // accumulator use counter, as in the MMA issuer
    uint32_t epi_uses = 0;       // loss mode: staging-buffer use counter (buffer = use % 2, barrier phase = (use / 2) % 2)
    float loss = 0.f;
    double dloss = 0.0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int m_blk, n_blk;
      tile_coord(tile, tiles_m, tiles_n, group_m, m_blk, n_blk);
      const int n0 = n_blk * BN;
#pragma unroll 1
      for (int h = 0; h < kMH; ++h, ++unit) {  // the CTA's h-th 128-row block: one accumulator
      const uint32_t acc = unit & 1u;
      if constexpr (!kLoss) {
        mbar_wait(&s.tmem_full[acc], (unit >> 1) & 1u);
        tc_fence_after();
      }
      const int row = m_blk * 2 * BMC + (int)rank * BMC + h * BM + quarter * 32 + lane;
      const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + acc * (uint32_t)BN;
      if constexpr (kLoss) {
        // loss mode: `ref` comes through the same staging buffers as the plain mode's output goes out - one TMA load per
        // 32 x 64 chunk, two chunks in flight, the first two issued BEFORE the accumulator is waited for (ref does not
        // depend on the MMAs) - instead of 16-byte loads scattered over 32 rows per instruction, whose latency sat in the
        // open once both accumulators belonged to one tile.
        const int row0 = m_blk * 2 * BMC + (int)rank * BMC + h * BM + quarter * 32;
        auto fetch = [&](int c) {  // chunk c of this accumulator -> buffer (c & 1)
          if (lane == 0) {
            uint64_t* bar = &s.epi_bar[quarter][c & 1];
            mbar_expect_tx(bar, kEpiBytes);
            tma_load_2d(s.epi[quarter][c & 1], &map_out, bar, n0 + c * kEpiCols, row0);
          }
        };
        fetch(0);  // both buffers are free: the previous accumulator's chunks were consumed in lock step
        fetch(1);
        mbar_wait(&s.tmem_full[acc], (unit >> 1) & 1u);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BN / kEpiCols; ++c, ++epi_uses) {
          const uint8_t* buf = s.epi[quarter][c & 1];
          mbar_wait(&s.epi_bar[quarter][c & 1], (epi_uses >> 1) & 1u);
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t r[32];
            tmem_ld32(taddr + (uint32_t)(c * kEpiCols + 32 * hh), r);
            if (row < M) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                const int col = n0 + c * kEpiCols + 32 * hh + j;
                if (col < N) {
                  float f[8];
#pragma unroll
                  for (int i = 0; i < 8; ++i) f[i] = kFp8 ? __fmul_rn(__uint_as_float(r[j + i]), alpha) : __uint_as_float(r[j + i]);
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
                  const int chunk = 4 * hh + (j >> 3);  // SWIZZLE_128B: 16-byte chunk index ^ (row % 8)
                  const uint4 rv = *reinterpret_cast<const uint4*>(buf + lane * 128 + ((chunk ^ (lane & 7)) << 4));
                  const T* rp = reinterpret_cast<const T*>(&rv);
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    T dt;
                    Elem<T>::store(&dt, __fsub_rn(Elem<T>::load(op + i), Elem<T>::load(rp + i)));
                    const float d = Elem<T>::load(&dt);
                    loss = __fmaf_rn(d, d, loss);
                  }
                }
              }
            }
          }
          __syncwarp();  // every lane has read this buffer
          if (c + 2 < BN / kEpiCols) fetch(c + 2);
        }
      } else {
        // plain mode: stage 32 rows x 64 columns per warp in swizzled shared memory, then one TMA store per chunk
        const int row0 = m_blk * 2 * BMC + (int)rank * BMC + h * BM + quarter * 32;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += kEpiCols) {
          uint8_t* buf = s.epi[quarter][(c0 / kEpiCols) & 1];  // BN / kEpiCols is even: the parity continues across halves
          // the TMA store that last read this buffer (two chunks ago) must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
#if defined(MO_GEMM_EXP) && (MO_GEMM_EXP & 2)  // lab: no epilogue work (accumulators are dropped)
          const bool live = false;
#else
          const bool live = n0 + c0 < N && row0 < M;  // a chunk wholly outside the tensor is neither staged nor stored
#endif
          if (live) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t r[32];
              tmem_ld32(taddr + (uint32_t)(c0 + 32 * h), r);
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                const int col = n0 + c0 + 32 * h + j;
                float f[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = kFp8 ? __fmul_rn(__uint_as_float(r[j + i]), alpha) : __uint_as_float(r[j + i]);
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
#if defined(MO_GEMM_EXP) && (MO_GEMM_EXP & 4)  // lab: TMEM is read, nothing is staged or stored
                if (o.x == 0x12345678u && o.y == 0x9abcdef0u) *reinterpret_cast<uint4*>(buf) = o;
#else
                *reinterpret_cast<uint4*>(buf + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = o;
#endif
              }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the TMA
          }
          __syncwarp();
          if (lane == 0) {
#if defined(MO_GEMM_EXP) && (MO_GEMM_EXP & 4)
            if (false)
#else
            if (live)
#endif
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                           ::"l"(&map_out), "r"(smem_u32(buf)), "r"(n0 + c0), "r"(row0)
                           : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");  // one group per chunk keeps the wait count exact
          }
        }
      }
      if constexpr (kLoss) { dloss += (double)loss; loss = 0.f; }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cta16(&s.tmem_empty[acc], 0);  // the leader's barrier collects both CTAs' 4 warps
      }
    }
    if constexpr (kLoss) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dloss += __shfl_xor_sync(0xffffffffu, dloss, o);
      if (lane == 0 && dloss != 0.0) atomicAdd(loss_acc, dloss * loss_scale);
    } else {
      if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // every store has landed before exit
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all16();  // the peer may still be reading / arriving: nobody frees TMEM or exits early
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// operand [rows, K] row-major (K-major), boxes of `box_rows` rows x one 128-byte swizzle row; esize 2 (bf16 / f16) or 1 (E4M3)
static bool make_map_pair(CUtensorMap* m, const void* base, int64_t rows, int64_t K, int box_rows, int esize, bool half) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)K * (cuuint64_t)esize};
  const cuuint32_t box[2] = {(cuuint32_t)(f16p::kRowBytes / esize), (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = esize == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                            : (half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  const CUresult r = fn(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, MO_LAB_L2PROMO, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return false; }
  return true;
}

// output [M, N] row-major, stored in boxes of 32 rows x 64 columns (one warp's staged chunk), 128-byte swizzle
static bool make_map16_out(CUtensorMap* m, void* base, int64_t M, int64_t N, bool half) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M};
  const cuuint64_t strides[1] = {(cuuint64_t)N * 2};
  const cuuint32_t box[2] = {(cuuint32_t)f16p::kEpiCols, 32u};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (output) failed (CUresult %d)", (int)r); return false; }
  return true;
}

template <bool kHalf, bool kLoss, int kMH, bool kFp8>
static int linear_pair_launch(const void* x, const void* w, const void* bias, void* out, const void* ref,
                              double* loss_acc, const float* scale_a, const float* scale_b, int M, int N, int K,
                              cudaStream_t s, const char* what) {
  CUtensorMap ma, mb, mo;
  const int esize = kFp8 ? 1 : 2;
  if (!make_map_pair(&ma, x, M, K, f16p::Cfg<kMH>::BMC, esize, kHalf) || !make_map_pair(&mb, w, N, K, f16p::BNH, esize, kHalf))
    return MO_ERR_ARG;
  // the [M, N] tensor the epilogue streams through its staging buffers: `out` (stored) or, in loss mode, `ref` (loaded)
  if (!make_map16_out(&mo, kLoss ? const_cast<void*>(ref) : out, M, N, kHalf)) return MO_ERR_ARG;
  auto kern = linear_pair_kernel<kHalf, kLoss, kMH, kFp8>;
  const int smem = (int)sizeof(f16p::Smem<kMH>) + 1024;
  static bool configured = false;
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e)); return (int)e; }
    configured = true;
  }
  constexpr int BMC = f16p::Cfg<kMH>::BMC;
  const int tiles = ((M + 2 * BMC - 1) / (2 * BMC)) * ((N + f16p::BN - 1) / f16p::BN);
  const int pairs = tiles < kNumSMs / 2 ? tiles : kNumSMs / 2;
  // B200 measurements (profiles/r02_gemm_lab_mh2.txt): a weight that fits in L2 next to the streaming A rows (N <= 4096
  // here) wants small groups - B stays resident, A is read once; wider weights want more M-blocks per pass over B
  int group_m = MO_LAB_GROUPM;
  if (group_m <= 0) group_m = (N <= 4096 ? 2 : 8) * (kMH == 2 ? 1 : 2);
  kern<<<2 * pairs, f16p::kThreadsG, smem, s>>>(ma, mb, mo, bias, out, ref, loss_acc, 1.0 / ((double)M * (double)N), scale_a,
                                                 scale_b, M, N, K, group_m);
  return check_launch(what);
}

// Which form runs (profiles/r02_gemm_lab_mh2.txt, r02_gemm_lab_loss.txt; sustained TFLOP/s at M = 16384 on B200):
//   plain: the 512 x 256 pair tile everywhere it was measured (+5 % at K = 4096, +13 % at K = 14336);
//   loss : its epilogue (TMEM read + `ref` from the staging buffers + the squared error) is about twice the plain one and
//          both accumulators of the 512 x 256 form belong to one tile, so it is only hidden by a long K loop: 512 x 256
//          from K = 8192 (+8 % at K = 14336), 256 x 256 below (where 512 x 256 measured -4 %).
template <bool kHalf, bool kLoss, bool kFp8>
static int linear_pair_pick(const void* x, const void* w, const void* bias, void* out, const void* ref, double* loss_acc,
                            const float* scale_a, const float* scale_b, int M, int N, int K, cudaStream_t s,
                            const char* what) {
  int mh = MO_LAB_MH;
  if (mh != 1 && mh != 2) mh = (kLoss && K < 8192) || M <= 256 ? 1 : 2;
  return mh == 2 ? linear_pair_launch<kHalf, kLoss, 2, kFp8>(x, w, bias, out, ref, loss_acc, scale_a, scale_b, M, N, K, s, what)
                 : linear_pair_launch<kHalf, kLoss, 1, kFp8>(x, w, bias, out, ref, loss_acc, scale_a, scale_b, M, N, K, s, what);
}

}  // namespace mo

using namespace mo;

extern "C" int mo_linear_pair(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                              int dtype, const void* ref, double* loss_acc, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_pair: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (dtype != MO_BF16 && dtype != MO_F16) { set_error("mo_linear_pair: bf16 / f16 only (dtype %d)", dtype); return MO_ERR_DTYPE; }
  const bool loss = ref != nullptr || loss_acc != nullptr;
  if (!x || !w || (loss ? (!ref || !loss_acc) : !out)) { set_error("mo_linear_pair: NULL pointer"); return MO_ERR_NULL; }
  if (K < 8 || K % 8 != 0 || N % 8 != 0) { set_error("mo_linear_pair: K and N must be multiples of 8"); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_pair: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(x) || mis(w) || mis(bias) || mis(out) || mis(ref)) { set_error("mo_linear_pair: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  const bool half = dtype == MO_F16;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int m = (int)M, n = (int)N, k = (int)K;
  const char* what = "mo_linear_pair";
  if (half) return loss ? linear_pair_pick<true, true, false>(x, w, bias, out, ref, loss_acc, nullptr, nullptr, m, n, k, s, what)
                        : linear_pair_pick<true, false, false>(x, w, bias, out, ref, loss_acc, nullptr, nullptr, m, n, k, s, what);
  return loss ? linear_pair_pick<false, true, false>(x, w, bias, out, ref, loss_acc, nullptr, nullptr, m, n, k, s, what)
              : linear_pair_pick<false, false, false>(x, w, bias, out, ref, loss_acc, nullptr, nullptr, m, n, k, s, what);
}

// FP8 per-tensor real-quant linear (the reference's `torch._scaled_mm` backend, backends/fp8_per_tensor_gemm.py:36-50) on the
// same kernel: kind::f8f6f4, K = 32 codes per instruction, out = T(acc * (scale_a * scale_b) + bias).
extern "C" int mo_linear_fp8_pair(const uint8_t* a, const uint8_t* b, const float* scale_a, const float* scale_b,
                                  const void* bias, void* out, int64_t M, int64_t N, int64_t K, int out_dtype,
                                  void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_fp8_pair: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (out_dtype != MO_BF16 && out_dtype != MO_F16) { set_error("mo_linear_fp8_pair: bf16 / f16 output only"); return MO_ERR_DTYPE; }
  if (!a || !b || !scale_a || !scale_b || !out) { set_error("mo_linear_fp8_pair: NULL pointer"); return MO_ERR_NULL; }
  if (K < 16 || K % 16 != 0 || N % 8 != 0) { set_error("mo_linear_fp8_pair: K %% 16 and N %% 8 required"); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_fp8_pair: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(a) || mis(b) || mis(bias) || mis(out)) { set_error("mo_linear_fp8_pair: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const char* what = "mo_linear_fp8_pair";
  const int m = (int)M, n = (int)N, k = (int)K;
  return out_dtype == MO_F16
             ? linear_pair_pick<true, false, true>(a, b, bias, out, nullptr, nullptr, scale_a, scale_b, m, n, k, s, what)
             : linear_pair_pick<false, false, true>(a, b, bias, out, nullptr, nullptr, scale_a, scale_b, m, n, k, s, what);
}
