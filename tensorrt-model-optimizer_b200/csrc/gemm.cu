// Quant-aware linear: out[M,N] = xq[M,K] . wq[N,K]^T (+ bias), bf16/f16 operands, fp32 accumulation in TMEM.
//
// Replaces the F.linear of QuantLinearConvBase.forward / _QuantLinear.quantized_linear_fn (reference
// nn/modules/quant_module.py:192-278, quant_linear.py:38-53) and the 11 GEMMs + loss reductions AWQ-lite runs per
// linear per batch (model_calib.py:1537-1551, :1485-1491): in loss mode the epilogue accumulates
// mean((T(out) - ref)^2) and never writes `out`.
//
// Shape of the kernel (sm_100a, hand-written tcgen05):
//   * persistent grid, one CTA per SM, 192 threads: warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (one elected
//     lane) and TMEM owner, warps 2..5 = epilogue (each reads its own 32-lane quarter of TMEM);
//   * CTA tile 128 x 256, K step 64 (one 128-byte swizzle row of bf16), 4-stage smem ring (4 x 48 KB) filled by
//     cp.async.bulk.tensor (SWIZZLE_128B) and released by tcgen05.commit;
//   * two 256-column fp32 accumulators in TMEM (all 512 columns): the epilogue of tile i overlaps the MMAs of tile i+1;
//   * tiles are walked in groups of 16 M-blocks x all N-blocks so that the group's activation rows and the whole
//     weight stay L2-resident.
// K must be a multiple of 8 and N a multiple of 8 (16-byte TMA strides / vector stores); M is arbitrary (TMA zero-
// fills out-of-range rows, the epilogue masks them).
#include <type_traits>

#include "qdq_ops.cuh"
#include "tc_utils.cuh"

namespace mo {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int UMMA_K = 16;
constexpr int kStages = 4;
constexpr int kGemmThreads = 192;
constexpr int kProThreads = 128;   // 4 extra warps that fake-quantize the A tile on its way into shared memory
constexpr uint32_t kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kGroupM = 16;

struct GemmSmem {
  uint8_t a[kStages][kABytes];
  uint8_t b[kStages][kBBytes];
  uint64_t full[kStages], empty[kStages], tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

// Operand-A prologue: the activation is read raw from global memory, multiplied by pre_quant_scale (optional) and
// fake-quantized per tensor (FP8 E4M3 or INT-N) in registers, then written into the SWIZZLE_128B tile the MMA reads.
struct ProArgs {
  const void* x;       // raw activation [M, K]
  const float* amax;   // fp32 scalar on the device
  const float* pre;    // fp32 [K] pre_quant_scale (values of the operand dtype) or NULL
  float maxb, minb;    // INT bounds
  int eager;           // FP8: MO_NUMERICS_CPU scale arithmetic
};

template <typename T>
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  V32 t;
  t.w[0] = v.x; t.w[1] = v.y; t.w[2] = v.z; t.w[3] = v.w;
  t.w[4] = t.w[5] = t.w[6] = t.w[7] = 0u;
  float g[16];
  Elem<T>::unpack(t, g);
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = g[i];
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float* f) {
  float g[16];
#pragma unroll
  for (int i = 0; i < 8; ++i) { g[i] = f[i]; g[i + 8] = 0.f; }
  V32 t;
  Elem<T>::pack(g, t);
  return make_uint4(t.w[0], t.w[1], t.w[2], t.w[3]);
}

struct TileCoord {
  int m_blk, n_blk;
};
__device__ __forceinline__ TileCoord tile_coord(int tile, int tiles_m, int tiles_n) {
  const int per_group = kGroupM * tiles_n;
  const int g = tile / per_group, r = tile - g * per_group;
  const int gm = min(kGroupM, tiles_m - g * kGroupM);
  TileCoord c;
  c.n_blk = r / gm;
  c.m_blk = g * kGroupM + (r - c.n_blk * gm);
  return c;
}

// kHalf: operands / output are f16 (else bf16).  kLoss: accumulate the squared error against `ref`, no store.
// kPro: 0 = A arrives quantized through TMA; 1 = FP8 prologue; 2 = INT-N prologue (then 4 more warps run).
template <bool kHalf, bool kLoss, int kPro>
__global__ void __launch_bounds__(kGemmThreads + (kPro ? kProThreads : 0), 1)
linear_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      const void* __restrict__ bias_v, void* __restrict__ out_v, const void* __restrict__ ref_v,
                      double* __restrict__ loss_acc, double loss_scale, int M, int N, int K, ProArgs pro) {
  using T = typename std::conditional<kHalf, __half, __nv_bfloat16>::type;
  extern __shared__ uint8_t smem_raw[];
  GemmSmem& s = *reinterpret_cast<GemmSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN, num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    // full: the TMA thread's expect_tx arrive (+ one arrive per prologue warp)
    for (int i = 0; i < kStages; ++i) { mbar_init(&s.full[i], kPro ? 1 + kProThreads / 32 : 1); mbar_init(&s.empty[i], 1); }
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
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(tile, tiles_m, tiles_n);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&s.empty[stage], phase ^ 1);
          mbar_expect_tx(&s.full[stage], kPro ? kBBytes : kStageBytes);
          if constexpr (kPro == 0) tma_load_2d(s.a[stage], &map_a, &s.full[stage], kb * BK, tc.m_blk * BM);
          tma_load_2d(s.b[stage], &map_b, &s.full[stage], kb * BK, tc.n_blk * BN);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor: D = F32, A/B = BF16 (1) or F16 (0), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
      const uint32_t fmt = kHalf ? 0u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
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
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // +32 bytes per UMMA_K step inside the 128-byte swizzle row: +2 in the (>>4) address field
            umma_f16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          }
          umma_commit(&s.empty[stage]);          // smem slot is free once these MMAs have read it
          if (kb == num_kb - 1) umma_commit(&s.tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (kPro != 0 && warp >= kGemmThreads / 32) {
    // ===== operand-A prologue: 128 threads, 8 per 128-byte row segment, 16 rows per pass, 8 passes per stage =====
    const int t = threadIdx.x - kGemmThreads;
    const int chunk = t & 7, r0 = t >> 3;
    const T* x = static_cast<const T*>(pro.x);
    Fp8Op fop;
    IntOp<false> iop;
    if constexpr (kPro == 1) {
      fop.ax.amax = pro.amax; fop.ax.mode = kPerTensor; fop.plain = false; fop.eager = pro.eager != 0;
      fop.init();
    } else {
      iop.ax.amax = pro.amax; iop.ax.mode = kPerTensor; iop.maxb = pro.maxb; iop.minb = pro.minb; iop.status = nullptr;
      iop.init();
    }
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const TileCoord tc = tile_coord(tile, tiles_m, tiles_n);
      const int m0 = tc.m_blk * BM;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int kcol = kb * BK + chunk * 8;
        uint4 v[8];
#pragma unroll
        for (int p = 0; p < 8; ++p) {  // global loads first: they do not depend on the smem slot
          const int row = m0 + r0 + 16 * p;
          v[p] = (row < M && kcol < K) ? __ldg(reinterpret_cast<const uint4*>(x + (int64_t)row * K + kcol))
                                       : make_uint4(0u, 0u, 0u, 0u);
        }
        float ps[8];
        if (pro.pre != nullptr && kcol < K) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(pro.pre + kcol));
          const float4 b = __ldg(reinterpret_cast<const float4*>(pro.pre + kcol + 4));
          ps[0] = a.x; ps[1] = a.y; ps[2] = a.z; ps[3] = a.w; ps[4] = b.x; ps[5] = b.y; ps[6] = b.z; ps[7] = b.w;
        }
        mbar_wait(&s.empty[stage], phase ^ 1);
        uint8_t* tile_a = s.a[stage];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          float f[8];
          unpack8<T>(v[p], f);
          if (pro.pre != nullptr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = __fmul_rn(f[i], ps[i]);
            const uint4 tq = pack8<T>(f);   // the product is a tensor of the operand dtype in the reference
            unpack8<T>(tq, f);
          }
          if constexpr (kPro == 1) fop.template vec<8>(f, 0); else iop.template vec<8>(f, 0);
          const int r = r0 + 16 * p;
          const uint32_t off = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((chunk ^ (r & 7)) << 4));
          *reinterpret_cast<uint4*>(tile_a + off) = pack8<T>(f);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the MMA
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.full[stage]);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue: warps 2..5 own TMEM lanes 32*(warp%4) .. +31 =====
    const int quarter = warp & 3;
    const T* bias = static_cast<const T*>(bias_v);
    T* out = static_cast<T*>(out_v);
    const T* ref = static_cast<const T*>(ref_v);
    int acc = 0;
    uint32_t acc_phase = 0;
    float loss = 0.f;
    double dloss = 0.0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const TileCoord tc = tile_coord(tile, tiles_m, tiles_n);
      mbar_wait(&s.tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = tc.m_blk * BM + quarter * 32 + lane;
      const int n0 = tc.n_blk * BN;
      const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + (uint32_t)c0, r);   // warp-collective: every lane takes part even when its row is masked
        if (row < M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const int col = n0 + c0 + j;
            if (col < N) {
              float f[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(r[j + i]);
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
              if constexpr (kLoss) {
                const uint4 rv = __ldg(reinterpret_cast<const uint4*>(ref + (int64_t)row * N + col));
                const T* rp = reinterpret_cast<const T*>(&rv);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  T dt;
                  Elem<T>::store(&dt, __fsub_rn(Elem<T>::load(op + i), Elem<T>::load(rp + i)));
                  const float d = Elem<T>::load(&dt);
                  loss = __fmaf_rn(d, d, loss);
                }
              } else {
                *reinterpret_cast<uint4*>(out + (int64_t)row * N + col) = o;
              }
            }
          }
        }
      }
      if constexpr (kLoss) { dloss += (double)loss; loss = 0.f; }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if constexpr (kLoss) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dloss += __shfl_xor_sync(0xffffffffu, dloss, o);
      if (lane == 0 && dloss != 0.0) atomicAdd(loss_acc, dloss * loss_scale);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
// [rows, K] row-major 16-bit matrix, box = [box_rows, 64] with the 128-byte swizzle.
static bool make_map(CUtensorMap* m, const void* base, int64_t rows, int64_t K, int box_rows, bool half) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return false; }
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(m, half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return false; }
  return true;
}

template <bool kHalf, bool kLoss, int kPro = 0>
static int linear_launch(const CUtensorMap& ma, const CUtensorMap& mb, const void* bias, void* out, const void* ref,
                         double* loss_acc, int M, int N, int K, cudaStream_t s, ProArgs pro = ProArgs{}) {
  auto kern = linear_tcgen05_kernel<kHalf, kLoss, kPro>;
  const int smem = (int)sizeof(GemmSmem) + 1024;
  static bool configured = false;
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mo_linear: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    configured = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < kNumSMs ? tiles : kNumSMs;
  kern<<<grid, kGemmThreads + (kPro ? kProThreads : 0), smem, s>>>(ma, mb, bias, out, ref, loss_acc,
                                                                    1.0 / ((double)M * (double)N), M, N, K, pro);
  return check_launch("mo_linear");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_linear(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                         int dtype, const void* ref, double* loss_acc, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (dtype != MO_BF16 && dtype != MO_F16) { set_error("mo_linear: bf16 / f16 only (dtype %d)", dtype); return MO_ERR_DTYPE; }
  const bool loss = ref != nullptr || loss_acc != nullptr;
  if (!x || !w || (loss ? (!ref || !loss_acc) : !out)) { set_error("mo_linear: NULL pointer"); return MO_ERR_NULL; }
  if (K < 8 || K % 8 != 0 || N % 8 != 0) { set_error("mo_linear: K and N must be multiples of 8 (K=%lld N=%lld)", (long long)K, (long long)N); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear: dimension too large"); return MO_ERR_SHAPE; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(x) || mis(w) || mis(bias) || mis(out) || mis(ref)) { set_error("mo_linear: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  const bool half = dtype == MO_F16;
  CUtensorMap ma, mb;
  if (!make_map(&ma, x, M, K, BM, half) || !make_map(&mb, w, N, K, BN, half)) return MO_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int m = (int)M, n = (int)N, k = (int)K;
  if (half) return loss ? linear_launch<true, true>(ma, mb, bias, out, ref, loss_acc, m, n, k, s)
                        : linear_launch<true, false>(ma, mb, bias, out, ref, loss_acc, m, n, k, s);
  return loss ? linear_launch<false, true>(ma, mb, bias, out, ref, loss_acc, m, n, k, s)
              : linear_launch<false, false>(ma, mb, bias, out, ref, loss_acc, m, n, k, s);
}

extern "C" int mo_linear_qdq(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                             int dtype, const float* x_amax, const float* prescale, int num_bits, int fp8,
                             int narrow_range, int numerics, void* stream) {
  if (M < 0 || N < 0 || K < 0) { set_error("mo_linear_qdq: bad shape"); return MO_ERR_SHAPE; }
  if (M == 0 || N == 0) return MO_OK;
  if (dtype != MO_BF16 && dtype != MO_F16) { set_error("mo_linear_qdq: bf16 / f16 only (dtype %d)", dtype); return MO_ERR_DTYPE; }
  if (!x || !w || !out || !x_amax) { set_error("mo_linear_qdq: NULL pointer"); return MO_ERR_NULL; }
  if (K < 8 || K % 8 != 0 || N % 8 != 0) { set_error("mo_linear_qdq: K and N must be multiples of 8"); return MO_ERR_SHAPE; }
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) { set_error("mo_linear_qdq: dimension too large"); return MO_ERR_SHAPE; }
  if (!fp8 && (num_bits < 2 || num_bits > 16)) { set_error("mo_linear_qdq: bad num_bits"); return MO_ERR_ARG; }
  auto mis = [](const void* p) { return p && (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(x) || mis(w) || mis(bias) || mis(out) || mis(prescale)) { set_error("mo_linear_qdq: pointers must be 16-byte aligned"); return MO_ERR_ARG; }
  const bool half = dtype == MO_F16;
  CUtensorMap ma, mb;
  if (!make_map(&ma, x, M, K, BM, half) || !make_map(&mb, w, N, K, BN, half)) return MO_ERR_ARG;
  ProArgs pro;
  pro.x = x; pro.amax = x_amax; pro.pre = prescale; pro.eager = numerics == MO_NUMERICS_CPU;
  pro.maxb = (float)((1 << (num_bits - 1)) - 1);
  pro.minb = -(pro.maxb + (narrow_range ? 0.f : 1.f));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int m = (int)M, n = (int)N, k = (int)K;
  if (fp8) return half ? linear_launch<true, false, 1>(ma, mb, bias, out, nullptr, nullptr, m, n, k, s, pro)
                       : linear_launch<false, false, 1>(ma, mb, bias, out, nullptr, nullptr, m, n, k, s, pro);
  return half ? linear_launch<true, false, 2>(ma, mb, bias, out, nullptr, nullptr, m, n, k, s, pro)
              : linear_launch<false, false, 2>(ma, mb, bias, out, nullptr, nullptr, m, n, k, s, pro);
}
