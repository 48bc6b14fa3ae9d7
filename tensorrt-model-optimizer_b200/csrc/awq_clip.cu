// AWQ-clip block search (reference model_calib.py:1800-1858, `_clip_search`, the blocked branch).
//
// For every weight block (output channel `co`, input block `nb`) and every clip ratio r the reference forms the whole
// clipped fake-quantised weight, multiplies it element-wise with the sampled tokens, sums over the block and compares
// with the same sum on the original weight:
//
//     loss[r][co][nb] += mean_t ( sum_j x[t, j] * Wq_r[co, j]  -  sum_j x[t, j] * W[co, j] )^2 ,   j in block nb
//
// which costs 11 full-weight QDQ passes and 2 * 11 [co_bsz, tokens, n_block, block] temporaries per co-batch.  Here
// one kernel does all ratios: a CTA owns one input block column and 64 output channels, stages the sampled tokens of
// that column once in shared memory (transposed, token-minor), and each warp takes one weight block at a time:
// the 32 lanes quantise the block for all ratios (the INT fake-quant of tensor_quant_gpu.cu:40-56, result rounded to
// the weight dtype like the reference's output tensor) and leave d_r[j] = Wq_r[j] - W[j] in shared memory; then every
// lane owns two tokens and accumulates sum_j x[t, j] * d_r[j] for all ratios with the d's read as broadcasts.
// The weight is read exactly once, nothing is written but the [ratios, co, n_blocks] losses.
//
// Summation is fp32 with the difference taken BEFORE the dot product; the reference takes it after two dots rounded to
// the tensor dtype.  For fp32 tensors the two agree to rounding; for bf16 the reference's own value carries the
// rounding noise of two bf16 sums, so only the selected clip values are compared there.
#include "common.cuh"
#include "qdq_ops.cuh"

namespace mo {

constexpr int kClipMaxRatios = 12;   // per launch; the host loops over longer lists
constexpr int kClipTok = 64;         // tokens staged per pass (two per lane)
constexpr int kClipPitch = kClipTok + 1;
constexpr int kClipWarps = 8;
constexpr int kClipRowsPerWarp = 8;
constexpr int kClipRows = kClipWarps * kClipRowsPerWarp;
constexpr int kClipMaxBlock = 256;

struct ClipRatios {
  float r[kClipMaxRatios];
  int n;
};

template <typename T>
__global__ void __launch_bounds__(kClipWarps * 32)
awq_clip_loss_kernel(const T* __restrict__ w, const T* __restrict__ x, int64_t co, int64_t ci, int block,
                     int64_t n_blocks, int64_t n_tok, int64_t tok_stride, const float* __restrict__ w_amax,
                     ClipRatios ratios, int amax_dtype, float minb, float maxb, float* __restrict__ loss) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* xT = reinterpret_cast<float*>(smem_raw);                                   // [block][kClipPitch]
  const int xt_words = (block * kClipPitch + 3) & ~3;
  float4* dall = reinterpret_cast<float4*>(xT + xt_words);                          // [warps][3][block]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4* dsm = dall + (size_t)warp * 3 * block;
  const int64_t nb = blockIdx.x;
  const int64_t col0 = nb * block;
  const int64_t row0 = (int64_t)blockIdx.y * kClipRows + warp * kClipRowsPerWarp;
  const float inv_tok = 1.0f / (float)n_tok;

  for (int64_t t0 = 0; t0 < n_tok; t0 += kClipTok) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < block * kClipTok; idx += kClipWarps * 32) {
      const int t = idx / block, j = idx - t * block;
      float v = 0.f;
      if (t0 + t < n_tok && col0 + j < ci) v = Elem<T>::load(x + (t0 + t) * tok_stride + col0 + j);
      xT[j * kClipPitch + t] = v;
    }
    __syncthreads();
    for (int i = 0; i < kClipRowsPerWarp; ++i) {
      const int64_t r = row0 + i;
      if (r >= co) break;
      const float am = __ldg(w_amax + r * n_blocks + nb);
      // --- quantise the block for every ratio, keep the differences ---
      for (int j = lane; j < block; j += 32) {
        const float wv = col0 + j < ci ? Elem<T>::load(w + r * ci + col0 + j) : 0.f;
        float d[kClipMaxRatios];
#pragma unroll
        for (int k = 0; k < kClipMaxRatios; ++k) {
          d[k] = 0.f;
          if (k < ratios.n) {
            float a = __fmul_rn(am, ratios.r[k]);                // `w_amax * shrink` (:1826), in the amax dtype
            if (amax_dtype == MO_BF16) a = __bfloat162float(__float2bfloat16_rn(a));
            else if (amax_dtype == MO_F16) a = __half2float(__float2half_rn(a));
            float o = 0.f;
            if (!(a < kEps)) {                                    // tensor_quant_gpu.cu:46
              const float s = __fdiv_rn(maxb, a);
              float v = __fmul_rn(wv, s);
              v = max_nan(min_nan(v, maxb), minb);
              v = __fsub_rn(__fadd_rn(v, 12582912.0f), 12582912.0f);
              o = __fdiv_rn(v, s);
            }
            T t;
            Elem<T>::store(&t, o);
            d[k] = __fsub_rn(Elem<T>::load(&t), wv);
          }
        }
        dsm[j] = make_float4(d[0], d[1], d[2], d[3]);
        dsm[block + j] = make_float4(d[4], d[5], d[6], d[7]);
        dsm[2 * block + j] = make_float4(d[8], d[9], d[10], d[11]);
      }
      __syncwarp();
      // --- per-token dot products: lane owns tokens `lane` and `lane + 32` of the chunk ---
      float a0[kClipMaxRatios], a1[kClipMaxRatios];
#pragma unroll
      for (int k = 0; k < kClipMaxRatios; ++k) a0[k] = a1[k] = 0.f;
      for (int j = 0; j < block; ++j) {
        const float x0 = xT[j * kClipPitch + lane], x1 = xT[j * kClipPitch + lane + 32];
        const float4 p = dsm[j], q = dsm[block + j], u = dsm[2 * block + j];
        const float dv[kClipMaxRatios] = {p.x, p.y, p.z, p.w, q.x, q.y, q.z, q.w, u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < kClipMaxRatios; ++k) {
          a0[k] = __fmaf_rn(x0, dv[k], a0[k]);
          a1[k] = __fmaf_rn(x1, dv[k], a1[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < kClipMaxRatios; ++k) {
        if (k >= ratios.n) break;
        float sq = __fmaf_rn(a0[k], a0[k], __fmul_rn(a1[k], a1[k]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if (lane == 0) loss[((int64_t)k * co + r) * n_blocks + nb] += sq * inv_tok;  // this warp owns the entry
      }
      __syncwarp();
    }
  }
}

template <typename T>
static int clip_launch(const void* w, const void* x, int64_t co, int64_t ci, int block, int64_t n_tok,
                       int64_t tok_stride, const float* w_amax, const ClipRatios& ratios, int amax_dtype, float minb,
                       float maxb, float* loss, cudaStream_t s) {
  const int64_t n_blocks = (ci + block - 1) / block;
  const size_t smem = (size_t)((block * kClipPitch + 3) & ~3) * sizeof(float) + (size_t)kClipWarps * 3 * block * sizeof(float4);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(awq_clip_loss_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_error("mo_awq_clip_loss: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e)); return (int)e; }
    configured = smem;
  }
  const int64_t row_groups = (co + kClipRows - 1) / kClipRows;
  if (row_groups > 65535) { set_error("mo_awq_clip_loss: too many output channels (%lld)", (long long)co); return MO_ERR_SHAPE; }
  dim3 grid((unsigned)n_blocks, (unsigned)row_groups);
  awq_clip_loss_kernel<T><<<grid, kClipWarps * 32, smem, s>>>(static_cast<const T*>(w), static_cast<const T*>(x), co, ci,
                                                             block, n_blocks, n_tok, tok_stride, w_amax, ratios,
                                                             amax_dtype, minb, maxb, loss);
  return check_launch("mo_awq_clip_loss");
}

}  // namespace mo

using namespace mo;

extern "C" int mo_awq_clip_loss(const void* w, const void* x, int64_t co, int64_t ci, int block, int64_t n_tok,
                                int64_t tok_stride, int dtype, const float* w_amax, int amax_dtype,
                                const float* ratios_host,
                                int n_ratio, int num_bits, int is_unsigned, int narrow_range, float* loss,
                                void* stream) {
  if (co < 0 || ci < 0 || n_tok < 0 || block < 1 || block > kClipMaxBlock || tok_stride < ci) {
    set_error("mo_awq_clip_loss: bad shape (co=%lld ci=%lld block=%d n_tok=%lld tok_stride=%lld; block <= %d)",
              (long long)co, (long long)ci, block, (long long)n_tok, (long long)tok_stride, kClipMaxBlock);
    return MO_ERR_SHAPE;
  }
  if (n_ratio < 1 || n_ratio > kClipMaxRatios) {
    set_error("mo_awq_clip_loss: n_ratio must be in [1, %d], got %d", kClipMaxRatios, n_ratio);
    return MO_ERR_ARG;
  }
  if (amax_dtype != MO_F32 && amax_dtype != MO_F16 && amax_dtype != MO_BF16) {
    set_error("mo_awq_clip_loss: unsupported amax_dtype %d", amax_dtype);
    return MO_ERR_DTYPE;
  }
  if (num_bits < 2 || num_bits > 16) { set_error("mo_awq_clip_loss: num_bits %d out of range", num_bits); return MO_ERR_ARG; }
  if (co == 0 || ci == 0 || n_tok == 0) return MO_OK;
  if (!w || !x || !w_amax || !ratios_host || !loss) { set_error("mo_awq_clip_loss: NULL pointer"); return MO_ERR_NULL; }
  ClipRatios ratios;
  for (int k = 0; k < kClipMaxRatios; ++k) ratios.r[k] = k < n_ratio ? ratios_host[k] : 0.f;
  ratios.n = n_ratio;
  const float bound = (float)((1 << (num_bits - 1 + (is_unsigned ? 1 : 0))) - 1);
  const float maxb = bound;
  const float minb = -(bound + (narrow_range ? 0.f : 1.f));  // tensor_quant_gpu.cu:69 (also for unsigned)
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case MO_F32: return clip_launch<float>(w, x, co, ci, block, n_tok, tok_stride, w_amax, ratios, amax_dtype, minb, maxb, loss, s);
    case MO_F16: return clip_launch<__half>(w, x, co, ci, block, n_tok, tok_stride, w_amax, ratios, amax_dtype, minb, maxb, loss, s);
    case MO_BF16: return clip_launch<__nv_bfloat16>(w, x, co, ci, block, n_tok, tok_stride, w_amax, ratios, amax_dtype, minb, maxb, loss, s);
  }
  set_error("mo_awq_clip_loss: unsupported dtype %d", dtype);
  return MO_ERR_DTYPE;
}
