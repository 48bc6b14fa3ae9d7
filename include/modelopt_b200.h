/* modelopt_b200 — C ABI of the B200 (sm_100a) quantization hot path.
 *
 * Drop-in boundary for ModelOpt's TensorQuantizer hot path: these are the entry points the
 * reference's Python binds today through three JIT-compiled pybind extensions and a set of Triton
 * kernels.  Each declaration cites the reference interface it replaces (paths relative to the
 * reference checkout).  No torch types cross this boundary: plain pointers, sizes, a dtype code and
 * a `cudaStream_t` passed as `void*`.  Unless a name ends in `_host`, every data pointer is a DEVICE
 * pointer valid on the current CUDA device and the call only enqueues work on `stream` (no
 * host-device synchronisation).  `_host` entry points take HOST pointers, stage through pinned
 * buffers owned by the library and return after the result is back in host memory.
 *
 * Return value: 0 on success; MO_ERR_* (<0) for a rejected argument; a positive `cudaError_t`
 * otherwise.  `mo_last_error()` returns a thread-local description of the last failure.
 */
#ifndef MODELOPT_B200_H_
#define MODELOPT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MO_ABI_VERSION 4

typedef enum { MO_F32 = 0, MO_F16 = 1, MO_BF16 = 2 } mo_dtype;

enum {
  MO_OK = 0,
  MO_ERR_DTYPE = -1,    /* unsupported dtype code */
  MO_ERR_SHAPE = -2,    /* inconsistent sizes */
  MO_ERR_ARG = -3,      /* bad scalar argument (num_bits, block size, ...) */
  MO_ERR_NULL = -4,     /* required pointer is NULL */
  MO_ERR_ALIGN = -5,    /* pointer not aligned to the element size */
  MO_ERR_NO_DEVICE = -6 /* no sm_100 device / driver */
};

/* Bits OR-ed into the optional device status word by kernels that can detect bad input. */
#define MO_FLAG_NEGATIVE_IN_UNSIGNED 1u /* tensor_quant_gpu.cu:65-67 CUDA_KERNEL_ASSERT(inputs >= 0) */
#define MO_FLAG_NEGATIVE_AMAX 2u        /* tensor_quant_gpu.cu:44   CUDA_KERNEL_ASSERT(amax >= 0)   */

/* Numerics mode: the reference itself computes slightly different scale arithmetic depending on
 * where it runs, and "bit-exact against the reference" has to say which.
 *   MO_NUMERICS_GPU  what the reference executes on an sm_100a GPU: div.full.f32 inside its Triton
 *                    NVFP4 kernels, `tensor / python_scalar` = multiply by the fp32 reciprocal on the
 *                    host side, true division 448.f / amax in the FP8 CUDA extension
 *                    (gemm/tensor_quant_gpu_fp8.cu:96-98).  DEFAULT for the product.
 *   MO_NUMERICS_CPU  what the reference executes on CPU tensors (the oracle): IEEE division everywhere,
 *                    and Python `448.0 / amax` = reciprocal(amax) * 448 (two roundings; torch
 *                    Tensor.__rtruediv__) in `_fp8_eager` (tensor_quant.py:46-59).                     */
#define MO_NUMERICS_GPU 0
#define MO_NUMERICS_CPU 1
#define MO_DIV_TRITON MO_NUMERICS_GPU
#define MO_DIV_IEEE MO_NUMERICS_CPU

int mo_abi_version(void);
const char* mo_last_error(void);
/* Name of device 0, its SM count and compute capability (major*10+minor); MO_ERR_NO_DEVICE if none. */
int mo_device_info(char* name, int name_len, int* sm_count, int* cc);

/* ------------------------------------------------------------------------------------------------
 * Fake quantize-dequantize (row a1-a4 of SURVEY.md section 8).
 *
 * Tensors are contiguous with `n` elements.  Per-tensor: amax has 1 element (axis_size = 1,
 * inner = n).  Per-axis / per-block: the tensor is viewed [outer, axis_size, inner] and element
 * idx uses amax[(idx / inner) % axis_size]; the reference's INT4 block-128 path is the view
 * [n/128, 128] with axis_size = n/128, inner = 128 (tensor_quant.py:107-111).
 * `amax` is fp32 on the device (the reference converts with amax.to(float), tensor_quant_gpu.cu:83).
 * In-place (y == x) is allowed.
 * ---------------------------------------------------------------------------------------------- */

/* Replaces cuda_ext.fake_tensor_quant / fake_tensor_quant_with_axis / fake_tensor_quant_
 * (gemm/tensor_quant.cpp:36-62, gemm/tensor_quant_gpu.cu:58-140).
 * status: optional device uint32 that receives MO_FLAG_* bits (may be NULL). */
int mo_qdq_int(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
               int64_t inner, int num_bits, int is_unsigned, int narrow_range, uint32_t* status,
               void* stream);

/* Replaces cuda_ext_fp8.fake_e4m3fy / fake_e4m3fy_with_axis (gemm/tensor_quant_gpu_fp8.cu:35-107)
 * including the host-side torch ops that build scale = 448/amax and inv_scale = 1/scale (:93-98).
 * amax == NULL: plain cast through E4M3 (tensor_quant.py:46-59 with amax=None). */
int mo_qdq_fp8(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size,
               int64_t inner, int numerics, void* stream);

/* Replaces Triton fp4_fake_quant_block (gemm/fp4_kernel_hopper.py:102-170): dynamic NVFP4 over the
 * last dim of a [rows, cols] view; `global_amax` is the calibrated fp32 scalar (device).
 * block in {16, 32, 64}; a ragged last block is zero-padded like the reference's tile loads. */
int mo_qdq_nvfp4_dynamic(const void* x, void* y, int64_t rows, int64_t cols, int dtype,
                         const float* global_amax, int block, int div_mode, void* stream);

/* Replaces Triton static_blockwise_fp4_fake_quant_kernel (gemm/fp4_kernel.py:193-214, scalar math
 * common/nvfp4_quant.py:66-101): x is n_blocks*block contiguous elements, `scale` is the fp32
 * per-block scale (device) as produced by compute_fp4_scales (:217-251; see mo_fp4_scales). */
int mo_qdq_nvfp4_static(const void* x, void* y, int64_t n_blocks, int block, int in_dtype,
                        int out_dtype, const float* scale, int div_mode, void* stream);

/* The two steps above in ONE launch: `amax` is the fp32 per-block amax, the scale is derived per thread
 * (static_blockwise_fp4_fake_quant = compute_fp4_scales + the kernel, gemm/fp4_kernel.py:217-316).  Same dtype in and
 * out, block in {16,32,64}, 32-byte aligned pointers; otherwise returns MO_ERR_ALIGN without launching and the caller
 * uses mo_fp4_scales + mo_qdq_nvfp4_static.  global_amax is required when quantize_block_scales != 0. */
int mo_qdq_nvfp4_static_amax(const void* x, void* y, int64_t n_blocks, int block, int dtype, const float* amax,
                             const float* global_amax, int quantize_block_scales, float fp8_max_for_normalization,
                             int div_mode, void* stream);

/* Replaces compute_fp4_scales (gemm/fp4_kernel.py:217-251): scale = amax/6, optionally round-tripped
 * through E4M3 under scale_amax = global_amax*(448/fp8_max)/6.  global_amax == NULL: reduce it from
 * `amax` first (two launches).  amax/scale fp32 device arrays of n_blocks. */
int mo_fp4_scales(const float* amax, float* scale, int64_t n_blocks, const float* global_amax,
                  int quantize_block_scales, float fp8_max_for_normalization, int div_mode,
                  uint32_t* scratch /* 4-byte device word, needed only when global_amax == NULL */,
                  void* stream);

/* Replaces Triton static_blockwise_fp4_cast_kernel (gemm/fp4_kernel.py:319-400). */
int mo_fp4_cast(const void* x, void* y, int64_t n, int in_dtype, int out_dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Amax reductions (rows a7, a8): replaces reduce_amax / reduce_block_amax
 * (quantization/utils/core_utils.py:43-89, :146-183) and the running max of MaxCalibrator.collect
 * (calib/max.py:52-86).
 *
 * The tensor is viewed [outer, axis_size, inner]; out[a] = max over (o, i) of |x[o, a, i]|, MERGED
 * into `amax_bits` with an unsigned max on the fp32 bit pattern (one atomic per CTA and channel).
 * The caller zero-fills amax_bits before the first batch and reinterprets it as fp32 afterwards:
 * max is exact and order-independent, NaN inputs survive as a NaN pattern (torch.max semantics) and
 * inf as inf, so the reference's three host-syncing asserts (max.py:69-77) become one check at
 * finish time.  Per-tensor: axis_size = 1.
 * ---------------------------------------------------------------------------------------------- */
int mo_amax(const void* x, int64_t outer, int64_t axis_size, int64_t inner, int dtype,
            uint32_t* amax_bits, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Histogram calibration (rows a9, a10): replaces HistogramCalibrator.collect and the three amax
 * searches (calib/histogram.py:77-130, :210-343).
 * ---------------------------------------------------------------------------------------------- */
/* One read of x: bins |x| with torch.histc's rule (bin = (int)(v * nbins / range_max), v in
 * [0, range_max], last bin closed; range_max == 0 -> range [-1, 1]) into `counts` (device uint64[nbins],
 * ADDED to: exact integer counts, where the reference's fp32 histc saturates at 2^24 per bin), and
 * merges max|x| into `amax_bits` (optional, as mo_amax).  range_max is a DEVICE fp32 scalar.
 * skip_zeros: histogram.py:107-108. */
int mo_hist(const void* x, int64_t n, int dtype, const float* range_max, int64_t nbins, uint64_t* counts,
            uint32_t* amax_bits, int skip_zeros, void* stream);

/* HistogramCalibrator.collect (histogram.py:77-130) WITHOUT a host round trip per batch.  The calibrator's parameters live
 * in `state` (device int32[MO_HIST_STATE_WORDS], zero-initialised by the caller) and `counts` has a fixed `capacity`
 * (bins beyond the current nbins stay zero).  Per batch the caller runs mo_hist_collect_begin, then - data-parallel
 * calibration only - an all-reduce(MAX) over state[MO_HIST_BATCH_AMAX] (one word), then mo_hist_collect_end:
 *   first = 1 : begin = max|x|;  end = nbins := init_bins, range := that max (histogram.py:116-118), bin x.
 *   first = 0 : begin = bin x under the current range and fold max|x| in the same read;  end = if the max exceeds the
 *               range, grow it exactly like histogram.py:121-126 (num_bins = ceil(x_max / width), range = last edge of
 *               arange(0, x_max + width, width); numerics = MO_NUMERICS_CPU evaluates arange like the CPU kernel, in
 *               fp64) and move the batch from the old bins to the new ones (one more read of x, only on such batches).
 * With the all-reduce every rank takes the same decisions, so its counts are bin-compatible with every other rank's and
 * their SUM equals one process calibrating on the concatenated batches.  A range needing more than `capacity` bins, or
 * a non-finite max, sets a flag in state[MO_HIST_FLAGS] and freezes the state (later launches are no-ops); the caller
 * reads the state back whenever it likes (compute_amax does).  smem_bins_hint: how many bins to keep in shared-memory
 * counters (any value is correct; nbins or a little more is fastest; 0 = capacity). */
enum {
  MO_HIST_NBINS = 0, MO_HIST_RANGE_MAX = 1, MO_HIST_WIDTH = 2, MO_HIST_LAST_END = 3, MO_HIST_FLAGS = 4,
  MO_HIST_FIRST_AMAX = 5, MO_HIST_BATCH_AMAX = 6, MO_HIST_GROWN = 7, MO_HIST_OLD_NBINS = 8, MO_HIST_OLD_RANGE_MAX = 9,
  MO_HIST_BATCHES = 10, MO_HIST_CAPACITY = 11, MO_HIST_STATE_WORDS = 16
};
#define MO_HIST_FLAG_OVERFLOW 1   /* the range would need more than `capacity` bins */
#define MO_HIST_FLAG_NONFINITE 2  /* max|x| was inf or NaN (torch.histc raises "range ... is not finite") */
int mo_hist_collect_begin(const void* x, int64_t n, int dtype, int32_t* state, uint64_t* counts, int64_t capacity,
                          int64_t smem_bins_hint, int skip_zeros, int first, void* stream);
int mo_hist_collect_end(const void* x, int64_t n, int dtype, int32_t* state, uint64_t* counts, int64_t capacity,
                        int64_t smem_bins_hint, int skip_zeros, int first, int init_bins, int numerics, void* stream);
/* begin + end in one call (single process: nothing to exchange between the halves). */
int mo_hist_collect(const void* x, int64_t n, int dtype, int32_t* state, uint64_t* counts, int64_t capacity,
                    int64_t smem_bins_hint, int skip_zeros, int first, int init_bins, int numerics, void* stream);

/* _compute_amax_percentile (histogram.py:328-343): *idx_out (device) = searchsorted(cumsum(hist/total),
 * percentile/100) in fp64, numpy's summation order; the amax is edges[idx]. */
int mo_hist_percentile(const uint64_t* counts, int64_t nbins, double percentile, int32_t* idx_out,
                       void* stream);

/* _compute_amax_mse (histogram.py:285-325): *idx_out = argmin over candidates i = start_bin + k*stride of
 * mean((fq(centres; amax = centre_i) - centres)^2 * counts); the amax is centre[start_bin + idx*stride].
 * as_written = 1 reproduces the reference's mis-slotted call (bias <- num_bits, num_bits <- unsigned,
 * :308); 0 evaluates the intended INT-N (fp8 = 0) or E4M3 (fp8 = 1) fake-quant.
 * edges: device fp32[nbins + 1]; loss_scratch: device double[number of candidates]. */
int mo_hist_mse(const uint64_t* counts, const float* edges, int64_t nbins, int num_bits, int is_unsigned,
                int fp8, int as_written, int stride, int start_bin, double* loss_scratch, int32_t* idx_out,
                void* stream);

/* _compute_amax_entropy (histogram.py:210-282): TensorRT-style KL search in fp64; *idx_out = index of the
 * LAST minimum over candidates i = start_bin, start_bin + stride, ..., <= nbins; the amax is
 * edges[start_bin + idx*stride].  scratch: device buffer of mo_hist_entropy_scratch_bytes(). */
int64_t mo_hist_entropy_scratch_bytes(int64_t nbins, int stride, int start_bin);
int mo_hist_entropy(const uint64_t* counts, int64_t nbins, int num_bits, int is_unsigned, int stride,
                    int start_bin, void* scratch, int32_t* idx_out, void* stream);

/* mo_hist_entropy for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  NO counterpart in the reference.  Same search, same
 * result as mo_hist_entropy; the scratch is a fixed file-scope device array inside the library (1.5 MiB, held by
 * every GPU that loads the library until the process exits), so the call performs no cudaMalloc / cudaFree and calls
 * no allocating library routine.  Limit: nbins <= mo_hist_entropy_static_max_bins() (65536); more returns
 * MO_ERR_ARG with a message - use mo_hist_entropy with your own scratch then.  All calls on one device share the
 * array (it keeps the last call's contents; nothing in it needs to start at zero): calls that overlap on the GPU
 * (different streams, parallel graph branches) race - serialise them. */
int64_t mo_hist_entropy_static_max_bins(void);
int mo_hist_entropy_static(const uint64_t* counts, int64_t nbins, int num_bits, int is_unsigned, int stride,
                           int start_bin, int32_t* idx_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * NVFP4 real-quant packer (row a16): replaces NVFP4QTensor.quantize / dequantize
 * (qtensor/nvfp4_tensor.py:253-407) and the Triton fp4_dequantize (gemm/fp4_kernel.py:54-190).
 * x: n_blocks * block contiguous elements (last dim already padded to a multiple of `block`,
 * reduce_block_padding :279).  sf2: device fp32 scalar = amax / (6 * 448) (:204-207).
 * codes: uint8[n_blocks * block / 2], element 2j in the LOW nibble, nibble = sign << 3 | e2m1 index (:230-250, :337).
 * sf_out: uint8[n_blocks] E4M3 bit patterns of the per-block scales (:169-201, :32-46); if sf_in != NULL the
 * given scales are used instead (weights_scaling_factor argument) and sf_out may be NULL.
 * ---------------------------------------------------------------------------------------------- */
int mo_nvfp4_pack(const void* x, int64_t n_blocks, int block, int dtype, const float* sf2, const uint8_t* sf_in,
                  uint8_t* codes, uint8_t* sf_out, void* stream);
/* fast = 0: value * (sf * sf2) (nvfp4_tensor.py:396-404); fast = 1: (value * sf) * sf2 (Triton fp4_dequantize). */
int mo_nvfp4_unpack(const uint8_t* codes, const uint8_t* sf, const float* sf2, int64_t n_blocks, int block,
                    int out_dtype, int fast, void* y, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Scale searches (row a11): the tensor is read once for every candidate.
 * ---------------------------------------------------------------------------------------------- */
/* Replaces Triton nvfp4_fp8_scale_sweep (gemm/nvfp4_fp8_sweep.py:57-162) and the 126-pass Python sweep of
 * NVFP4MSECalibrator (calib/mse.py:175-311): for each block of `block` elements try the 126 positive finite
 * E4M3 values c = e4m3/448 (gemm/_fp8_scale_candidates.py:26-31) as scale = c*global_amax/6, keep the FIRST
 * minimum of sum((|w| - e2m1(|w|/scale)*scale)^2); best_amax[b] = global_amax * c_best (fp32, device). */
int mo_nvfp4_fp8_scale_sweep(const void* x, int64_t n_blocks, int block, int dtype, const float* global_amax,
                             float* best_amax, int numerics, void* stream);
/* Replaces the K fake-quant + mse_loss + reduce passes of MseCalibrator.collect (calib/mse.py:83-119):
 * losses[k*axis_size + a] += sum over the [outer, axis_size, inner] view of (x - fq(x, cand_amax[k*axis_size + a]))^2,
 * fp64 accumulators (device, caller zero-fills before the first batch).  cand_amax is the [n_cand, axis_size] fp32
 * table initial_amax * candidate the caller forms (calib/mse.py:76-81).  fp8 = 1: E4M3; else INT-N. */
int mo_mse_sweep(const void* x, int64_t outer, int64_t axis_size, int64_t inner, int dtype, const float* cand_amax,
                 int n_cand, int num_bits, int is_unsigned, int narrow_range, int fp8, int numerics, double* losses,
                 void* stream);

/* ------------------------------------------------------------------------------------------------
 * SmoothQuant / AWQ-lite support (rows a13, a14) and the fused pre_quant_scale paths of row a5.
 * ---------------------------------------------------------------------------------------------- */
/* INT-N fake quant whose amax is the max|.| of every `block` consecutive elements of the data itself
 * (TensorQuantizer with block_sizes and no calibrated amax: tensor_quantizer.py:736-751 -> reduce_amax on the
 * [n/block, block] view), with an optional per-column multiply in front (pre_quant_scale, :1143; `prescale`:
 * [cols] values IN THE TENSOR'S DTYPE, cols % block == 0).  Replaces the `w*s -> reshape -> amax -> QDQ` chain
 * AWQ-lite runs 11x per linear per batch (model_calib.py:1537-1551).  amax_out (nullable): fp32 [n_blocks]. */
int mo_qdq_int_block_dynamic(const void* x, void* y, int64_t n_blocks, int block, int dtype, const void* prescale,
                             int64_t cols, int num_bits, int is_unsigned, int narrow_range, float* amax_out,
                             void* stream);
/* mo_qdq_int / mo_qdq_fp8 with TensorQuantizer's `inputs * pre_quant_scale` (tensor_quantizer.py:1143) folded into
 * the same pass: p = T(x * prescale[i % cols]) is quantized instead of x.  prescale: [cols] values in the tensor's
 * dtype (`dtype`; one 32-byte load per vector instead of 64 bytes of fp32), or NULL (identical to the plain entry
 * points). */
int mo_qdq_int_pqs(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size, int64_t inner,
                   int num_bits, int is_unsigned, int narrow_range, uint32_t* status, const void* prescale,
                   int64_t cols, void* stream);
int mo_qdq_fp8_pqs(const void* x, void* y, int64_t n, int dtype, const float* amax, int64_t axis_size, int64_t inner,
                   int numerics, const void* prescale, int64_t cols, void* stream);
/* y = T(x * scale[col]) over a [rows, cols] matrix (y == x allowed), and in the same pass the amax of y:
 * amax_mode 0 none, 1 per tensor -> amax_bits[0], 2 per row -> amax_bits[row] (bit patterns, merged with max).
 * Replaces the weight fold + max re-calibration of apply_pre_quant_scale_and_smooth / _apply_weight_pre_quant_scale
 * (model_calib.py:1208-1270) and the stand-alone `inputs * pre_quant_scale` pass (tensor_quantizer.py:1143). */
int mo_scale_cols_amax(const void* x, void* y, int64_t rows, int64_t cols, int dtype, const float* scale,
                       uint32_t* amax_bits, int amax_mode, void* stream);
/* AWQ get_weight_scale (model_calib.py:1453-1469): partial[s*cols + c] = sum over the rows of split s of
 * T(|w[r,c]| / T(blockmax(r, c/block) + tiny(T))); the caller adds the `splits` rows and divides by `rows`. */
int mo_awq_weight_scale(const void* w, int64_t rows, int64_t cols, int dtype, int block, float* partial, int splits,
                        void* stream);
/* AWQ get_act_scale (model_calib.py:1471-1472): partial[s*cols + c] = sum over the rows of split s of |x[r,c]|. */
int mo_abs_col_sum(const void* x, int64_t rows, int64_t cols, int dtype, float* partial, int splits, void* stream);
/* AWQ update_loss (model_calib.py:1485-1491): acc[0] += mean((T(a - b))^2) (fp64 device accumulator). */
int mo_sq_diff_mean(const void* a, const void* b, int64_t n, int dtype, double* acc, void* stream);
/* AWQ-clip block search (model_calib.py:1800-1858, `_clip_search`, blocked branch): for every weight block
 * (row r of `co`, block nb of ceil(ci/block)) and every clip ratio k < n_ratio (<= 12 per call),
 *   loss[(k*co + r)*n_blocks + nb] += mean over the n_tok sampled tokens of
 *       ( sum_j x[t, j] * (QDQ(w[r, j]; amax = w_amax[r, nb] * ratios[k]) - w[r, j]) )^2 ,  j in block nb,
 * QDQ being the INT fake-quant of mo_qdq_int rounded to `dtype`; the product w_amax * ratio is rounded to `amax_dtype`
 * first (the reference multiplies the amax tensor in ITS dtype: bf16 for a bf16 model).  Token t is the row x + t*tok_stride (the
 * reference's `inputs[0::stride]` sampling is a stride, not a copy).  `ratios` is a HOST array; w, x, w_amax, loss are
 * device pointers; block <= 256; columns past ci count as zeros (the reference pads). */
int mo_awq_clip_loss(const void* w, const void* x, int64_t co, int64_t ci, int block, int64_t n_tok,
                     int64_t tok_stride, int dtype, const float* w_amax, int amax_dtype, const float* ratios,
                     int n_ratio,
                     int num_bits, int is_unsigned, int narrow_range, float* loss, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Quant-aware linear (row a15): hand-written tcgen05 GEMM, fp32 accumulation in TMEM.
 * ---------------------------------------------------------------------------------------------- */
/* out[M,N] = T(x[M,K] . w[N,K]^T + bias[N]) with x, w, bias, out of `dtype` (MO_BF16 / MO_F16), row-major.
 * Replaces the F.linear inside QuantLinearConvBase.forward / _QuantLinear.quantized_linear_fn
 * (nn/modules/quant_module.py:192-278, quant_linear.py:38-53); x and w are the fake-quantized operands.
 * Loss mode (ref != NULL): nothing is stored, loss_acc[0] += mean((T(out) - ref)^2) (fp64 device accumulator) -
 * the GEMM + update_loss pair of the AWQ-lite search (model_calib.py:1537-1551, :1485-1491).
 * K and N must be multiples of 8, pointers 16-byte aligned; bias may be NULL. */
int mo_linear(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K, int dtype,
              const void* ref, double* loss_acc, void* stream);

/* mo_linear with the input quantizer fused into the GEMM's operand-A prologue: x is the RAW activation; four extra
 * warps read each 128x64 tile from global memory, apply `prescale` (pre_quant_scale, fp32 [K], may be NULL) and the
 * per-tensor fake-quant (fp8 = 1: E4M3 with `numerics`; else INT-`num_bits`, signed) in registers - the same arithmetic
 * as mo_qdq_fp8_pqs / mo_qdq_int_pqs - and store the result straight into the swizzled shared-memory tile the MMA
 * reads: `input_q(x)` never goes to HBM.  w is the (cached) fake-quantized weight.  x_amax: fp32 scalar on the device. */
int mo_linear_qdq(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K, int dtype,
                  const float* x_amax, const float* prescale, int num_bits, int fp8, int narrow_range, int numerics,
                  void* stream);

/* NVFP4 W4A4 linear on the block-scaled tensor-core path (tcgen05.mma.kind::mxf4nvf4.block_scale), the role of the
 * reference's real-quant NVFP4 GEMM backend (backends/nvfp4_gemm.py:28-125):
 *   out[m,n] = T(a_gs*b_gs * sum_k (a[m,k]*a_sf[m,k/16]) * (b[n,k]*b_sf[n,k/16]) + bias[n])
 * a_codes [M, K/2], b_codes [N, K/2]: packed E2M1 as mo_nvfp4_pack writes them; a_sf / b_sf: the E4M3 block scales
 * re-ordered by mo_nvfp4_sf_swizzle (row_pad 128 for a, 256 for b); a_gs / b_gs: fp32 global scales on the device.
 * K % 256 == 0, N % 8 == 0; out / bias dtype MO_BF16 or MO_F16. */
int mo_linear_nvfp4(const void* a_codes, const void* a_sf, const float* a_gs, const void* b_codes, const void* b_sf,
                    const float* b_gs, const void* bias, void* out, int64_t M, int64_t N, int64_t K, int out_dtype,
                    void* stream);
/* The same product (same operands, same results) on CTA pairs: tcgen05 cta_group::2, 256 x 256 pair tile, each CTA holds
 * half of the B tile (the single-CTA kernel is bound by shared-memory bandwidth), scale factors carried into tensor
 * memory by dedicated warps, overlapped accumulators, TMA-store epilogue.  Same reference role as mo_linear_nvfp4
 * (backends/nvfp4_gemm.py:28-125). */
int mo_linear_nvfp4_pair(const void* a_codes, const void* a_sf, const float* a_gs, const void* b_codes, const void* b_sf,
                         const float* b_gs, const void* bias, void* out, int64_t M, int64_t N, int64_t K, int out_dtype,
                         void* stream);
/* E4M3 block scales [rows, kblocks] row-major (mo_nvfp4_pack's sf_out) -> the 128-row x 4-block tile order of
 * tcgen05 scale factors, rows zero-padded to a multiple of row_pad (128 or 256), kblocks to a multiple of 4. */
int64_t mo_nvfp4_sf_swizzle_bytes(int64_t rows, int64_t kblocks, int row_pad);
/* mo_nvfp4_pack (block 16) for a [rows, cols] GEMM operand, cols % 64 == 0, writing the block scales directly in the
 * swizzled order (buffer of mo_nvfp4_sf_swizzle_bytes(rows, cols/16, pad) bytes; the caller zero-fills it when rows is
 * not a multiple of the pad): activation quantization for mo_linear_nvfp4 in ONE kernel. */
int mo_nvfp4_pack_gemm(const void* x, int64_t rows, int64_t cols, int dtype, const float* sf2, uint8_t* codes,
                       uint8_t* sf_swizzled, void* stream);
int mo_nvfp4_sf_swizzle(const void* sf, int64_t rows, int64_t kblocks, int row_pad, void* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * MX formats and the remaining real-quant packers (rank 1).
 * ---------------------------------------------------------------------------------------------- */
/* Element / scale format codes = the reference's `Types` enum (gemm/tensor_quant_mx.h:39). */
enum {
  MO_FMT_E4M3 = 0, MO_FMT_E5M2 = 1, MO_FMT_INT8 = 2, MO_FMT_E0M3 = 3, MO_FMT_E1M2 = 4, MO_FMT_E3M0 = 5,
  MO_FMT_E2M1 = 6, MO_FMT_E3M2 = 7, MO_FMT_E2M3 = 8, MO_FMT_E8M0 = 9
};

/* Replaces cuda_ext_mx.fused_amax_convert(inputs, block_size, format, scale_format, global_amax)
 * (gemm/tensor_quant_mx.cu:347-387; called from tensor_quant.py:184-191 for MXFP8 / MXFP6 / MXFP4 / MXINT8 and every
 * other dynamic block format): per block of `block` in {8,16,32} elements of the last dim (a short last block is
 * zero-padded) compute amax, derive (scale, unscale) - E8M0: unscale = 2^ceil(log2(amax / format_max)) (:104-132);
 * other scale formats: (:135-181) with the optional fp32 device scalar global_amax - and write
 * y = sign(x) * format(|x| * scale) * unscale.  x / y: [rows, cols] contiguous, y == x allowed.
 * numerics: MO_NUMERICS_GPU = the fp32 quotients are div.approx.ftz.f32 as in the reference's --use_fast_math build
 * (quantization/extensions.py:69); MO_NUMERICS_CPU = IEEE quotients. */
int mo_qdq_mx(const void* x, void* y, int64_t rows, int64_t cols, int dtype, int block, int format,
              int scale_format, const float* global_amax, int numerics, void* stream);

/* Replaces MXFP8QTensor.quantize / quantize_with_scale / get_weights_scaling_factor (qtensor/mxfp8_tensor.py:41-229).
 * x: [rows, cols]; q: uint8 E4M3 codes [rows, cols]; sf: uint8 E8M0 bytes [rows, ceil(cols / 32)], written unless
 * scale_given != 0 (then read).  q may be NULL with scale_given == 0 to compute only the scales. */
int mo_mxfp8_pack(const void* x, uint8_t* q, uint8_t* sf, int64_t rows, int64_t cols, int dtype, int scale_given,
                  int numerics, void* stream);
/* Replaces MXFP4QTensor.quantize (qtensor/mxfp4_tensor.py:36-83).  x: n_blocks * block elements; q: uint8[n / 2], even
 * element in the LOW nibble, nibble = sign << 3 | #{e2m1 bounds < |x / 2^e|} (a zero element carries the sign bit, ties
 * round down, as written there); sf: uint8[n_blocks] = ceil(log2(amax / 6)) + 127. */
int mo_mxfp4_pack(const void* x, uint8_t* q, uint8_t* sf, int64_t n_blocks, int block, int dtype, int numerics,
                  void* stream);
/* Replaces MXFP8QTensor.dequantize (kind 0, q = E4M3 codes [rows, cols]) and MXFP4QTensor.dequantize (kind 1, q =
 * packed nibbles [rows, cols / 2]) (mxfp8_tensor.py:231-262, mxfp4_tensor.py:85-144): value * 2^(sf - 127) cast to
 * out_dtype; sf: [rows, ceil(cols / block)]. */
int mo_mx_unpack(const uint8_t* q, const uint8_t* sf, void* y, int64_t rows, int64_t cols, int block, int kind,
                 int out_dtype, void* stream);

/* Replaces INT4QTensor.quantize (qtensor/int4_tensor.py:39-88) = reduce_amax + `7 / amax` + cuda_ext.INT4_quantize
 * (gemm/tensor_quant_gpu.cu:313-366) in ONE pass.  x: n_blocks * block elements of `dtype` (already padded);
 * scales (out): n_blocks elements of the SAME dtype = T(T(1 / amax) * 7); q: uint8[n / 2], EVEN element in the HIGH
 * nibble, nibble = code + 8.  numerics: MO_NUMERICS_GPU = the CUDA kernel's roundf(T(clamp(T(x * s), -8, 7) + 8)) in
 * the tensor's own arithmetic; MO_NUMERICS_CPU = the eager round(T(x * s)).clamp(-8, 7) + 8. */
int mo_int4_pack(const void* x, int64_t n_blocks, int block, int dtype, void* scales, uint8_t* q, int numerics,
                 void* stream);
/* Replaces cuda_ext.INT4_dequantize / the eager branch (int4_tensor.py:90-130): y[2i], y[2i+1] = (nibble - 8) / scale
 * in the scales' dtype. */
int mo_int4_unpack(const uint8_t* q, const void* scales, int64_t n_blocks, int block, int dtype, void* y, void* stream);
/* Replaces NF4QTensor.quantize's first stage (qtensor/nf4_tensor.py:75-124, cuda_ext.NF4_quantize
 * tensor_quant_gpu.cu:192-247): scales (out) = per-block amax in `dtype`, q = first-argmin table index of x / amax, even
 * element in the high nibble.  numerics as for mo_int4_pack (kernel: fp32 table literals; eager: bf16 table values). */
int mo_nf4_pack(const void* x, int64_t n_blocks, int block, int dtype, void* scales, uint8_t* q, int numerics,
                void* stream);
/* Replaces cuda_ext.NF4_dequantize (bf16_form = 1: bf16 table * bf16 scale -> bf16, out_dtype must be MO_BF16) and
 * the eager branch (bf16_form = 0: bf16 table value * fp32 scale). scales: fp32 [n_blocks], already double-dequantized. */
int mo_nf4_unpack(const uint8_t* q, const float* scales, int64_t n_blocks, int block, int out_dtype, int bf16_form,
                  void* y, void* stream);

/* Replaces INT8QTensor.quantize (kind 0, qtensor/int8_tensor.py:35-93) and FP8QTensor.quantize (kind 1,
 * qtensor/fp8_tensor.py:38-115) for per-tensor, per-axis and N-D block scales.  x is viewed [O, A, I]; element (o, a, i)
 * uses scales[(o / bo) * so + (a / ba) * sa + (i / bi) * si]; dims = {O, A, I, bo, ba, bi, so, sa, si} (host array).
 * scales have x's dtype.  q = int8(clamp(round(T(x / s)), -128, 127)) or e4m3(T(x / s)); kind 2 = `_to_fp8`
 * (backends/fp8_per_tensor_gemm.py:31-33, a compiled function): scales are FP32, the quotient stays in fp32, is clamped
 * to +-448 (the saturating conversion) and converted to E4M3. */
int mo_q8_pack(const void* x, uint8_t* q, const void* scales, const int64_t* dims, int dtype, int kind, void* stream);
/* Replaces INT8QTensor.dequantize / FP8QTensor.dequantize (int8_tensor.py:95-124, fp8_tensor.py:117-151):
 * y = O(O(q) * O(scale)). */
int mo_q8_unpack(const uint8_t* q, const void* scales, void* y, const int64_t* dims, int scale_dtype, int out_dtype,
                 int kind, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Local-Hessian weight calibration.
 * ---------------------------------------------------------------------------------------------- */
/* Replaces _LocalHessianAccumulator.accumulate (model_calib.py:848-862): hessian[n] += X_n^T X_n for every block n of
 * `block` (= 16) input columns, X = x viewed [tokens, cin] (any float dtype, products in fp32).  hessian: fp32
 * [cin / 16, 16, 16], accumulated in place; scratch: mo_block_hessian_scratch_bytes(tokens, cin) bytes of device memory
 * (partial sums of the token slices, added in a fixed order). */
int64_t mo_block_hessian_scratch_bytes(int64_t tokens, int64_t cin);
int mo_block_hessian(const void* x, int64_t tokens, int64_t cin, int dtype, int block, float* hessian, void* scratch,
                     void* stream);
/* Replaces Triton nvfp4_fp8_scale_sweep_hessian (gemm/nvfp4_fp8_sweep.py:173-290): w is [cout, n_cin_blocks * 16]
 * row-major (block b = row * n_cin_blocks + n), hessian fp32 [n_cin_blocks, 16, 16] (normalised by the sample count by
 * the caller); per block the candidate among the 126 FP8 scales (as quantised by compute_fp4_scales, :263-266) that
 * minimises dw^T H dw, dw = w - quant(w); first minimum; best_amax[b] = candidate_amax[best] (fp32 [cout * n_cin_blocks]). */
int mo_nvfp4_fp8_scale_sweep_hessian(const void* w, int64_t cout, int64_t n_cin_blocks, int block, int dtype,
                                     const float* hessian, const float* global_amax, float* best_amax, int numerics,
                                     void* stream);

/* Replaces Triton gptq_fused_block_scalar (gemm/gptq_fused_kernel.py:43-136): the GPTQ column loop of ONE block of
 * `block` in {32,64,128,256} columns with the static NVFP4 quantizer fused in.  w: fp32 [rows, block] working weights of
 * the block; qw / err (out): fp32 [rows, block] quantized weights and propagated errors; amax: fp32
 * [rows, n_amax_blocks] per-block amax of the whole weight; global_scale = global_amax / (6 * 448); hinv: fp32
 * [block, block] block of the upper Cholesky factor of the inverse Hessian; block_start: column offset of the block. */
int mo_gptq_block_nvfp4(const float* w, float* qw, float* err, const float* amax, float global_scale, const float* hinv,
                        int64_t rows, int block, int n_amax_blocks, int quant_block, int block_start, int numerics,
                        void* stream);

/* Replaces the FP8 per-tensor real-quant GEMM backend (backends/fp8_per_tensor_gemm.py:36-50: torch._scaled_mm with
 * per-tensor scale_a / scale_b): out[M,N] = (a[M,K] . b[N,K]^T) * scale_a * scale_b (+ bias), a / b = E4M3 codes
 * (mo_q8_pack kind 2 = `(x / scale).clamp(-448, 448).to(e4m3)`, :31-33), scale_a / scale_b fp32 device scalars, bias and
 * out in out_dtype (bf16 / f16).  tcgen05.mma.kind::f8f6f4, fp32 accumulation.  K % 16 == 0, N % 8 == 0. */
int mo_linear_fp8(const uint8_t* a, const uint8_t* b, const float* scale_a, const float* scale_b, const void* bias,
                  void* out, int64_t M, int64_t N, int64_t K, int out_dtype, void* stream);
/* mo_linear (plain and loss modes) on CTA pairs: tcgen05 cta_group::2, 512 x 256 (or, for short-K loss searches and
 * small M, 256 x 256) tile per 2-CTA cluster, TMA-store epilogue (csrc/gemm_2cta.cu).  Same arguments, same results bit
 * for bit. */
int mo_linear_pair(const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K, int dtype,
                   const void* ref, double* loss_acc, void* stream);
/* Same contract on CTA pairs (tcgen05 cta_group::2, 512 x 256 or 256 x 256 tile per 2-CTA cluster; csrc/gemm_2cta.cu). */
int mo_linear_fp8_pair(const uint8_t* a, const uint8_t* b, const float* scale_a, const float* scale_b, const void* bias,
                       void* out, int64_t M, int64_t N, int64_t K, int out_dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Device self-tests of the numerics tricks the kernels rely on (run by the GPU test-suite).
 * Each returns the number of mismatching cases in *mismatches (host pointer), synchronously.
 * ---------------------------------------------------------------------------------------------- */
/* cvt.rn.satfinite.e2m1x2.f32 round trip == the reference compare chain, over all 2^31 |fp32|. */
int mo_selftest_e2m1(uint64_t* mismatches);
/* ExactDiv (hoisted reciprocal) == div.rn.f32 for q in [-q_max, q_max] x n_scales random scales. */
int mo_selftest_exact_div(int q_max, int n_scales, uint64_t seed, uint64_t* mismatches);

/* ------------------------------------------------------------------------------------------------
 * HOST-buffer entry points (the end-to-end path bench.py reports as `e2e`).  x/y are host pointers
 * (pinned or pageable); the call stages chunks through library-owned pinned buffers on two streams
 * so H2D, the kernel and D2H overlap, and returns when y is complete.
 * ---------------------------------------------------------------------------------------------- */
int mo_qdq_fp8_host(const void* x, void* y, int64_t n, int dtype, float amax, int numerics);
int mo_qdq_int_host(const void* x, void* y, int64_t n, int dtype, float amax, int num_bits,
                    int is_unsigned, int narrow_range);
int mo_qdq_nvfp4_dynamic_host(const void* x, void* y, int64_t rows, int64_t cols, int dtype,
                              float global_amax, int block, int div_mode);
/* amax of a host tensor (per-tensor); *amax_out receives the fp32 value. */
int mo_amax_host(const void* x, int64_t n, int dtype, float* amax_out);
/* Collect + fake-quant of a BATCH of host tensors (a TensorQuantizer that is calibrating and quantizing in the same
 * forward, tensor_quantizer.py:1181-1200; equal to mo_amax_host then mo_qdq_*_host per tensor): running_amax[i] =
 * max(running_amax[i], amax(x_i)), y_i = fake-quant(x_i) under it.  kind 0 = NVFP4 dynamic (`block`), 1 = FP8 per tensor.
 * Each tensor crosses the bus once per direction (it stays resident on the device between its amax and its fake-quant) and
 * the upload of tensor i+1 overlaps the fake-quant + download of tensor i.  Host pointers should be pinned.
 * xs[i] may also be a DEVICE pointer (a weight that already lives on the GPU: read in place, no copy).  ys[i] may be a
 * device pointer (written in place) or NULL = calibration-only: the fake-quantized tensor stays on the device (in the
 * pipeline's buffer, for the consumer kernel) and only running_amax[i] returns - what max_calibrate
 * (model_calib.py:311-498) needs back.  One pipeline per device ordinal (the current device's). */
int mo_calib_qdq_host_batch(int n_tensors, const void* const* xs, void* const* ys, const int64_t* rows, const int64_t* cols,
                            int dtype, float* running_amax, int kind, int block, int numerics);

#ifdef __cplusplus
}
#endif
#endif /* MODELOPT_B200_H_ */
