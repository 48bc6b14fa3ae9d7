// Amax searches over a collected histogram: percentile, MSE and entropy (KL) — the device versions of
// _compute_amax_percentile / _compute_amax_mse / _compute_amax_entropy (reference
// calib/histogram.py:328-343, :285-325, :210-282), which run as host numpy / scipy loops (entropy is
// O(bins^2) Python: 40 s for a 12k-bin histogram on this box; the mse search launches one fake-quant
// kernel and one .cpu() sync per candidate).
//
// Every kernel returns an INDEX (device int32); the host looks the amax up in its edge / centre array
// exactly like the reference does, so the returned amax is bit-identical whenever the index agrees.
//   percentile : fp64, sequential cumulative sum in numpy's order -> index agrees exactly.
//   mse        : one CTA per candidate, the reference's fp32 per-element arithmetic, fp64 accumulation
//                (the reference's fp32 .mean() order is device-dependent already).
//   entropy    : one CTA per candidate, fp64 throughout, fixed-order tree sums.
#include <math.h>

#include "common.cuh"

namespace mo {

constexpr int kSearchThreads = 256;

__device__ __forceinline__ double block_sum_f64(double v, double* scratch) {
  // fixed-order tree: deterministic for a given block size
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kSearchThreads / 32; ++w) t += scratch[w];
  __syncthreads();
  if (threadIdx.x == 0) scratch[0] = t;
  __syncthreads();
  t = scratch[0];
  __syncthreads();
  return t;
}

// ------------------------------------------------------------------------------------------------
// percentile: total = sum(hist); cdf = cumsum(hist / total); idx = searchsorted(cdf, p / 100)
// ------------------------------------------------------------------------------------------------
__global__ void hist_percentile_kernel(const unsigned long long* __restrict__ counts, int nbins, double q,
                                       int* __restrict__ idx_out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  unsigned long long total = 0;
  for (int i = 0; i < nbins; ++i) total += counts[i];
  const double tot = (double)total;
  double cdf = 0.0;
  int idx = nbins;  // searchsorted returns len(cdf) when every entry is < q
  for (int i = 0; i < nbins; ++i) {
    cdf = cdf + (double)counts[i] / tot;  // numpy cumsum: sequential fp64 adds
    if (cdf >= q) { idx = i; break; }     // side='left': first i with cdf[i] >= q  (NaN: never)
  }
  *idx_out = idx;
}

// ------------------------------------------------------------------------------------------------
// mse.  The reference evaluates, per candidate i (amax = centre[i]):
//     mean_j( (fq(centre_j) - centre_j)^2 * count_j )
// where fq is _tensor_quant (tensor_quant.py:607-645) reached with MIS-SLOTTED arguments
// (histogram.py:308: bias <- num_bits, num_bits <- unsigned): as_written = 1 reproduces that,
// as_written = 0 evaluates the intended fake quantisation (INT-N, or FP8 E4M3 when fp8 = 1).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float legacy_tensor_quant(float x, float amax, int num_bits, bool is_unsigned,
                                                     bool narrow) {
  const float maxb = ldexpf(1.0f, num_bits - 1 + (is_unsigned ? 1 : 0)) - 1.0f;
  const float minb = is_unsigned ? 0.0f : (narrow ? -maxb : -maxb - 1.0f);
  const float scale = __fdiv_rn(maxb, amax);
  float v = rintf(__fmul_rn(x, scale));
  // torch.clamp(v, min, max) == min(max(v, min), max) with NaN propagated: when min > max the result is max
  if (!isnan(v)) v = fminf(fmaxf(v, minb), maxb);
  return __fdiv_rn(v, scale);
}

__global__ void __launch_bounds__(kSearchThreads) hist_mse_kernel(const unsigned long long* __restrict__ counts,
                                                                  const float* __restrict__ edges, int nbins,
                                                                  int num_bits, int is_unsigned, int fp8,
                                                                  int as_written, int stride, int start_bin,
                                                                  double* __restrict__ loss) {
  __shared__ double scratch[kSearchThreads / 32];
  const int cand = blockIdx.x;
  const int i = start_bin + cand * stride;
  const float amax = __fmul_rn(__fadd_rn(edges[i + 1], edges[i]), 0.5f);  // centres = (e[1:] + e[:-1]) / 2
  double acc = 0.0;
  float sc = 1.f, inv = 1.f;
  if (fp8) {  // _fp8_eager on the host device of the search (CPU numerics: reciprocal * 448)
    const float safe = amax <= kEps ? 1.0f : amax;
    sc = __fmul_rn(__fdiv_rn(1.0f, safe), 448.0f);
    inv = __fdiv_rn(1.0f, sc);
  }
  for (int j = threadIdx.x; j < nbins; j += kSearchThreads) {
    const float c = __fmul_rn(__fadd_rn(edges[j + 1], edges[j]), 0.5f);
    float q;
    if (as_written) {
      const float bias = (float)num_bits;
      q = __fadd_rn(legacy_tensor_quant(__fsub_rn(c, bias), amax, is_unsigned ? 1 : 0, false, true), bias);
    } else if (fp8) {
      q = __fmul_rn(e4m3_roundtrip1(__fmul_rn(c, sc)), inv);
    } else {
      q = legacy_tensor_quant(c, amax, num_bits, is_unsigned != 0, true);
    }
    const float d = __fsub_rn(q, c);
    const float e = __fmul_rn(__fmul_rn(d, d), (float)counts[j]);  // (q - c) ** 2 * counts, fp32 like the reference
    acc += (double)e;
  }
  const double tot = block_sum_f64(acc, scratch);
  if (threadIdx.x == 0) loss[cand] = (double)(float)(tot / (double)nbins);  // .mean() result is fp32
}

// first minimum; any NaN -> first NaN (numpy.argmin).  last = 1: LAST minimum (entropy's reversed argmin;
// with NaN present numpy picks the first NaN of the REVERSED array = the last NaN).
__global__ void __launch_bounds__(kSearchThreads) argmin_kernel(const double* __restrict__ v, int n, int last,
                                                                int* __restrict__ idx_out) {
  __shared__ double sv[kSearchThreads];
  __shared__ int si[kSearchThreads];
  double best = 0.0;
  int bi = -1;
  bool nan_seen = false;
  auto better = [&](double a, int ia, bool an, double b, int ib, bool bn) {
    // is (a, ia) preferred over (b, ib)?
    if (ib < 0) return true;
    if (ia < 0) return false;
    if (an != bn) return an;  // NaN wins
    if (an) return last ? ia > ib : ia < ib;
    if (a != b) return a < b;
    return last ? ia > ib : ia < ib;
  };
  for (int i = threadIdx.x; i < n; i += kSearchThreads) {
    const double x = v[i];
    const bool xn = isnan(x);
    if (better(x, i, xn, best, bi, nan_seen)) { best = x; bi = i; nan_seen = xn; }
  }
  sv[threadIdx.x] = best;
  si[threadIdx.x] = bi;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int t = 1; t < kSearchThreads; ++t) {
      const bool tn = si[t] >= 0 && isnan(sv[t]);
      if (better(sv[t], si[t], tn, best, bi, nan_seen)) { best = sv[t]; bi = si[t]; nan_seen = tn; }
    }
    *idx_out = bi;
  }
}

// ------------------------------------------------------------------------------------------------
// entropy (KL) search.  Candidate i keeps the first i bins: they are merged into G = 2^(bits-1+unsigned)
// equal-width groups (numpy: digitize(range(i), linspace(0, i, G + 1)) - 1); each NON-EMPTY bin of a
// group takes the group's mean over its non-empty bins (q); the reference distribution p is the first i
// bins with the clipped tail added to bin i-1; divergence = scipy.stats.entropy(p, q).
// `h` is the histogram AFTER the reference's hist[0] = hist[1] (done by the launcher kernel below).
// ------------------------------------------------------------------------------------------------
constexpr int kMaxGroups = 2048;  // num_bits + unsigned <= 12; static shared memory stays under 48 KB

__global__ void __launch_bounds__(kSearchThreads) hist_entropy_kernel(const unsigned long long* __restrict__ h,
                                                                      int nbins, int G, int stride, int start_bin,
                                                                      const unsigned long long* __restrict__ suffix,
                                                                      double* __restrict__ kl) {
  __shared__ double scratch[kSearchThreads / 32];
  __shared__ double mean[kMaxGroups];
  __shared__ int jstart[kMaxGroups + 1];
  const int cand = blockIdx.x;
  const int i = start_bin + cand * stride;
  const double step = (double)i / (double)G;
  // group g holds the bins j with space[g] <= j < space[g + 1], space = numpy.linspace(0, i, G + 1)
  // (space[k] = k * step in fp64, space[G] = i): the first such j is ceil(space[g]), exact in fp64
  for (int g = threadIdx.x; g <= G; g += kSearchThreads)
    jstart[g] = g == G ? i : min(i, (int)ceil((double)g * step));
  __syncthreads();
  // pass 1: group means over the non-empty bins; sum(q) accumulated group by group
  double qsum_part = 0.0;
  for (int g = threadIdx.x; g < G; g += kSearchThreads) {
    double s = 0.0;
    int cnt = 0;
    for (int j = jstart[g]; j < jstart[g + 1]; ++j) {
      const unsigned long long c = h[j];
      if (c) { s += (double)c; ++cnt; }  // integer-valued partial sums: exact in fp64
    }
    const double m = cnt ? s / (double)cnt : 0.0;
    mean[g] = m;
    qsum_part += m * (double)cnt;
  }
  const double qsum = block_sum_f64(qsum_part, scratch);
  const unsigned long long tail = suffix[i];  // sum(h[i:])
  const double psum = (double)suffix[0];       // sum(p) = every count (exact)
  // pass 2: sum_j rel_entr(p_j / psum, q_j / qsum)
  double acc = 0.0;
  for (int j = threadIdx.x; j < i; j += kSearchThreads) {
    int g = min(G - 1, (int)((double)j / step));
    while (jstart[g] > j) --g;
    while (jstart[g + 1] <= j) ++g;
    unsigned long long pc = h[j];
    const double qv = pc ? mean[g] : 0.0;
    if (j == i - 1) pc += tail;
    const double pk = (double)pc / psum;
    const double qk = qv / qsum;
    double t;
    if (pk > 0.0 && qk > 0.0) t = pk * log(pk / qk);
    else if (pk == 0.0 && qk >= 0.0) t = 0.0;
    else t = INFINITY;
    acc += t;
  }
  const double tot = block_sum_f64(acc, scratch);
  if (threadIdx.x == 0) kl[cand] = tot;
}

// h = counts with h[0] = h[1]; suffix[i] = sum(h[i:]) (suffix[nbins] = 0)
__global__ void hist_entropy_prep_kernel(const unsigned long long* __restrict__ counts, int nbins,
                                         unsigned long long* __restrict__ h, unsigned long long* __restrict__ suffix) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int i = 0; i < nbins; ++i) h[i] = counts[i];
  if (nbins > 1) h[0] = h[1];
  unsigned long long s = 0;
  suffix[nbins] = 0;
  for (int i = nbins - 1; i >= 0; --i) { s += h[i]; suffix[i] = s; }
}

}  // namespace mo

using namespace mo;

extern "C" int mo_hist_percentile(const uint64_t* counts, int64_t nbins, double percentile, int32_t* idx_out,
                                  void* stream) {
  if (!counts || !idx_out) { set_error("mo_hist_percentile: NULL pointer"); return MO_ERR_NULL; }
  if (nbins < 1 || !(percentile >= 0.0 && percentile <= 100.0)) {
    set_error("mo_hist_percentile: percentile must be in [0, 100], nbins >= 1");
    return MO_ERR_ARG;
  }
  hist_percentile_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(counts), (int)nbins, percentile / 100, idx_out);
  return check_launch("mo_hist_percentile");
}

extern "C" int mo_hist_mse(const uint64_t* counts, const float* edges, int64_t nbins, int num_bits, int is_unsigned,
                           int fp8, int as_written, int stride, int start_bin, double* loss_scratch,
                           int32_t* idx_out, void* stream) {
  if (!counts || !edges || !loss_scratch || !idx_out) { set_error("mo_hist_mse: NULL pointer"); return MO_ERR_NULL; }
  if (stride < 1 || start_bin < 0 || start_bin >= nbins) { set_error("mo_hist_mse: bad stride/start_bin"); return MO_ERR_ARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int ncand = (int)((nbins - start_bin + stride - 1) / stride);  // range(start_bin, len(centers), stride)
  hist_mse_kernel<<<ncand, kSearchThreads, 0, s>>>(reinterpret_cast<const unsigned long long*>(counts), edges,
                                                   (int)nbins, num_bits, is_unsigned, fp8, as_written, stride,
                                                   start_bin, loss_scratch);
  argmin_kernel<<<1, kSearchThreads, 0, s>>>(loss_scratch, ncand, 0, idx_out);
  return check_launch("mo_hist_mse");
}

extern "C" int64_t mo_hist_entropy_scratch_bytes(int64_t nbins, int stride, int start_bin) {
  if (nbins < 1 || stride < 1 || start_bin < 0 || start_bin > nbins) return 0;
  const int64_t ncand = (nbins - start_bin) / stride + 1;  // range(start_bin, nbins + 1, stride)
  return (2 * nbins + 1) * 8 + ncand * 8;
}

extern "C" int mo_hist_entropy(const uint64_t* counts, int64_t nbins, int num_bits, int is_unsigned, int stride,
                               int start_bin, void* scratch, int32_t* idx_out, void* stream) {
  if (!counts || !scratch || !idx_out) { set_error("mo_hist_entropy: NULL pointer"); return MO_ERR_NULL; }
  if (stride < 1 || start_bin < 1 || start_bin > nbins || num_bits < 1 || num_bits + (is_unsigned ? 1 : 0) > 16) {
    set_error("mo_hist_entropy: bad stride/start_bin/num_bits");
    return MO_ERR_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int ncand = (int)((nbins - start_bin) / stride + 1);
  unsigned long long* h = static_cast<unsigned long long*>(scratch);
  unsigned long long* suffix = h + nbins;
  double* kl = reinterpret_cast<double*>(suffix + nbins + 1);
  const int G = 1 << (num_bits - 1 + (is_unsigned ? 1 : 0));
  if (G > kMaxGroups) { set_error("mo_hist_entropy: 2^(num_bits-1+unsigned) = %d groups exceeds %d", G, kMaxGroups); return MO_ERR_ARG; }
  hist_entropy_prep_kernel<<<1, 32, 0, s>>>(reinterpret_cast<const unsigned long long*>(counts), (int)nbins, h, suffix);
  hist_entropy_kernel<<<ncand, kSearchThreads, 0, s>>>(h, (int)nbins, G, stride, start_bin, suffix, kl);
  argmin_kernel<<<1, kSearchThreads, 0, s>>>(kl, ncand, 1, idx_out);
  return check_launch("mo_hist_entropy");
}

// ---------------------------------------------------------------------------------------------------------------------
// mo_hist_entropy_static: the same search for callers that cannot allocate (CUDA graph capture, serving loops that
// forbid cudaMalloc).  No counterpart in the reference.  The scratch
// (h, suffix sums, one KL word per candidate) lives in a file-scope device array of fixed size, so the call performs
// no cudaMalloc / cudaFree and calls no allocating library routine.  Every GPU that loads the library holds the
// array (kStaticEntropyBytes = 1.5 MiB) until the process exits.
//
// In host code `g_entropy_scratch` is only a host-side stand-in for the symbol, not a device address: its device
// address is looked up once per device (with that device current) by cudaGetSymbolAddress and cached.
//
// All calls on a device share the array and it keeps the last call's contents.  Nothing in it has to start at zero:
// the prep kernel rewrites h[0..nbins) and suffix[0..nbins], and every candidate's KL word is written before the
// argmin reads it.  Calls that overlap on the GPU (different streams, or graph branches) race on it: serialise them.
// ---------------------------------------------------------------------------------------------------------------------
namespace mo {
constexpr int64_t kStaticEntropyMaxBins = 65536;
constexpr int64_t kStaticEntropyBytes = (3 * kStaticEntropyMaxBins + 2) * 8;  // h + suffix(+1) + one double per candidate
static __device__ __align__(256) unsigned char g_entropy_scratch[kStaticEntropyBytes];

constexpr int kMaxDevices = 64;
static void* g_entropy_scratch_ptr[kMaxDevices];  // device address of the array, per device ordinal (0 = not looked up)
static std::mutex g_entropy_scratch_mu;

// device address of g_entropy_scratch on the CURRENT device
static int static_entropy_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("mo_hist_entropy_static: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
  if (dev < 0 || dev >= kMaxDevices) { set_error("mo_hist_entropy_static: device ordinal %d out of range", dev); return MO_ERR_ARG; }
  std::lock_guard<std::mutex> lock(g_entropy_scratch_mu);
  if (!g_entropy_scratch_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_entropy_scratch);
    if (e != cudaSuccess || !p) {
      set_error("mo_hist_entropy_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : MO_ERR_ARG;
    }
    g_entropy_scratch_ptr[dev] = p;
  }
  *out = g_entropy_scratch_ptr[dev];
  return MO_OK;
}
}  // namespace mo

extern "C" int64_t mo_hist_entropy_static_max_bins(void) { return kStaticEntropyMaxBins; }

extern "C" int mo_hist_entropy_static(const uint64_t* counts, int64_t nbins, int num_bits, int is_unsigned, int stride,
                                      int start_bin, int32_t* idx_out, void* stream) {
  if (!counts || !idx_out) { set_error("mo_hist_entropy_static: NULL pointer"); return MO_ERR_NULL; }
  if (nbins < 1 || nbins > kStaticEntropyMaxBins ||
      mo_hist_entropy_scratch_bytes(nbins, stride < 1 ? 1 : stride, start_bin < 0 ? 0 : start_bin) > kStaticEntropyBytes) {
    set_error("mo_hist_entropy_static: nbins = %lld exceeds the static scratch (at most %lld bins); use mo_hist_entropy "
              "with a caller-provided scratch buffer", (long long)nbins, (long long)kStaticEntropyMaxBins);
    return MO_ERR_ARG;
  }
  void* scratch = nullptr;
  const int rc = static_entropy_scratch(&scratch);
  if (rc != MO_OK) return rc;
  return mo_hist_entropy(counts, nbins, num_bits, is_unsigned, stride, start_bin, scratch, idx_out, stream);
}
