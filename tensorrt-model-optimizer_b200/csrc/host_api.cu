// HOST-buffer entry points: the end-to-end form of the hot path (bench.py's `e2e`).
//
// A call splits the tensor into chunks and runs H2D copy -> kernel -> D2H copy for each chunk on a
// ring of kSlots streams, so the copy engines (both directions) and the SMs overlap.  Registered
// (pinned) user memory is copied directly; pageable memory is staged through library-owned pinned
// buffers.  Device and pinned buffers are allocated once per process and reused.
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace mo {

constexpr int kSlots = 3;
constexpr size_t kChunkBytes = 32u << 20;  // per-direction chunk

struct HostPipe {
  bool ready = false;
  cudaStream_t stream[kSlots];
  void* d_in[kSlots];
  void* d_out[kSlots];
  void* p_in[kSlots];   // pinned staging (pageable user memory only)
  void* p_out[kSlots];
  float* d_scalar = nullptr;  // amax / global_amax parameter, one per slot + result word
  std::mutex mu;

  int init() {
    if (ready) return MO_OK;
    for (int i = 0; i < kSlots; ++i) {
      cudaError_t e = cudaStreamCreateWithFlags(&stream[i], cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaMalloc(&d_in[i], kChunkBytes);
      if (e == cudaSuccess) e = cudaMalloc(&d_out[i], kChunkBytes);
      if (e == cudaSuccess) e = cudaMallocHost(&p_in[i], kChunkBytes);
      if (e == cudaSuccess) e = cudaMallocHost(&p_out[i], kChunkBytes);
      if (e != cudaSuccess) { set_error("host pipeline init: %s", cudaGetErrorString(e)); return (int)e; }
    }
    cudaError_t e = cudaMalloc(&d_scalar, 64);
    if (e != cudaSuccess) { set_error("host pipeline init: %s", cudaGetErrorString(e)); return (int)e; }
    ready = true;
    return MO_OK;
  }
};

static HostPipe g_pipe;

static bool is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

static int esize(int dtype) { return dtype == MO_F32 ? 4 : (dtype == MO_F16 || dtype == MO_BF16) ? 2 : 0; }

// Generic chunk loop.  `launch(d_in, d_out, elems, stream)` enqueues the kernel for one chunk;
// has_output == false skips the D2H leg (reductions).
template <typename Launch>
static int run_chunks(const void* x, void* y, int64_t n, int64_t chunk_elems, int es, bool has_output, Launch launch) {
  HostPipe& P = g_pipe;
  const bool pin_x = is_pinned(x), pin_y = has_output && is_pinned(y);
  const char* xb = static_cast<const char*>(x);
  char* yb = static_cast<char*>(y);
  int64_t pending_off[kSlots], pending_n[kSlots];
  for (int i = 0; i < kSlots; ++i) pending_n[i] = 0;
  int64_t off = 0;
  int rc = MO_OK;
  for (int64_t c = 0; off < n; ++c) {
    const int s = (int)(c % kSlots);
    const int64_t m = (n - off) < chunk_elems ? (n - off) : chunk_elems;
    cudaStreamSynchronize(P.stream[s]);  // slot free (its previous chunk is complete)
    if (pending_n[s] && has_output && !pin_y) memcpy(yb + pending_off[s] * es, P.p_out[s], pending_n[s] * es);
    pending_n[s] = 0;
    const void* src = xb + off * es;
    if (!pin_x) { memcpy(P.p_in[s], src, m * es); src = P.p_in[s]; }
    cudaMemcpyAsync(P.d_in[s], src, m * es, cudaMemcpyHostToDevice, P.stream[s]);
    rc = launch(P.d_in[s], P.d_out[s], m, P.stream[s]);
    if (rc) break;
    if (has_output) {
      void* dst = pin_y ? (void*)(yb + off * es) : P.p_out[s];
      cudaMemcpyAsync(dst, P.d_out[s], m * es, cudaMemcpyDeviceToHost, P.stream[s]);
      pending_off[s] = off;
      pending_n[s] = m;
    }
    off += m;
  }
  for (int s = 0; s < kSlots; ++s) {
    cudaStreamSynchronize(P.stream[s]);
    if (pending_n[s] && has_output && !pin_y) memcpy(yb + pending_off[s] * es, P.p_out[s], pending_n[s] * es);
  }
  if (rc) return rc;
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("host pipeline: %s", cudaGetErrorString(e)); return (int)e; }
  return MO_OK;
}

static int put_scalar(float v, int slot) {
  // synchronous 4-byte H2D before the pipeline starts
  const cudaError_t e = cudaMemcpy(g_pipe.d_scalar + slot, &v, 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { set_error("scalar upload: %s", cudaGetErrorString(e)); return (int)e; }
  return MO_OK;
}

}  // namespace mo

using namespace mo;

#define MO_HOST_PROLOGUE(name)                                                         \
  const int es = esize(dtype);                                                         \
  if (!es) { set_error(name ": unsupported dtype %d", dtype); return MO_ERR_DTYPE; }   \
  if (!x) { set_error(name ": NULL pointer"); return MO_ERR_NULL; }                    \
  std::lock_guard<std::mutex> lock(g_pipe.mu);                                         \
  int rc = g_pipe.init();                                                              \
  if (rc) return rc;

extern "C" int mo_qdq_fp8_host(const void* x, void* y, int64_t n, int dtype, float amax, int numerics) {
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_qdq_fp8_host")
  if (!y) { set_error("mo_qdq_fp8_host: NULL pointer"); return MO_ERR_NULL; }
  if ((rc = put_scalar(amax, 0))) return rc;
  const float* d_amax = g_pipe.d_scalar;
  return run_chunks(x, y, n, (int64_t)(kChunkBytes / es), es, true,
                    [&](void* di, void* dout, int64_t m, cudaStream_t s) {
                      return mo_qdq_fp8(di, dout, m, dtype, d_amax, 1, m, numerics, s);
                    });
}

extern "C" int mo_qdq_int_host(const void* x, void* y, int64_t n, int dtype, float amax, int num_bits, int is_unsigned,
                               int narrow_range) {
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_qdq_int_host")
  if (!y) { set_error("mo_qdq_int_host: NULL pointer"); return MO_ERR_NULL; }
  if ((rc = put_scalar(amax, 0))) return rc;
  const float* d_amax = g_pipe.d_scalar;
  return run_chunks(x, y, n, (int64_t)(kChunkBytes / es), es, true,
                    [&](void* di, void* dout, int64_t m, cudaStream_t s) {
                      return mo_qdq_int(di, dout, m, dtype, d_amax, 1, m, num_bits, is_unsigned, narrow_range,
                                        nullptr, s);
                    });
}

extern "C" int mo_qdq_nvfp4_dynamic_host(const void* x, void* y, int64_t rows, int64_t cols, int dtype,
                                         float global_amax, int block, int div_mode) {
  if (rows == 0 || cols == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_qdq_nvfp4_dynamic_host")
  if (!y) { set_error("mo_qdq_nvfp4_dynamic_host: NULL pointer"); return MO_ERR_NULL; }
  if ((int64_t)cols * es > (int64_t)kChunkBytes) {
    set_error("mo_qdq_nvfp4_dynamic_host: a row (%lld bytes) exceeds the %zu-byte staging chunk",
              (long long)cols * es, kChunkBytes);
    return MO_ERR_SHAPE;
  }
  if ((rc = put_scalar(global_amax, 0))) return rc;
  const float* d_amax = g_pipe.d_scalar;
  const int64_t rows_per_chunk = (int64_t)(kChunkBytes / es) / cols;
  return run_chunks(x, y, rows * cols, rows_per_chunk * cols, es, true,
                    [&](void* di, void* dout, int64_t m, cudaStream_t s) {
                      return mo_qdq_nvfp4_dynamic(di, dout, m / cols, cols, dtype, d_amax, block, div_mode, s);
                    });
}

extern "C" int mo_amax_host(const void* x, int64_t n, int dtype, float* amax_out) {
  if (!amax_out) { set_error("mo_amax_host: NULL pointer"); return MO_ERR_NULL; }
  *amax_out = 0.0f;
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_amax_host")
  uint32_t* d_bits = reinterpret_cast<uint32_t*>(g_pipe.d_scalar + 8);
  { const uint32_t z = 0; cudaMemcpy(d_bits, &z, 4, cudaMemcpyHostToDevice); }
  rc = run_chunks(x, nullptr, n, (int64_t)(kChunkBytes / es), es, false,
                  [&](void* di, void*, int64_t m, cudaStream_t s) { return mo_amax(di, 1, 1, m, dtype, d_bits, s); });
  if (rc) return rc;
  const cudaError_t e = cudaMemcpy(amax_out, d_bits, 4, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { set_error("mo_amax_host: %s", cudaGetErrorString(e)); return (int)e; }
  return MO_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Batched "collect + fake-quant" of HOST tensors: for each tensor i, running_amax[i] = max(running_amax[i], amax(x_i)) and
// y_i = NVFP4 / FP8 fake-quant of x_i under that amax - what TensorQuantizer.forward does for a quantizer that is
// collecting and quantizing in the same pass (tensor_quantizer.py:1181-1200), and what mo_amax_host followed by
// mo_qdq_*_host computes, but every tensor crosses the bus ONCE in each direction: the tensor is kept resident on the
// device between its amax and its fake-quant (two ping-pong buffers), the upload of tensor i+1 runs on one stream while
// the fake-quant and download of tensor i run on another, so both copy engines stay busy.
// ---------------------------------------------------------------------------------------------------------------------
namespace mo {

struct BatchPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ready[2], freed[2];
  void* res[2] = {nullptr, nullptr};   // resident inputs
  void* outb[2] = {nullptr, nullptr};  // fake-quantized outputs awaiting download
  size_t cap = 0;
  uint32_t* d_bits = nullptr;  // per-tensor amax bit patterns
  float* d_amax = nullptr;     // per-tensor running amax (in: uploaded, out: merged)
  int n_cap = 0;
  int ensure(size_t bytes, int n) {
    cudaError_t e = cudaSuccess;
    if (!s_in) {
      e = cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking);
      for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&ready[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&freed[i], cudaEventDisableTiming);
      }
    }
    if (e == cudaSuccess && bytes > cap) {
      for (int i = 0; i < 2; ++i) { cudaFree(res[i]); cudaFree(outb[i]); res[i] = outb[i] = nullptr; }
      for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaMalloc(&res[i], bytes);
        if (e == cudaSuccess) e = cudaMalloc(&outb[i], bytes);
      }
      cap = e == cudaSuccess ? bytes : 0;
    }
    if (e == cudaSuccess && n > n_cap) {
      cudaFree(d_bits); cudaFree(d_amax);
      e = cudaMalloc(&d_bits, (size_t)n * 4);
      if (e == cudaSuccess) e = cudaMalloc(&d_amax, (size_t)n * 4);
      n_cap = e == cudaSuccess ? n : 0;
    }
    if (e != cudaSuccess) { set_error("batch host pipeline: %s", cudaGetErrorString(e)); return (int)e; }
    return MO_OK;
  }
};
static BatchPipe g_batch;

// running = max(running, this tensor's amax); amax bit patterns of non-negative floats order like the floats
__global__ void merge_running_amax_kernel(const uint32_t* bits, float* running) {
  const uint32_t r = __float_as_uint(*running);
  *running = __uint_as_float(max(r, *bits));
}

}  // namespace mo

// kind: 0 = NVFP4 dynamic (block = 16 / 32 / 64), 1 = FP8 E4M3 per tensor
extern "C" int mo_calib_qdq_host_batch(int n_tensors, const void* const* xs, void* const* ys, const int64_t* rows,
                                       const int64_t* cols, int dtype, float* running_amax, int kind, int block,
                                       int numerics) {
  if (n_tensors == 0) return MO_OK;
  const int es = esize(dtype);
  if (!es) { set_error("mo_calib_qdq_host_batch: unsupported dtype %d", dtype); return MO_ERR_DTYPE; }
  if (!xs || !ys || !rows || !cols || !running_amax || n_tensors < 0) { set_error("mo_calib_qdq_host_batch: NULL pointer"); return MO_ERR_NULL; }
  if (kind != 0 && kind != 1) { set_error("mo_calib_qdq_host_batch: kind must be 0 (nvfp4) or 1 (fp8)"); return MO_ERR_ARG; }
  size_t max_bytes = 0;
  for (int i = 0; i < n_tensors; ++i) {
    if (rows[i] < 0 || cols[i] < 0 || (rows[i] * cols[i] > 0 && (!xs[i] || !ys[i]))) { set_error("mo_calib_qdq_host_batch: bad tensor %d", i); return MO_ERR_ARG; }
    const size_t b = (size_t)rows[i] * (size_t)cols[i] * es;
    if (b > max_bytes) max_bytes = b;
  }
  std::lock_guard<std::mutex> lock(g_pipe.mu);
  BatchPipe& P = g_batch;
  int rc = P.ensure(max_bytes ? max_bytes : 32, n_tensors);
  if (rc) return rc;
  cudaMemcpyAsync(P.d_amax, running_amax, (size_t)n_tensors * 4, cudaMemcpyHostToDevice, P.s_in);
  cudaMemsetAsync(P.d_bits, 0, (size_t)n_tensors * 4, P.s_in);
  for (int i = 0; i < n_tensors && rc == MO_OK; ++i) {
    const int b = i & 1;
    const int64_t n = rows[i] * cols[i];
    if (n == 0) continue;
    // ---- upload + amax on s_in -------------------------------------------------------------------------------
    if (i >= 2) cudaStreamWaitEvent(P.s_in, P.freed[b], 0);  // the fake-quant of tensor i-2 has read res[b]
    cudaMemcpyAsync(P.res[b], xs[i], (size_t)n * es, cudaMemcpyHostToDevice, P.s_in);
    rc = mo_amax(P.res[b], 1, 1, n, dtype, P.d_bits + i, P.s_in);
    if (rc) break;
    merge_running_amax_kernel<<<1, 1, 0, P.s_in>>>(P.d_bits + i, P.d_amax + i);
    cudaEventRecord(P.ready[b], P.s_in);
    // ---- fake-quant + download on s_out (stream order also protects outb[b] from tensor i-2's download) ------
    cudaStreamWaitEvent(P.s_out, P.ready[b], 0);
    if (kind == 0) rc = mo_qdq_nvfp4_dynamic(P.res[b], P.outb[b], rows[i], cols[i], dtype, P.d_amax + i, block, numerics, P.s_out);
    else rc = mo_qdq_fp8(P.res[b], P.outb[b], n, dtype, P.d_amax + i, 1, n, numerics, P.s_out);
    if (rc) break;
    cudaEventRecord(P.freed[b], P.s_out);
    cudaMemcpyAsync(ys[i], P.outb[b], (size_t)n * es, cudaMemcpyDeviceToHost, P.s_out);
  }
  cudaStreamSynchronize(P.s_in);
  cudaStreamSynchronize(P.s_out);
  if (rc) return rc;
  cudaError_t e = cudaMemcpy(running_amax, P.d_amax, (size_t)n_tensors * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("mo_calib_qdq_host_batch: %s", cudaGetErrorString(e)); return (int)e; }
  return MO_OK;
}
