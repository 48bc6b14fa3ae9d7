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

// one pipeline per device ordinal: staging buffers and streams belong to the device that was current when they were
// created, so a process that drives several GPUs gets one set per GPU (created on first use with that GPU current)
constexpr int kMaxPipeDevices = 64;
static HostPipe g_pipes[kMaxPipeDevices];
static std::mutex g_pipe_mu;  // serialises host-buffer calls (they share the staging buffers)

static HostPipe* current_pipe() {
  int dev = 0;
  const cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxPipeDevices) {
    set_error("host pipeline: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  return &g_pipes[dev];
}

static bool is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

static int esize(int dtype) { return dtype == MO_F32 ? 4 : (dtype == MO_F16 || dtype == MO_BF16) ? 2 : 0; }

// Generic chunk loop.  `launch(d_in, d_out, elems, stream)` enqueues the kernel for one chunk;
// has_output == false skips the D2H leg (reductions).
template <typename Launch>
static int run_chunks(HostPipe& P, const void* x, void* y, int64_t n, int64_t chunk_elems, int es, bool has_output,
                      Launch launch) {
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
    cudaError_t ce = cudaStreamSynchronize(P.stream[s]);  // slot free (its previous chunk is complete)
    if (ce != cudaSuccess) { set_error("host pipeline: %s", cudaGetErrorString(ce)); rc = (int)ce; break; }
    if (pending_n[s] && has_output && !pin_y) memcpy(yb + pending_off[s] * es, P.p_out[s], pending_n[s] * es);
    pending_n[s] = 0;
    const void* src = xb + off * es;
    if (!pin_x) { memcpy(P.p_in[s], src, m * es); src = P.p_in[s]; }
    ce = cudaMemcpyAsync(P.d_in[s], src, m * es, cudaMemcpyHostToDevice, P.stream[s]);
    if (ce != cudaSuccess) { set_error("host pipeline: H2D copy: %s", cudaGetErrorString(ce)); rc = (int)ce; break; }
    rc = launch(P.d_in[s], P.d_out[s], m, P.stream[s]);
    if (rc) break;
    if (has_output) {
      void* dst = pin_y ? (void*)(yb + off * es) : P.p_out[s];
      ce = cudaMemcpyAsync(dst, P.d_out[s], m * es, cudaMemcpyDeviceToHost, P.stream[s]);
      if (ce != cudaSuccess) { set_error("host pipeline: D2H copy: %s", cudaGetErrorString(ce)); rc = (int)ce; break; }
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

// Parameter word(s) for the chunk kernels.  The pipeline's streams are non-blocking (they do not order against the
// legacy stream), so the upload is issued ON stream[0] and the other slots wait for it through an event: a kernel can
// never read a stale amax or an un-zeroed accumulator word.
static int put_words(HostPipe& P, const void* host_words, int first_word, int n_words) {
  cudaError_t e = cudaMemcpyAsync(P.d_scalar + first_word, host_words, (size_t)n_words * 4, cudaMemcpyHostToDevice, P.stream[0]);
  if (e == cudaSuccess) e = cudaStreamSynchronize(P.stream[0]);  // the source is a stack variable: complete before returning
  if (e != cudaSuccess) { set_error("scalar upload: %s", cudaGetErrorString(e)); return (int)e; }
  return MO_OK;
}
static int put_scalar(HostPipe& P, float v, int slot) { return put_words(P, &v, slot, 1); }

}  // namespace mo

using namespace mo;

#define MO_HOST_PROLOGUE(name)                                                         \
  const int es = esize(dtype);                                                         \
  if (!es) { set_error(name ": unsupported dtype %d", dtype); return MO_ERR_DTYPE; }   \
  if (!x) { set_error(name ": NULL pointer"); return MO_ERR_NULL; }                    \
  std::lock_guard<std::mutex> lock(g_pipe_mu);                                         \
  HostPipe* pipe_ = current_pipe();                                                    \
  if (!pipe_) return MO_ERR_ARG;                                                       \
  HostPipe& P = *pipe_;                                                                \
  int rc = P.init();                                                                   \
  if (rc) return rc;

extern "C" int mo_qdq_fp8_host(const void* x, void* y, int64_t n, int dtype, float amax, int numerics) {
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_qdq_fp8_host")
  if (!y) { set_error("mo_qdq_fp8_host: NULL pointer"); return MO_ERR_NULL; }
  if ((rc = put_scalar(P, amax, 0))) return rc;
  const float* d_amax = P.d_scalar;
  return run_chunks(P, x, y, n, (int64_t)(kChunkBytes / es), es, true,
                    [&](void* di, void* dout, int64_t m, cudaStream_t s) {
                      return mo_qdq_fp8(di, dout, m, dtype, d_amax, 1, m, numerics, s);
                    });
}

extern "C" int mo_qdq_int_host(const void* x, void* y, int64_t n, int dtype, float amax, int num_bits, int is_unsigned,
                               int narrow_range) {
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_qdq_int_host")
  if (!y) { set_error("mo_qdq_int_host: NULL pointer"); return MO_ERR_NULL; }
  if ((rc = put_scalar(P, amax, 0))) return rc;
  const float* d_amax = P.d_scalar;
  return run_chunks(P, x, y, n, (int64_t)(kChunkBytes / es), es, true,
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
  if ((rc = put_scalar(P, global_amax, 0))) return rc;
  const float* d_amax = P.d_scalar;
  const int64_t rows_per_chunk = (int64_t)(kChunkBytes / es) / cols;
  return run_chunks(P, x, y, rows * cols, rows_per_chunk * cols, es, true,
                    [&](void* di, void* dout, int64_t m, cudaStream_t s) {
                      return mo_qdq_nvfp4_dynamic(di, dout, m / cols, cols, dtype, d_amax, block, div_mode, s);
                    });
}

extern "C" int mo_amax_host(const void* x, int64_t n, int dtype, float* amax_out) {
  if (!amax_out) { set_error("mo_amax_host: NULL pointer"); return MO_ERR_NULL; }
  *amax_out = 0.0f;
  if (n == 0) return MO_OK;
  MO_HOST_PROLOGUE("mo_amax_host")
  uint32_t* d_bits = reinterpret_cast<uint32_t*>(P.d_scalar + 8);
  { const uint32_t z = 0; if ((rc = put_words(P, &z, 8, 1))) return rc; }
  rc = run_chunks(P, x, nullptr, n, (int64_t)(kChunkBytes / es), es, false,
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
static BatchPipe g_batches[kMaxPipeDevices];  // per device ordinal, like g_pipes

// running = max(running, this tensor's amax); amax bit patterns of non-negative floats order like the floats
__global__ void merge_running_amax_kernel(const uint32_t* bits, float* running) {
  const uint32_t r = __float_as_uint(*running);
  *running = __uint_as_float(max(r, *bits));
}

}  // namespace mo

// kind: 0 = NVFP4 dynamic (block = 16 / 32 / 64), 1 = FP8 E4M3 per tensor.
// xs[i] may be a HOST pointer (uploaded once, kept resident between its amax and its fake-quant) or a DEVICE pointer
// (weights that already live on the GPU: no copy).  ys[i] may be a host pointer (the fake-quantized tensor is
// downloaded), a device pointer (written in place, no copy) or NULL (calibration-only: the fake-quantized tensor stays
// in the pipeline's device buffer for the consumer kernel and only running_amax[i] comes back - what max_calibrate needs).
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
    if (rows[i] < 0 || cols[i] < 0 || (rows[i] * cols[i] > 0 && !xs[i])) { set_error("mo_calib_qdq_host_batch: bad tensor %d", i); return MO_ERR_ARG; }
    const size_t b = (size_t)rows[i] * (size_t)cols[i] * es;
    if (b > max_bytes) max_bytes = b;
  }
  std::lock_guard<std::mutex> lock(g_pipe_mu);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxPipeDevices) { set_error("mo_calib_qdq_host_batch: cudaGetDevice: %s", cudaGetErrorString(e)); return e != cudaSuccess ? (int)e : MO_ERR_ARG; }
  BatchPipe& P = g_batches[dev];
  int rc = P.ensure(max_bytes ? max_bytes : 32, n_tensors);
  if (rc) return rc;
  auto on_device = [](const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
  };
#define MO_CK(call)                                                                                     \
  do {                                                                                                  \
    const cudaError_t ce_ = (call);                                                                     \
    if (ce_ != cudaSuccess && rc == MO_OK) {                                                            \
      set_error("mo_calib_qdq_host_batch: %s: %s", #call, cudaGetErrorString(ce_));                     \
      rc = (int)ce_;                                                                                    \
    }                                                                                                   \
  } while (0)
  MO_CK(cudaMemcpyAsync(P.d_amax, running_amax, (size_t)n_tensors * 4, cudaMemcpyHostToDevice, P.s_in));
  MO_CK(cudaMemsetAsync(P.d_bits, 0, (size_t)n_tensors * 4, P.s_in));
  bool used[2] = {false, false};
  for (int i = 0; i < n_tensors && rc == MO_OK; ++i) {
    const int b = i & 1;
    const int64_t n = rows[i] * cols[i];
    if (n == 0) continue;
    const bool x_dev = on_device(xs[i]);
    const bool y_dev = ys[i] != nullptr && on_device(ys[i]);
    // ---- upload + amax on s_in -------------------------------------------------------------------------------
    const void* src = xs[i];
    if (!x_dev) {
      if (used[b]) MO_CK(cudaStreamWaitEvent(P.s_in, P.freed[b], 0));  // the last fake-quant that read res[b] is done
      MO_CK(cudaMemcpyAsync(P.res[b], xs[i], (size_t)n * es, cudaMemcpyHostToDevice, P.s_in));
      src = P.res[b];
    }
    if (rc) break;
    rc = mo_amax(src, 1, 1, n, dtype, P.d_bits + i, P.s_in);
    if (rc) break;
    merge_running_amax_kernel<<<1, 1, 0, P.s_in>>>(P.d_bits + i, P.d_amax + i);
    MO_CK(cudaEventRecord(P.ready[b], P.s_in));
    // ---- fake-quant (+ download) on s_out: stream order also protects outb[b] from the previous download ----------
    MO_CK(cudaStreamWaitEvent(P.s_out, P.ready[b], 0));
    if (rc) break;
    void* dst = y_dev ? ys[i] : P.outb[b];
    if (kind == 0) rc = mo_qdq_nvfp4_dynamic(src, dst, rows[i], cols[i], dtype, P.d_amax + i, block, numerics, P.s_out);
    else rc = mo_qdq_fp8(src, dst, n, dtype, P.d_amax + i, 1, n, numerics, P.s_out);
    if (rc) break;
    MO_CK(cudaEventRecord(P.freed[b], P.s_out));
    used[b] = true;
    if (ys[i] != nullptr && !y_dev) MO_CK(cudaMemcpyAsync(ys[i], P.outb[b], (size_t)n * es, cudaMemcpyDeviceToHost, P.s_out));
  }
  MO_CK(cudaStreamSynchronize(P.s_in));
  MO_CK(cudaStreamSynchronize(P.s_out));
  if (rc) return rc;
  MO_CK(cudaMemcpy(running_amax, P.d_amax, (size_t)n_tensors * 4, cudaMemcpyDeviceToHost));
  MO_CK(cudaGetLastError());
#undef MO_CK
  return rc;
}
