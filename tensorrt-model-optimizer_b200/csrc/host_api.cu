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
