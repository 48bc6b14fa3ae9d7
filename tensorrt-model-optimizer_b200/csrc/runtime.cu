// Library runtime: error reporting, device info, device self-tests of the numerics tricks.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"

namespace mo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return MO_OK;
}

// ---- self-test kernels ---------------------------------------------------------------------------
__global__ void selftest_e2m1_kernel(unsigned long long* bad) {
  // every non-negative fp32 bit pattern (incl. inf / NaN), two at a time through the x2 converter
  unsigned long long local = 0;
  const uint64_t total = 1ull << 31;
  for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < total;
       i += (uint64_t)gridDim.x * blockDim.x * 2) {
    float a = __uint_as_float((uint32_t)i), b = __uint_as_float((uint32_t)i + 1);
    const float wa = e2m1_round_chain(a), wb = e2m1_round_chain(b);
    e2m1_roundtrip2(a, b);
    local += (a != wa) + (b != wb);
  }
  if (local) atomicAdd(bad, local);
}

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

__global__ void selftest_exact_div_kernel(int q_max, int n_scales, uint32_t seed, unsigned long long* bad) {
  unsigned long long local = 0;
  for (int si = blockIdx.x; si < n_scales; si += gridDim.x) {
    // random positive scale with exponent in [-70, 70] so both the fast window and the fallback run
    const uint32_t h = hash32(seed + 0x9e3779b9u * (uint32_t)si);
    const int e = (int)(hash32(h) % 141u) - 70;
    const float scale = __uint_as_float(((uint32_t)(e + 127) << 23) | (h & 0x7fffffu));
    const ExactDiv d(scale);
    for (int q = -q_max + (int)threadIdx.x; q <= q_max; q += blockDim.x) {
      const float a = (float)q;
      const float want = __fdiv_rn(a, scale), got = d(a);
      // -0 vs +0 for q == 0 is repaired by the caller's copysign; compare values
      local += !(want == got);
    }
  }
  if (local) atomicAdd(bad, local);
}

}  // namespace mo

using namespace mo;

extern "C" int mo_abi_version(void) { return MO_ABI_VERSION; }

extern "C" const char* mo_last_error(void) { return g_err; }

extern "C" int mo_device_info(char* name, int name_len, int* sm_count, int* cc) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    set_error("no CUDA device");
    return MO_ERR_NO_DEVICE;
  }
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceProp p;
  const cudaError_t e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) { set_error("cudaGetDeviceProperties: %s", cudaGetErrorString(e)); return (int)e; }
  if (name && name_len > 0) { strncpy(name, p.name, name_len - 1); name[name_len - 1] = 0; }
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc) *cc = p.major * 10 + p.minor;
  return MO_OK;
}

static int run_selftest_result(unsigned long long* d_bad, uint64_t* mismatches, const char* what) {
  int rc = check_launch(what);
  unsigned long long h = 0;
  cudaError_t e = cudaMemcpy(&h, d_bad, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d_bad);
  if (rc) return rc;
  if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return (int)e; }
  *mismatches = h;
  return MO_OK;
}

extern "C" int mo_selftest_e2m1(uint64_t* mismatches) {
  if (!mismatches) { set_error("mo_selftest_e2m1: NULL"); return MO_ERR_NULL; }
  unsigned long long* d_bad = nullptr;
  cudaError_t e = cudaMalloc(&d_bad, sizeof(*d_bad));
  if (e != cudaSuccess) { set_error("cudaMalloc: %s", cudaGetErrorString(e)); return (int)e; }
  cudaMemset(d_bad, 0, sizeof(*d_bad));
  selftest_e2m1_kernel<<<kNumSMs * 8, 256>>>(d_bad);
  return run_selftest_result(d_bad, mismatches, "mo_selftest_e2m1");
}

extern "C" int mo_selftest_exact_div(int q_max, int n_scales, uint64_t seed, uint64_t* mismatches) {
  if (!mismatches) { set_error("mo_selftest_exact_div: NULL"); return MO_ERR_NULL; }
  if (q_max < 1 || n_scales < 1) { set_error("mo_selftest_exact_div: bad args"); return MO_ERR_ARG; }
  unsigned long long* d_bad = nullptr;
  cudaError_t e = cudaMalloc(&d_bad, sizeof(*d_bad));
  if (e != cudaSuccess) { set_error("cudaMalloc: %s", cudaGetErrorString(e)); return (int)e; }
  cudaMemset(d_bad, 0, sizeof(*d_bad));
  const int grid = n_scales < kNumSMs * 8 ? n_scales : kNumSMs * 8;
  selftest_exact_div_kernel<<<grid, 256>>>(q_max, n_scales, (uint32_t)seed, d_bad);
  return run_selftest_result(d_bad, mismatches, "mo_selftest_exact_div");
}
