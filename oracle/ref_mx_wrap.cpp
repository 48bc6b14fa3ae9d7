// TEST INFRASTRUCTURE ONLY.  Exposes the host side of the reference's own MX conversion header
// (modelopt/torch/kernels/quantization/gemm/tensor_quant_mx.h, compiled where it lies - not copied) through a C ABI so
// tests can pin oracle/mx.py's element rounding against it.  Built by oracle/build_ref.py into oracle/_ref/.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <stdexcept>

using std::isinf;
using std::isnan;

#include "tensor_quant_mx.h"

extern "C" {
float ref_mx_convert(float x, int fmt) { return convert_to_types(x, static_cast<Types>(fmt)); }
float ref_mx_format_max(int fmt) { return get_format_max(static_cast<Types>(fmt)); }
int ref_mx_lg2_max(int fmt) { return get_format_lg2_max(static_cast<Types>(fmt)); }
void ref_mx_convert_many(const float* x, float* y, long n, int fmt) {
  for (long i = 0; i < n; ++i) y[i] = convert_to_types(x[i], static_cast<Types>(fmt));
}
}
