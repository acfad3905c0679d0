// peak.cu — FP64 FMA peak probe (measurement aid for bench.py's roofline denominator;
// MEASURED_PEAKS.json has no FP64 entry).  Each thread runs `iters` rounds of 16 independent
// DFMA chains; flops = blocks*threads*iters*16*2.
#include "common.cuh"
#include "../../include/concrete_b200.h"

namespace {
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, u32 iters) {
  double a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = 1.0 + 1e-9 * (threadIdx.x + i);
  const double m = 1.0000001, c = 1e-12;
  for (u32 it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fma(a[i], m, c);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
} // namespace

extern "C" void concrete_b200_fp64_peak(void *stream, uint32_t gpu_index, void *scratch,
                                        uint32_t blocks, uint32_t iters) {
  CB_CHECK(cudaSetDevice(gpu_index));
  fp64_peak_kernel<<<blocks, 256, 0, as_stream(stream)>>>((double *)scratch, iters);
  CB_CHECK(cudaGetLastError());
}
