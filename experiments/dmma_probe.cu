// dmma_probe.cu — is the FP64 tensor path (mma.sync ... f64, SASS DMMA) a second FP64 pipe on B200?
// BASELINE.json's north star allows "FP64 DMMA tensor cores ... only if ncu shows that product is a
// real dense contraction"; before asking that question this measures what the instruction is worth:
//   dfma   : 16 independent DFMA chains per thread                       (the roofline denominator)
//   dmma84 : 8 independent mma.m8n8k4.f64 accumulators per warp
//   dmma16 : 4 independent mma.m16n8k16.f64 accumulators per warp
//   mixed  : the dfma and dmma16 bodies interleaved in one instruction stream
// If mixed takes max(dfma, dmma16) the two are separate pipes; if it takes the sum they share one.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o experiments/dmma_probe experiments/dmma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma84(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0, %1, %2, %3}, {%4, %5, %6, %7, %8, %9, "
      "%10, %11}, {%12, %13, %14, %15}, {%0, %1, %2, %3};"
      : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
      : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
        "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int MODE> __global__ void __launch_bounds__(256) probe(double *out, unsigned iters) {
  double f[16], c2[8][2], c4[4][4], a8[8], b4[4];
#pragma unroll
  for (int i = 0; i < 16; ++i) f[i] = 1.0 + 1e-9 * (threadIdx.x + i);
#pragma unroll
  for (int i = 0; i < 8; ++i) c2[i][0] = c2[i][1] = 1e-3 * i, a8[i] = 1e-3 + 1e-9 * threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    b4[i] = 1e-3;
#pragma unroll
    for (int j = 0; j < 4; ++j) c4[i][j] = 1e-3 * (i + j);
  }
  const double m = 1.0000001, c = 1e-12;
  for (unsigned it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 3) {
#pragma unroll
      for (int i = 0; i < 16; ++i) f[i] = fma(f[i], m, c);
    }
    if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i) dmma84(c2[i], a8[0], b4[0]);
    }
    if (MODE == 2 || MODE == 3) {
#pragma unroll
      for (int i = 0; i < 4; ++i) dmma16(c4[i], a8, b4);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += f[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c2[i][0] + c2[i][1];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += c4[i][j];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> float run(double *d, unsigned iters, int blocks) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  probe<MODE><<<blocks, 256>>>(d, iters / 10);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<MODE><<<blocks, 256>>>(d, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 8;
  const unsigned iters = 200000;
  double *d;
  cudaMalloc(&d, (size_t)blocks * 256 * 8);
  const double threads = (double)blocks * 256, warps = threads / 32;
  const float t0 = run<0>(d, iters, blocks), t1 = run<1>(d, iters, blocks), t2 = run<2>(d, iters, blocks),
              t3 = run<3>(d, iters, blocks);
  const double f0 = threads * iters * 16 * 2, f1 = warps * iters * 8 * 512.0, f2 = warps * iters * 4 * 4096.0;
  printf("{\"sms\": %d, \"dfma_ms\": %.3f, \"dfma_tflops\": %.2f, \"dmma_m8n8k4_ms\": %.3f, "
         "\"dmma_m8n8k4_tflops\": %.2f, \"dmma_m16n8k16_ms\": %.3f, \"dmma_m16n8k16_tflops\": %.2f, "
         "\"mixed_ms\": %.3f, \"mixed_tflops\": %.2f, \"err\": \"%s\"}\n",
         sms, t0, f0 / t0 * 1e-9, t1, f1 / t1 * 1e-9, t2, f2 / t2 * 1e-9, t3, (f0 + f2) / t3 * 1e-9,
         cudaGetErrorString(cudaGetLastError()));
  return 0;
}
