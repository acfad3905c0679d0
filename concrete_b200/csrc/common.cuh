// common.cuh — error handling + shared types for the B200 TFHE hot-path library.
// Error convention follows the reference: print to stderr and abort()
// (third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda/include/device.h:14-28).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <atomic>

typedef uint64_t u64;
typedef int64_t i64;
typedef uint32_t u32;

#define CB_CHECK(ans)                                                                    \
  do {                                                                                   \
    cudaError_t cb_err__ = (ans);                                                        \
    if (cb_err__ != cudaSuccess) {                                                       \
      std::fprintf(stderr, "Cuda error: %s %s %d\n", cudaGetErrorString(cb_err__),       \
                   __FILE__, __LINE__);                                                  \
      std::abort();                                                                      \
    }                                                                                    \
  } while (0)

#define CB_PANIC(format, ...)                                                            \
  do {                                                                                   \
    std::fprintf(stderr, "%s::%d::%s: panic.\n" format "\n", __FILE__, __LINE__,         \
                 __func__, ##__VA_ARGS__);                                               \
    std::abort();                                                                        \
  } while (0)

// counts kernels launched by this library (bench.py reports it as gpu_launches)
extern std::atomic<uint64_t> g_cb_kernel_launches;
#define CB_LAUNCHED() (g_cb_kernel_launches.fetch_add(1, std::memory_order_relaxed))

static inline cudaStream_t as_stream(void *s) { return static_cast<cudaStream_t>(s); }
