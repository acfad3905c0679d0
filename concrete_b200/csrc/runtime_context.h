// runtime_context.h — minimal stand-in for mlir::concretelang::RuntimeContext
// (compilers/concrete-compiler/compiler/include/concretelang/Runtime/context.h:42-154).
//
// The real class needs Cap'n Proto keysets, concrete-cpu and LLVM headers, none of which can be
// built here.  The wrappers in wrappers.cpp use exactly two of its members on the GPU path,
//   get_ksk_gpu(level, input_lwe_dim, output_lwe_dim, gpu_idx, stream, ksk_idx)   context.h:117-145
//   get_bsk_gpu(input_lwe_dim, poly_size, level, glwe_dim, gpu_idx, stream, bsk_idx) context.h:86-115
// with the same names, argument order and caching semantics (per (gpu, key) device copy behind a
// mutex, stream synchronised before publication).
//
// Porting note (honest scope): wrappers.cpp / sdfg.cpp also use what this shim ADDS to the class —
// slot(), active_devices, device_base, the seeded / packing key registries — so they do not build
// against the reference's context.h unchanged.  A maintainer either adds those members to the
// real RuntimeContext (they are self-contained: the GpuSlot struct below plus three accessors) or
// keeps them in a side table keyed by the context pointer.  The device-pointer C ABI
// (include/concrete_b200.h) has no such dependency and is the drop-in boundary proper.
//
// What the shim adds is the per-GPU state the reference re-creates on every call
// (wrappers.cpp:110-256: stream, index arrays, PBS scratch, device buffers): here it is created
// once per GPU and reused — the "batched-op dispatch" rewrite.
#pragma once
#include "../../include/concrete_b200.h"
#include "../../include/concrete_b200_runtime.h"
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include <memory>
#include <mutex>
#include <vector>

namespace mlir {
namespace concretelang {

// A key is held either in full or in SEEDED form (bodies + the 128-bit compression seed, as the
// keyset ships it: Common/Keys.cpp:193-213); `len` is always the FULL key's word count, which is
// what sizes the device buffer (context.h:101-105).
struct HostKeyswitchKey {
  const uint64_t *data;
  size_t len; // words of the full key
  bool seeded = false;
  uint64_t seed_lo = 0, seed_hi = 0;
};
struct HostBootstrapKey {
  const uint64_t *data; // standard domain [n][l][k+1][k+1][N], or seeded [n][l][k+1][N]
  size_t len;           // words of the full key
  bool seeded = false;
  uint64_t seed_lo = 0, seed_hi = 0;
};

// grow-only device buffer owned by a GpuSlot
struct DevBuf {
  void *ptr = nullptr;
  uint64_t cap = 0;
};

struct GpuSlot {
  uint32_t gpu = 0;
  void *stream = nullptr;
  std::mutex mutex; // one batch at a time per GPU
  DevBuf ct_in, ct_mid, ct_out, lut, lut_idx, iota;
  uint32_t iota_len = 0;
  int8_t *pbs_buffer = nullptr;
  uint32_t pbs_k = 0, pbs_N = 0, pbs_l = 0, pbs_count = 0;
  std::vector<uint64_t> host_idx; // staging for lut indexes (must outlive the async copy)
  // copy/compute overlap inside one batched call: chunk c+1 uploads and chunk c-1 downloads while
  // chunk c computes (two extra streams on the two copy engines, one event pair per chunk)
  void *h2d_stream = nullptr, *d2h_stream = nullptr;
  std::vector<cudaEvent_t> ev_in, ev_done;
};

typedef struct RuntimeContext {
  RuntimeContext() {
    num_devices = cuda_get_number_of_gpus();
    const char *env = std::getenv("SDFG_NUM_GPUS"); // same knob as GPUDFG.cpp:1749-1755
    active_devices = 1;
    if (env) {
      int v = std::atoi(env);
      if (v >= 1 && v <= num_devices) active_devices = v;
    }
    for (int i = 0; i < num_devices; ++i) {
      bsk_gpu_mutex.push_back(std::make_unique<std::mutex>());
      ksk_gpu_mutex.push_back(std::make_unique<std::mutex>());
      slots.push_back(std::make_unique<GpuSlot>());
      slots.back()->gpu = (uint32_t)i;
    }
    bsk_gpu.resize(num_devices);
    ksk_gpu.resize(num_devices);
    pksk_gpu.resize(num_devices);
  }

  virtual ~RuntimeContext() {
    for (int i = 0; i < num_devices; ++i) {
      for (auto k : bsk_gpu[i])
        if (k != nullptr) cuda_drop(k, i);
      for (auto k : ksk_gpu[i])
        if (k != nullptr) cuda_drop(k, i);
      for (auto k : pksk_gpu[i])
        if (k != nullptr) cuda_drop(k, i);
      GpuSlot &s = *slots[i];
      if (s.stream) {
        cuda_synchronize_stream(s.stream, i);
        if (s.pbs_buffer) cleanup_cuda_programmable_bootstrap(s.stream, i, &s.pbs_buffer);
        for (DevBuf *b : {&s.ct_in, &s.ct_mid, &s.ct_out, &s.lut, &s.lut_idx, &s.iota})
          if (b->ptr) cuda_drop(b->ptr, i);
        if (s.h2d_stream) cuda_destroy_stream(s.h2d_stream, i);
        if (s.d2h_stream) cuda_destroy_stream(s.d2h_stream, i);
        for (auto &v : {&s.ev_in, &s.ev_done})
          for (cudaEvent_t e : *v) cudaEventDestroy(e);
        cuda_destroy_stream(s.stream, i);
      }
    }
  }

  size_t add_keyswitch_key(const uint64_t *data, size_t len) {
    host_ksks.push_back({data, len});
    for (auto &v : ksk_gpu) v.push_back(nullptr);
    return host_ksks.size() - 1;
  }
  size_t add_bootstrap_key(const uint64_t *data, size_t len) {
    host_bsks.push_back({data, len});
    for (auto &v : bsk_gpu) v.push_back(nullptr);
    return host_bsks.size() - 1;
  }

  // seeded forms: the masks are regenerated on each device (csrc/seeded.cu) instead of
  // decompressing on the host and uploading the full key
  size_t add_seeded_keyswitch_key(const uint64_t *bodies, size_t full_len, uint64_t seed_lo,
                                  uint64_t seed_hi) {
    size_t id = add_keyswitch_key(bodies, full_len);
    host_ksks[id].seeded = true;
    host_ksks[id].seed_lo = seed_lo;
    host_ksks[id].seed_hi = seed_hi;
    return id;
  }
  size_t add_seeded_bootstrap_key(const uint64_t *bodies, size_t full_len, uint64_t seed_lo,
                                  uint64_t seed_hi) {
    size_t id = add_bootstrap_key(bodies, full_len);
    host_bsks[id].seeded = true;
    host_bsks[id].seed_lo = seed_lo;
    host_bsks[id].seed_hi = seed_hi;
    return id;
  }

  // packing keyswitch keys of the circuit bootstrap (context.h fp_keyswitch_key_buffer)
  size_t add_packing_keyswitch_key(const uint64_t *data, size_t len) {
    host_pksks.push_back({data, len});
    for (auto &v : pksk_gpu) v.push_back(nullptr);
    return host_pksks.size() - 1;
  }
  // raw upload, cached per device like get_ksk_gpu
  void *get_pksk_gpu(uint32_t gpu_idx, void *stream, uint32_t pksk_idx) {
    if (pksk_gpu[gpu_idx][pksk_idx] != nullptr) return pksk_gpu[gpu_idx][pksk_idx];
    const std::lock_guard<std::mutex> guard(*ksk_gpu_mutex[gpu_idx]);
    if (pksk_gpu[gpu_idx][pksk_idx] != nullptr) return pksk_gpu[gpu_idx][pksk_idx];
    const HostKeyswitchKey &key = host_pksks[pksk_idx];
    void *tmp = cuda_malloc_async(sizeof(uint64_t) * key.len, stream, gpu_idx);
    cuda_memcpy_async_to_gpu(tmp, const_cast<uint64_t *>(key.data), sizeof(uint64_t) * key.len,
                             stream, gpu_idx);
    cuda_synchronize_stream(stream, gpu_idx);
    pksk_gpu[gpu_idx][pksk_idx] = tmp;
    return tmp;
  }

  virtual const uint64_t *keyswitch_key_buffer(size_t keyId) { return host_ksks[keyId].data; }

  // context.h:86-115
  void *get_bsk_gpu(uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
                    uint32_t glwe_dim, uint32_t gpu_idx, void *stream, uint32_t bsk_idx) {
    if (bsk_gpu[gpu_idx][bsk_idx] != nullptr) return bsk_gpu[gpu_idx][bsk_idx];
    const std::lock_guard<std::mutex> guard(*bsk_gpu_mutex[gpu_idx]);
    if (bsk_gpu[gpu_idx][bsk_idx] != nullptr) return bsk_gpu[gpu_idx][bsk_idx];
    const HostBootstrapKey &bsk = host_bsks[bsk_idx];
    size_t bsk_gpu_buffer_size = bsk.len * sizeof(double);
    void *bsk_gpu_tmp = cuda_malloc_async(bsk_gpu_buffer_size, stream, gpu_idx);
    if (bsk.seeded)
      cuda_convert_seeded_lwe_programmable_bootstrap_key_64(stream, gpu_idx, bsk_gpu_tmp, bsk.data,
                                                            input_lwe_dim, glwe_dim, level,
                                                            poly_size, bsk.seed_lo, bsk.seed_hi);
    else
      cuda_convert_lwe_programmable_bootstrap_key_64(stream, gpu_idx, bsk_gpu_tmp,
                                                     const_cast<uint64_t *>(bsk.data),
                                                     input_lwe_dim, glwe_dim, level, poly_size);
    cuda_synchronize_stream(stream, gpu_idx);
    bsk_gpu[gpu_idx][bsk_idx] = bsk_gpu_tmp;
    return bsk_gpu_tmp;
  }

  // context.h:117-145
  void *get_ksk_gpu(uint32_t level, uint32_t input_lwe_dim, uint32_t output_lwe_dim,
                    uint32_t gpu_idx, void *stream, uint32_t ksk_idx) {
    if (ksk_gpu[gpu_idx][ksk_idx] != nullptr) return ksk_gpu[gpu_idx][ksk_idx];
    const std::lock_guard<std::mutex> guard(*ksk_gpu_mutex[gpu_idx]);
    if (ksk_gpu[gpu_idx][ksk_idx] != nullptr) return ksk_gpu[gpu_idx][ksk_idx];
    const HostKeyswitchKey &ksk = host_ksks[ksk_idx];
    size_t ksk_buffer_size = sizeof(uint64_t) * ksk.len;
    void *ksk_gpu_tmp = cuda_malloc_async(ksk_buffer_size, stream, gpu_idx);
    if (ksk.seeded)
      cuda_decompress_seeded_lwe_keyswitch_key_64(stream, gpu_idx, ksk_gpu_tmp, ksk.data,
                                                  input_lwe_dim, output_lwe_dim, level, ksk.seed_lo,
                                                  ksk.seed_hi);
    else
      cuda_memcpy_async_to_gpu(ksk_gpu_tmp, const_cast<uint64_t *>(ksk.data), ksk_buffer_size,
                               stream, gpu_idx);
    cuda_synchronize_stream(stream, gpu_idx);
    ksk_gpu[gpu_idx][ksk_idx] = ksk_gpu_tmp;
    return ksk_gpu_tmp;
  }

  GpuSlot &slot(uint32_t gpu_idx) {
    GpuSlot &s = *slots[gpu_idx];
    if (!s.stream) s.stream = cuda_create_stream(gpu_idx);
    return s;
  }

  int num_devices = 0;
  int active_devices = 1;
  int device_base = 0; // first device of this context (one process per GPU: LOCAL_RANK)

private:
  std::vector<HostKeyswitchKey> host_ksks;
  std::vector<HostBootstrapKey> host_bsks;
  std::vector<HostKeyswitchKey> host_pksks;
  std::vector<std::vector<void *>> pksk_gpu;
  std::vector<std::unique_ptr<std::mutex>> bsk_gpu_mutex;
  std::vector<std::vector<void *>> bsk_gpu;
  std::vector<std::unique_ptr<std::mutex>> ksk_gpu_mutex;
  std::vector<std::vector<void *>> ksk_gpu;
  std::vector<std::unique_ptr<GpuSlot>> slots;
} RuntimeContext;

} // namespace concretelang
} // namespace mlir
