// wrappers.cpp — the runtime's batched-op dispatch, rebuilt on top of the C ABI.
//
// Same prototypes (MLIR memref ABI) and the same argument meaning / assertions as
//   compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:70-106   (single ciphertext)
//   ...:110-162  memref_batched_keyswitch_lwe_cuda_u64
//   ...:164-256  memref_batched_bootstrap_lwe_cuda_u64
//   ...:258-363  memref_batched_mapped_bootstrap_lwe_cuda_u64
//   ...:745-849  memref_batched_{keyswitch,bootstrap,mapped_bootstrap}_lwe_u64 (CPU-named; here
//                they run the same GPU path: the north star asks for no CPU fallback).
// What changes relative to the reference implementation of these functions:
//   * nothing is created per call: stream, device buffers, index arrays and the PBS scratch live
//     in the RuntimeContext's per-GPU slot (the reference creates a stream, cudaMallocs inputs,
//     outputs and indexes, and allocates + frees the PBS scratch on every call, :125-160,:176-255);
//   * the trivial GLWE accumulator is assembled on the device (memset + strided copy of the
//     N-word table) instead of on the host (:199-213);
//   * batches are sharded over SDFG_NUM_GPUS devices with one host thread per device, contiguous
//     split S_g = floor(S/G) (+1 for g < S mod G) — the reference hard-wires gpu 0 (:122,:177);
//   * an extra fused entry point chains KS -> PBS without the host round trip.
#include "runtime_context.h"
#include <cassert>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include <functional>
#include <thread>
#include <vector>

using mlir::concretelang::DevBuf;
using mlir::concretelang::GpuSlot;
using mlir::concretelang::RuntimeContext;

namespace {

#define W_CHECK(ans)                                                                     \
  do {                                                                                   \
    cudaError_t e__ = (ans);                                                             \
    if (e__ != cudaSuccess) {                                                            \
      std::fprintf(stderr, "Cuda error: %s %s %d\n", cudaGetErrorString(e__), __FILE__,  \
                   __LINE__);                                                            \
      std::abort();                                                                      \
    }                                                                                    \
  } while (0)

void *ensure(DevBuf &b, uint64_t bytes, GpuSlot &s) {
  if (b.cap >= bytes && b.ptr) return b.ptr;
  if (b.ptr) {
    cuda_synchronize_stream(s.stream, s.gpu);
    cuda_drop(b.ptr, s.gpu);
  }
  uint64_t cap = bytes + bytes / 4 + 256;
  b.ptr = cuda_malloc(cap, s.gpu);
  b.cap = cap;
  return b.ptr;
}

__global__ void k_fill_iota(uint64_t *p, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    p[i] = i;
}

void *get_iota(GpuSlot &s, uint32_t n) {
  if (s.iota_len >= n && s.iota.ptr) return s.iota.ptr;
  uint32_t len = 1024;
  while (len < n) len <<= 1;
  ensure(s.iota, (uint64_t)len * 8, s);
  W_CHECK(cudaSetDevice(s.gpu));
  k_fill_iota<<<(len + 255) / 256, 256, 0, (cudaStream_t)s.stream>>>((uint64_t *)s.iota.ptr, len);
  W_CHECK(cudaGetLastError());
  s.iota_len = len;
  return s.iota.ptr;
}

void ensure_pbs(GpuSlot &s, uint32_t k, uint32_t N, uint32_t l, uint32_t count) {
  if (s.pbs_buffer && s.pbs_k == k && s.pbs_N == N && s.pbs_l == l && s.pbs_count >= count) return;
  if (s.pbs_buffer) cleanup_cuda_programmable_bootstrap(s.stream, s.gpu, &s.pbs_buffer);
  scratch_cuda_programmable_bootstrap_64(s.stream, s.gpu, &s.pbs_buffer, k, N, l, count, true);
  s.pbs_k = k;
  s.pbs_N = N;
  s.pbs_l = l;
  s.pbs_count = count;
}

// contiguous split of [0,total) over the active devices, one host thread per device
// (GPUDFG.cpp:852-895; split rule helper_multi_gpu.cu:53-83)
void for_each_shard(RuntimeContext *ctx, uint64_t total,
                    const std::function<void(uint32_t gpu, uint64_t begin, uint64_t count)> &f) {
  const uint64_t G = (uint64_t)std::max(1, std::min<int>(ctx->active_devices, (int)total));
  const uint32_t base = (uint32_t)ctx->device_base;
  if (G == 1) {
    f(base, 0, total);
    return;
  }
  std::vector<std::thread> threads;
  uint64_t begin = 0;
  for (uint64_t g = 0; g < G; ++g) {
    uint64_t cnt = total / G + (g < total % G ? 1 : 0);
    threads.emplace_back(f, base + (uint32_t)g, begin, cnt);
    begin += cnt;
  }
  for (auto &t : threads) t.join();
}

// Upload `num_luts` tables of N words as trivial GLWE accumulators [(k+1)*N] (wrappers.cpp:199-213)
void *upload_luts(GpuSlot &s, const uint64_t *tlu, uint64_t num_luts, uint64_t tlu_row_stride,
                  uint32_t k, uint32_t N) {
  const uint64_t glwe_words = (uint64_t)(k + 1) * N;
  void *d = ensure(s.lut, num_luts * glwe_words * 8, s);
  cuda_memset_async(d, 0, num_luts * glwe_words * 8, s.stream, s.gpu);
  W_CHECK(cudaSetDevice(s.gpu));
  W_CHECK(cudaMemcpy2DAsync((uint64_t *)d + (uint64_t)k * N, glwe_words * 8, tlu,
                            tlu_row_stride * 8, (uint64_t)N * 8, num_luts,
                            cudaMemcpyHostToDevice, (cudaStream_t)s.stream));
  return d;
}

// Chunk boundaries of one shard for the copy/compute pipeline: up to 4 chunks, each a whole
// number of `quantum` samples (one wave of the persistent PBS grid / one GEMM tile row), so that
// chunking adds no partially filled wave.  Small batches stay one chunk.
std::vector<uint64_t> chunk_bounds(uint64_t cnt, uint64_t quantum) {
  std::vector<uint64_t> b{0};
  if (cnt < 4 * quantum) {
    b.push_back(cnt);
    return b;
  }
  const uint64_t per = ((cnt + 3) / 4 + quantum - 1) / quantum * quantum;
  for (uint64_t o = per; o < cnt; o += per) b.push_back(o);
  b.push_back(cnt);
  return b;
}

void ensure_pipe(GpuSlot &s, size_t chunks) {
  W_CHECK(cudaSetDevice(s.gpu));
  if (!s.h2d_stream) s.h2d_stream = cuda_create_stream(s.gpu);
  if (!s.d2h_stream) s.d2h_stream = cuda_create_stream(s.gpu);
  while (s.ev_in.size() < chunks) {
    cudaEvent_t a, b;
    W_CHECK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
    W_CHECK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
    s.ev_in.push_back(a);
    s.ev_done.push_back(b);
  }
}

// H2D of every chunk on the upload stream, compute(chunk) on the slot's stream once its input
// landed, D2H of the chunk on the download stream once it is computed; returns after all three
// streams drained.  in/out rows are in_w / out_w words.
void run_pipelined(GpuSlot &s, const std::vector<uint64_t> &bounds, const uint64_t *h_in,
                   uint64_t in_w, void *d_in, uint64_t *h_out, uint64_t out_w, void *d_out,
                   const std::function<void(uint64_t off, uint64_t n)> &compute) {
  const size_t C = bounds.size() - 1;
  const uint32_t gpu = s.gpu;
  if (C == 1) {
    cuda_memcpy_async_to_gpu(d_in, const_cast<uint64_t *>(h_in), bounds[1] * in_w * 8, s.stream, gpu);
    compute(0, bounds[1]);
    cuda_memcpy_async_to_cpu(h_out, d_out, bounds[1] * out_w * 8, s.stream, gpu);
    cuda_synchronize_stream(s.stream, gpu);
    return;
  }
  ensure_pipe(s, C);
  // whatever the slot's stream still holds (key conversion, LUT upload) precedes the pipeline
  for (size_t c = 0; c < C; ++c) {
    const uint64_t off = bounds[c], n = bounds[c + 1] - off;
    cuda_memcpy_async_to_gpu((uint64_t *)d_in + off * in_w, const_cast<uint64_t *>(h_in + off * in_w),
                             n * in_w * 8, s.h2d_stream, gpu);
    W_CHECK(cudaEventRecord(s.ev_in[c], (cudaStream_t)s.h2d_stream));
  }
  for (size_t c = 0; c < C; ++c) {
    const uint64_t off = bounds[c], n = bounds[c + 1] - off;
    W_CHECK(cudaStreamWaitEvent((cudaStream_t)s.stream, s.ev_in[c], 0));
    compute(off, n);
    W_CHECK(cudaEventRecord(s.ev_done[c], (cudaStream_t)s.stream));
    W_CHECK(cudaStreamWaitEvent((cudaStream_t)s.d2h_stream, s.ev_done[c], 0));
    cuda_memcpy_async_to_cpu(h_out + off * out_w, (uint64_t *)d_out + off * out_w, n * out_w * 8,
                             s.d2h_stream, gpu);
  }
  cuda_synchronize_stream(s.d2h_stream, gpu);
  cuda_synchronize_stream(s.stream, gpu);
}

uint64_t pbs_wave(uint32_t gpu) { // samples per wave of the persistent PBS grid (6 per SM)
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, (int)gpu);
  return (uint64_t)sms * 6;
}

void bootstrap_impl(uint64_t *out, const uint64_t *ct0, uint64_t num_samples, uint64_t ct0_size1,
                    uint64_t out_size1, const uint64_t *tlu, uint64_t num_luts,
                    uint64_t tlu_row_stride, uint32_t input_lwe_dim, uint32_t poly_size,
                    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
                    RuntimeContext *context) {
  for_each_shard(context, num_samples, [&](uint32_t gpu, uint64_t begin, uint64_t cnt) {
    if (cnt == 0) return;
    GpuSlot &s = context->slot(gpu);
    const std::lock_guard<std::mutex> guard(s.mutex);
    void *fbsk = context->get_bsk_gpu(input_lwe_dim, poly_size, level, glwe_dim, gpu, s.stream,
                                      bsk_index);
    void *d_in = ensure(s.ct_in, cnt * ct0_size1 * 8, s);
    void *d_out = ensure(s.ct_out, cnt * out_size1 * 8, s);
    void *d_lut, *d_lut_idx;
    if (num_luts == 1) {
      d_lut = upload_luts(s, tlu, 1, tlu_row_stride, glwe_dim, poly_size);
      d_lut_idx = ensure(s.lut_idx, cnt * 8, s);
      cuda_memset_async(d_lut_idx, 0, cnt * 8, s.stream, gpu);
    } else {
      // one table per sample: this shard's tables, indexes 0..cnt-1
      d_lut = upload_luts(s, tlu + begin * tlu_row_stride, cnt, tlu_row_stride, glwe_dim,
                          poly_size);
      d_lut_idx = get_iota(s, (uint32_t)cnt);
    }
    void *iota = get_iota(s, (uint32_t)cnt);
    ensure_pbs(s, glwe_dim, poly_size, level, (uint32_t)cnt);
    run_pipelined(s, chunk_bounds(cnt, pbs_wave(gpu)), ct0 + begin * ct0_size1, ct0_size1, d_in,
                  out + begin * out_size1, out_size1, d_out, [&](uint64_t off, uint64_t n) {
                    cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
                        s.stream, gpu, (uint64_t *)d_out + off * out_size1, iota, d_lut,
                        (uint64_t *)d_lut_idx + off, (uint64_t *)d_in + off * ct0_size1, iota, fbsk,
                        s.pbs_buffer, input_lwe_dim, glwe_dim, poly_size, base_log, level,
                        (uint32_t)n, 1, 1);
                  });
  });
}

} // namespace

extern "C" {

// ---- context factory (the real runtime builds RuntimeContext from a ServerKeyset,
//      lib/Runtime/context.cpp:32-57; host key buffers stay owned by the caller) ----------------
void *concrete_b200_context_create(void) { return new RuntimeContext(); }
void concrete_b200_context_destroy(void *ctx) { delete static_cast<RuntimeContext *>(ctx); }
uint64_t concrete_b200_context_add_ksk(void *ctx, const uint64_t *data, uint64_t len) {
  return static_cast<RuntimeContext *>(ctx)->add_keyswitch_key(data, len);
}
uint64_t concrete_b200_context_add_bsk(void *ctx, const uint64_t *data, uint64_t len) {
  return static_cast<RuntimeContext *>(ctx)->add_bootstrap_key(data, len);
}
uint64_t concrete_b200_context_add_seeded_ksk(void *ctx, const uint64_t *bodies,
                                              uint64_t full_len_words, uint64_t seed_lo,
                                              uint64_t seed_hi) {
  return static_cast<RuntimeContext *>(ctx)->add_seeded_keyswitch_key(bodies, full_len_words,
                                                                       seed_lo, seed_hi);
}
uint64_t concrete_b200_context_add_seeded_bsk(void *ctx, const uint64_t *bodies,
                                              uint64_t full_len_words, uint64_t seed_lo,
                                              uint64_t seed_hi) {
  return static_cast<RuntimeContext *>(ctx)->add_seeded_bootstrap_key(bodies, full_len_words,
                                                                       seed_lo, seed_hi);
}
void concrete_b200_context_set_num_gpus(void *ctx, int n) {
  RuntimeContext *c = static_cast<RuntimeContext *>(ctx);
  if (n >= 1 && c->device_base + n <= c->num_devices) c->active_devices = n;
}
void concrete_b200_context_set_device_base(void *ctx, int base) {
  RuntimeContext *c = static_cast<RuntimeContext *>(ctx);
  if (base >= 0 && base + c->active_devices <= c->num_devices) c->device_base = base;
}

void memref_batched_keyswitch_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint32_t level, uint32_t base_log,
    uint32_t input_lwe_dim, uint32_t output_lwe_dim, uint32_t ksk_index,
    RuntimeContext *context) {
  (void)out_allocated; (void)ct0_allocated; (void)out_stride0; (void)out_stride1;
  (void)ct0_stride0; (void)ct0_stride1;
  assert(out_size0 == ct0_size0);
  assert(out_size1 == output_lwe_dim + 1);
  assert(ct0_size1 == input_lwe_dim + 1);
  uint64_t *out = out_aligned + out_offset;
  const uint64_t *ct0 = ct0_aligned + ct0_offset;
  for_each_shard(context, out_size0, [&](uint32_t gpu, uint64_t begin, uint64_t cnt) {
    if (cnt == 0) return;
    GpuSlot &s = context->slot(gpu);
    const std::lock_guard<std::mutex> guard(s.mutex);
    void *ksk = context->get_ksk_gpu(level, input_lwe_dim, output_lwe_dim, gpu, s.stream, ksk_index);
    void *d_in = ensure(s.ct_in, cnt * ct0_size1 * 8, s);
    void *d_out = ensure(s.ct_out, cnt * out_size1 * 8, s);
    void *iota = get_iota(s, (uint32_t)cnt);
    run_pipelined(s, chunk_bounds(cnt, 1024), ct0 + begin * ct0_size1, ct0_size1, d_in,
                  out + begin * out_size1, out_size1, d_out, [&](uint64_t off, uint64_t n) {
                    cuda_keyswitch_lwe_ciphertext_vector_64(
                        s.stream, gpu, (uint64_t *)d_out + off * out_size1, iota,
                        (uint64_t *)d_in + off * ct0_size1, iota, ksk, input_lwe_dim,
                        output_lwe_dim, base_log, level, (uint32_t)n);
                  });
  });
}

void memref_batched_bootstrap_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride, uint32_t input_lwe_dim,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    RuntimeContext *context) {
  (void)out_allocated; (void)ct0_allocated; (void)tlu_allocated; (void)out_stride0;
  (void)out_stride1; (void)ct0_stride0; (void)ct0_stride1; (void)tlu_stride;
  assert(out_size0 == ct0_size0);
  assert(out_size1 == glwe_dim * poly_size + 1);
  assert(tlu_size == poly_size);
  (void)tlu_size;
  bootstrap_impl(out_aligned + out_offset, ct0_aligned + ct0_offset, out_size0, ct0_size1,
                 out_size1, tlu_aligned + tlu_offset, 1, poly_size, input_lwe_dim, poly_size,
                 level, base_log, glwe_dim, bsk_index, context);
}

void memref_batched_mapped_bootstrap_lwe_cuda_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size0, uint64_t tlu_size1, uint64_t tlu_stride0,
    uint64_t tlu_stride1, uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
    uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index, RuntimeContext *context) {
  (void)out_allocated; (void)ct0_allocated; (void)tlu_allocated; (void)out_stride0;
  (void)out_stride1; (void)ct0_stride0; (void)ct0_stride1; (void)tlu_stride0; (void)tlu_stride1;
  assert(out_size0 == ct0_size0);
  assert(out_size1 == glwe_dim * poly_size + 1);
  assert((out_size0 == tlu_size0 || tlu_size0 == 1) && "Number of LUTs does not match batch size");
  bootstrap_impl(out_aligned + out_offset, ct0_aligned + ct0_offset, out_size0, ct0_size1,
                 out_size1, tlu_aligned + tlu_offset, tlu_size0, tlu_size1, input_lwe_dim,
                 poly_size, level, base_log, glwe_dim, bsk_index, context);
}

// wrappers.cpp:70-106
void memref_keyswitch_lwe_cuda_u64(uint64_t *out_allocated, uint64_t *out_aligned,
                                   uint64_t out_offset, uint64_t out_size, uint64_t out_stride,
                                   uint64_t *ct0_allocated, uint64_t *ct0_aligned,
                                   uint64_t ct0_offset, uint64_t ct0_size, uint64_t ct0_stride,
                                   uint32_t level, uint32_t base_log, uint32_t input_lwe_dim,
                                   uint32_t output_lwe_dim, uint32_t ksk_index,
                                   RuntimeContext *context) {
  assert(out_stride == 1);
  assert(ct0_stride == 1);
  memref_batched_keyswitch_lwe_cuda_u64(out_allocated, out_aligned, out_offset, 1, out_size,
                                        out_size, out_stride, ct0_allocated, ct0_aligned,
                                        ct0_offset, 1, ct0_size, ct0_size, ct0_stride, level,
                                        base_log, input_lwe_dim, output_lwe_dim, ksk_index,
                                        context);
}

void memref_bootstrap_lwe_cuda_u64(uint64_t *out_allocated, uint64_t *out_aligned,
                                   uint64_t out_offset, uint64_t out_size, uint64_t out_stride,
                                   uint64_t *ct0_allocated, uint64_t *ct0_aligned,
                                   uint64_t ct0_offset, uint64_t ct0_size, uint64_t ct0_stride,
                                   uint64_t *tlu_allocated, uint64_t *tlu_aligned,
                                   uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride,
                                   uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
                                   uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
                                   RuntimeContext *context) {
  memref_batched_bootstrap_lwe_cuda_u64(out_allocated, out_aligned, out_offset, 1, out_size,
                                        out_size, out_stride, ct0_allocated, ct0_aligned,
                                        ct0_offset, 1, ct0_size, ct0_size, ct0_stride,
                                        tlu_allocated, tlu_aligned, tlu_offset, tlu_size,
                                        tlu_stride, input_lwe_dim, poly_size, level, base_log,
                                        glwe_dim, bsk_index, context);
}

// CPU-named wrappers (wrappers.cpp:745-849): same prototypes, same GPU path.
void memref_batched_keyswitch_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint32_t level, uint32_t base_log,
    uint32_t input_lwe_dim, uint32_t output_lwe_dim, uint32_t ksk_index,
    RuntimeContext *context) {
  memref_batched_keyswitch_lwe_cuda_u64(out_allocated, out_aligned, out_offset, out_size0,
                                        out_size1, out_stride0, out_stride1, ct0_allocated,
                                        ct0_aligned, ct0_offset, ct0_size0, ct0_size1,
                                        ct0_stride0, ct0_stride1, level, base_log, input_lwe_dim,
                                        output_lwe_dim, ksk_index, context);
}

void memref_batched_bootstrap_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size, uint64_t tlu_stride, uint32_t input_lwe_dim,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    RuntimeContext *context) {
  memref_batched_bootstrap_lwe_cuda_u64(out_allocated, out_aligned, out_offset, out_size0,
                                        out_size1, out_stride0, out_stride1, ct0_allocated,
                                        ct0_aligned, ct0_offset, ct0_size0, ct0_size1,
                                        ct0_stride0, ct0_stride1, tlu_allocated, tlu_aligned,
                                        tlu_offset, tlu_size, tlu_stride, input_lwe_dim,
                                        poly_size, level, base_log, glwe_dim, bsk_index, context);
}

void memref_batched_mapped_bootstrap_lwe_u64(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0,
    uint64_t out_size1, uint64_t out_stride0, uint64_t out_stride1, uint64_t *ct0_allocated,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t ct0_stride0, uint64_t ct0_stride1, uint64_t *tlu_allocated, uint64_t *tlu_aligned,
    uint64_t tlu_offset, uint64_t tlu_size0, uint64_t tlu_size1, uint64_t tlu_stride0,
    uint64_t tlu_stride1, uint32_t input_lwe_dim, uint32_t poly_size, uint32_t level,
    uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index, RuntimeContext *context) {
  memref_batched_mapped_bootstrap_lwe_cuda_u64(
      out_allocated, out_aligned, out_offset, out_size0, out_size1, out_stride0, out_stride1,
      ct0_allocated, ct0_aligned, ct0_offset, ct0_size0, ct0_size1, ct0_stride0, ct0_stride1,
      tlu_allocated, tlu_aligned, tlu_offset, tlu_size0, tlu_size1, tlu_stride0, tlu_stride1,
      input_lwe_dim, poly_size, level, base_log, glwe_dim, bsk_index, context);
}

// Extension: KS -> PBS with the intermediate ciphertexts kept on the device (what the SDFG
// runtime achieves per chunk, GPUDFG.cpp:1085,1209).  ct0: [S][big+1], out: [S][k*N+1].
void memref_batched_keyswitch_bootstrap_lwe_cuda_u64(
    uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size0, uint64_t out_size1,
    uint64_t *ct0_aligned, uint64_t ct0_offset, uint64_t ct0_size0, uint64_t ct0_size1,
    uint64_t *tlu_aligned, uint64_t tlu_offset, uint64_t tlu_size, uint32_t ks_level,
    uint32_t ks_base_log, uint32_t big_lwe_dim, uint32_t small_lwe_dim, uint32_t ksk_index,
    uint32_t poly_size, uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t bsk_index,
    RuntimeContext *context) {
  assert(out_size0 == ct0_size0);
  assert(out_size1 == glwe_dim * poly_size + 1);
  assert(ct0_size1 == big_lwe_dim + 1);
  assert(tlu_size == poly_size);
  (void)tlu_size;
  uint64_t *out = out_aligned + out_offset;
  const uint64_t *ct0 = ct0_aligned + ct0_offset;
  const uint64_t *tlu = tlu_aligned + tlu_offset;
  for_each_shard(context, out_size0, [&](uint32_t gpu, uint64_t begin, uint64_t cnt) {
    if (cnt == 0) return;
    GpuSlot &s = context->slot(gpu);
    const std::lock_guard<std::mutex> guard(s.mutex);
    void *ksk = context->get_ksk_gpu(ks_level, big_lwe_dim, small_lwe_dim, gpu, s.stream, ksk_index);
    void *fbsk = context->get_bsk_gpu(small_lwe_dim, poly_size, level, glwe_dim, gpu, s.stream,
                                      bsk_index);
    void *d_in = ensure(s.ct_in, cnt * ct0_size1 * 8, s);
    void *d_mid = ensure(s.ct_mid, cnt * (uint64_t)(small_lwe_dim + 1) * 8, s);
    void *d_out = ensure(s.ct_out, cnt * out_size1 * 8, s);
    void *d_lut = upload_luts(s, tlu, 1, poly_size, glwe_dim, poly_size);
    void *d_lut_idx = ensure(s.lut_idx, cnt * 8, s);
    cuda_memset_async(d_lut_idx, 0, cnt * 8, s.stream, gpu);
    ensure_pbs(s, glwe_dim, poly_size, level, (uint32_t)cnt);
    // chunk pipeline: big ciphertexts of chunk c+1 upload while chunk c runs KS -> PBS on the device
    // and chunk c-1 downloads; the small ciphertexts never leave the device
    const uint64_t mid_w = (uint64_t)small_lwe_dim + 1;
    run_pipelined(s, chunk_bounds(cnt, pbs_wave(gpu)), ct0 + begin * ct0_size1, ct0_size1, d_in,
                  out + begin * out_size1, out_size1, d_out, [&](uint64_t off, uint64_t n) {
                    cuda_keyswitch_bootstrap_lwe_ciphertext_vector_64(
                        s.stream, gpu, (uint64_t *)d_out + off * out_size1, d_lut,
                        (uint64_t *)d_lut_idx + off, (uint64_t *)d_in + off * ct0_size1, ksk, fbsk,
                        s.pbs_buffer, (uint64_t *)d_mid + off * mid_w, big_lwe_dim, small_lwe_dim,
                        ks_base_log, ks_level, glwe_dim, poly_size, base_log, level, (uint32_t)n);
                  });
  });
}

uint64_t concrete_b200_context_add_pksk(void *ctx, const uint64_t *data, uint64_t len) {
  return static_cast<RuntimeContext *>(ctx)->add_packing_keyswitch_key(data, len);
}

// WoP-PBS over CRT blocks, HOST memrefs — same prototype and argument meaning as the reference's
// CPU-only wrapper (compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:855-997); the
// whole chain (bit extraction, circuit bootstrap, vertical packing) runs on the first device of
// the context, operands go up once, only the `B` output blocks come back.
void memref_wop_pbs_crt_buffer(
    uint64_t *out_allocated, uint64_t *out_aligned, uint64_t out_offset, uint64_t out_size_0,
    uint64_t out_size_1, uint64_t out_stride_0, uint64_t out_stride_1, uint64_t *in_allocated,
    uint64_t *in_aligned, uint64_t in_offset, uint64_t in_size_0, uint64_t in_size_1,
    uint64_t in_stride_0, uint64_t in_stride_1, uint64_t *lut_ct_allocated,
    uint64_t *lut_ct_aligned, uint64_t lut_ct_offset, uint64_t lut_ct_size0, uint64_t lut_ct_size1,
    uint64_t lut_ct_stride0, uint64_t lut_ct_stride1, uint64_t *crt_decomp_allocated,
    uint64_t *crt_decomp_aligned, uint64_t crt_decomp_offset, uint64_t crt_decomp_size,
    uint64_t crt_decomp_stride, uint32_t lwe_small_dim, uint32_t cbs_level_count,
    uint32_t cbs_base_log, uint32_t ksk_level_count, uint32_t ksk_base_log,
    uint32_t bsk_level_count, uint32_t bsk_base_log, uint32_t fpksk_level_count,
    uint32_t fpksk_base_log, uint32_t polynomial_size, uint32_t ksk_index, uint32_t bsk_index,
    uint32_t pksk_index, RuntimeContext *context) {
  (void)out_allocated; (void)in_allocated; (void)lut_ct_allocated; (void)crt_decomp_allocated;
  (void)out_stride_0; (void)in_stride_1; (void)lut_ct_stride0; (void)lut_ct_stride1;
  (void)crt_decomp_stride;
  assert(out_stride_1 == 1);
  assert(in_stride_0 == in_size_1);
  assert(out_size_0 == in_size_0 && out_size_0 == crt_decomp_size);
  assert(out_size_1 == in_size_1);
  (void)out_stride_1; (void)in_stride_0; (void)in_size_0;
  const uint64_t big_size = in_size_1, big_dim = big_size - 1;
  assert(big_dim % polynomial_size == 0);
  const uint32_t glwe_dim = (uint32_t)(big_dim / polynomial_size);
  const uint32_t blocks = (uint32_t)crt_decomp_size;
  const uint64_t *crt = crt_decomp_aligned + crt_decomp_offset;
  assert(lut_ct_size0 == blocks);
  const uint32_t gpu = (uint32_t)context->device_base;
  GpuSlot &s = context->slot(gpu);
  const std::lock_guard<std::mutex> guard(s.mutex);
  void *ksk = context->get_ksk_gpu(ksk_level_count, (uint32_t)big_dim, lwe_small_dim, gpu, s.stream,
                                   ksk_index);
  void *fbsk = context->get_bsk_gpu(lwe_small_dim, polynomial_size, bsk_level_count, glwe_dim, gpu,
                                    s.stream, bsk_index);
  void *pksk = context->get_pksk_gpu(gpu, s.stream, pksk_index);
  const uint64_t ct_bytes = (uint64_t)blocks * big_size * 8, lut_bytes = lut_ct_size0 * lut_ct_size1 * 8;
  void *d_in = ensure(s.ct_in, ct_bytes, s);
  void *d_out = ensure(s.ct_out, ct_bytes, s);
  void *d_lut = ensure(s.lut, lut_bytes, s);
  cuda_memcpy_async_to_gpu(d_in, in_aligned + in_offset, ct_bytes, s.stream, gpu);
  cuda_memcpy_async_to_gpu(d_lut, lut_ct_aligned + lut_ct_offset, lut_bytes, s.stream, gpu);
  cuda_wop_pbs_crt_lwe_ciphertext_vector_64(
      s.stream, gpu, d_out, d_in, d_lut, crt, blocks, fbsk, ksk, pksk, lwe_small_dim, glwe_dim,
      polynomial_size, bsk_level_count, bsk_base_log, ksk_level_count, ksk_base_log,
      fpksk_level_count, fpksk_base_log, cbs_level_count, cbs_base_log);
  cuda_memcpy_async_to_cpu(out_aligned + out_offset, d_out, ct_bytes, s.stream, gpu);
  cuda_synchronize_stream(s.stream, gpu);
}

} // extern "C"
