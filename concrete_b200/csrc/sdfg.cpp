// sdfg.cpp — the SDFG ("stream emulator") GPU scheduler, rebuilt GPU-only on top of the C ABI.
//
// Same entry points and prototypes as the reference's code-generation interface
//   compilers/concrete-compiler/compiler/include/concretelang/Runtime/stream_emulator_api.h:29-104
// implemented in the reference by
//   compilers/concrete-compiler/compiler/lib/Runtime/GPUDFG.cpp:1467-1799  (entry points)
//   ...:262-480 (Dependence), :523-956 (Stream, schedule_work :600-911), :1057-1457 (processes).
// This is the path `fhe.Compiler(..., use_gpu=True)` takes by default (SDFG lowering): the compiled
// program builds a graph of processes connected by streams, PUTs host memrefs into the input
// streams and GETs host memrefs from the output streams.
//
// What is kept: the graph model (streams with one producer, processes with 1-2 inputs and one
// output, demand-driven evaluation at GET, generations), the argument conventions (plaintext /
// cleartext operands as {1, num} or {1, 1} memrefs, one table {1, N} or one table per sample
// {num, N}, `data_ownership`), SDFG_NUM_GPUS, one host thread and one stream per device.
// What is different:
//   * GPU only: no CPU/GPU work splitting, no per-chunk host fallback (north star: no CPU path);
//   * a GET evaluates the whole pending sub-graph per device shard: every process of the chain
//     runs on the device on the shard's rows, intermediates never visit the host (the reference
//     gets there per chunk, GPUDFG.cpp:1085,1209, but re-allocates every buffer, index array and
//     PBS scratch per process call, :1091-1096,:1200-1205);
//   * inside a shard the rows go through a 3-stream pipeline in up to 4 chunks: chunk c+1 uploads
//     while chunk c computes and chunk c-1 downloads;
//   * device buffers, index arrays, PBS scratch and key copies are per-GPU state that lives as
//     long as the graph / the RuntimeContext; the trivial GLWE accumulator is assembled on the
//     device from the N-word table.
#include "runtime_context.h"
#include <algorithm>
#include <cassert>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <vector>

using mlir::concretelang::GpuSlot;
using mlir::concretelang::RuntimeContext;

typedef enum stream_type {
  TS_STREAM_TYPE_X86_TO_TOPO_LSAP,
  TS_STREAM_TYPE_TOPO_TO_TOPO_LSAP,
  TS_STREAM_TYPE_TOPO_TO_X86_LSAP,
  TS_STREAM_TYPE_TOPO_TO_BOTH,
  TS_STREAM_TYPE_X86_TO_X86_LSAP
} stream_type;

namespace {

#define S_CHECK(ans)                                                                     \
  do {                                                                                   \
    cudaError_t e__ = (ans);                                                             \
    if (e__ != cudaSuccess) {                                                            \
      std::fprintf(stderr, "Cuda error: %s %s %d\n", cudaGetErrorString(e__), __FILE__,  \
                   __LINE__);                                                            \
      std::abort();                                                                      \
    }                                                                                    \
  } while (0)
#define S_PANIC(...)                                                                     \
  do {                                                                                   \
    std::fprintf(stderr, "SDFG runtime: " __VA_ARGS__);                                  \
    std::fprintf(stderr, "\n");                                                          \
    std::abort();                                                                        \
  } while (0)

struct Process;
struct Dfg;

// Device memory of the scheduler comes from a library-owned stream-ordered pool that keeps what it
// has mapped: the default pool trims itself at every synchronisation, and re-mapping the
// ~0.6 GB a P2048 x 16 384 evaluation needs costs tens of milliseconds per GET.
cudaMemPool_t sdfg_pool(int gpu) {
  static std::mutex mu;
  static std::map<int, cudaMemPool_t> pools;
  std::lock_guard<std::mutex> g(mu);
  auto it = pools.find(gpu);
  if (it != pools.end()) return it->second;
  cudaMemPoolProps props = {};
  props.allocType = cudaMemAllocationTypePinned;
  props.handleTypes = cudaMemHandleTypeNone;
  props.location.type = cudaMemLocationTypeDevice;
  props.location.id = gpu;
  cudaMemPool_t pool;
  S_CHECK(cudaMemPoolCreate(&pool, &props));
  unsigned long long keep = ~0ull;
  S_CHECK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  pools[gpu] = pool;
  return pool;
}
uint64_t *pool_alloc(uint64_t words, cudaStream_t stream, int gpu) {
  void *p = nullptr;
  S_CHECK(cudaMallocFromPoolAsync(&p, std::max<uint64_t>(words, 1) * 8, sdfg_pool(gpu), stream));
  return static_cast<uint64_t *>(p);
}

// device copy of the rows [begin, begin + count) of a stream value on one GPU
struct Shard {
  uint32_t gpu = 0;
  uint64_t begin = 0, count = 0;
  uint64_t *dev = nullptr;
};

struct Stream {
  std::string name;
  stream_type type;
  Dfg *dfg = nullptr;
  Process *producer = nullptr;
  bool batched = false, ct = false, is_const = false;
  // current value
  uint64_t generation = 0;   // bumped by every PUT / every evaluation of the producer
  uint64_t rows = 0, cols = 0;
  uint64_t *host = nullptr;  // contiguous rows x cols, nullptr if the value lives on devices only
  bool host_owned = false;
  std::vector<Shard> shards; // device copies (ciphertext streams computed on the GPU)

  void drop_host() {
    if (host && host_owned) std::free(host);
    host = nullptr;
    host_owned = false;
  }
  void drop_device() {
    for (Shard &s : shards)
      if (s.dev) cuda_drop(s.dev, s.gpu);
    shards.clear();
  }
  ~Stream() {
    drop_host();
    drop_device();
  }
};

enum Kind { K_ADD, K_ADD_PLAIN, K_MUL_CLEAR, K_NEGATE, K_KEYSWITCH, K_BOOTSTRAP };

struct Process {
  Kind kind;
  Stream *in[2] = {nullptr, nullptr};
  Stream *out = nullptr;
  uint32_t level = 0, base_log = 0, input_lwe_dim = 0, output_lwe_dim = 0, poly_size = 0,
           glwe_dim = 0, key_index = 0, output_size = 0;
  RuntimeContext *ctx = nullptr;
  uint64_t in_gen[2] = {0, 0}; // generations of the inputs the current output was computed from
  bool ever_run = false;
};

struct Dfg {
  std::vector<Process *> procs;
  std::vector<Stream *> streams; // streams attached to a process of this graph
  int num_gpus = 1;
  ~Dfg() {
    for (Process *p : procs) delete p;
    for (Stream *s : streams) delete s;
  }
  void attach(Stream *s) {
    if (!s->dfg) {
      s->dfg = this;
      streams.push_back(s);
    }
  }
};

Process *make_process(void *dfg, Kind kind, void *sin1, void *sin2, void *sout) {
  Dfg *g = static_cast<Dfg *>(dfg);
  Process *p = new Process();
  p->kind = kind;
  p->in[0] = static_cast<Stream *>(sin1);
  p->in[1] = static_cast<Stream *>(sin2);
  p->out = static_cast<Stream *>(sout);
  if (p->out->producer) S_PANIC("stream '%s' already has a producer", p->out->name.c_str());
  p->out->producer = p;
  for (Stream *s : {p->in[0], p->in[1], p->out})
    if (s) g->attach(s);
  g->procs.push_back(p);
  return p;
}

// ---- PUT / host side -----------------------------------------------------------------------------
// A PUT without ownership must not keep a reference to the caller's memory (the reference copies:
// memref_copy_alloc, GPUDFG.cpp:1676,1704).  For a batch of ciphertexts that is only ever read as
// a ciphertext operand the copy goes straight to where the rows will be consumed: the device
// shards of the evaluation (same contiguous split), one blocking upload, no host copy at all.
bool upload_at_put(Stream *s, const uint64_t *src, uint64_t rows, uint64_t cols, uint64_t stride0) {
  Dfg *g = s->dfg;
  if (!g || rows < 2 || rows * cols < (1u << 16)) return false;
  RuntimeContext *ctx = nullptr;
  bool used = false;
  for (Process *p : g->procs) {
    const bool second_is_ct = p->kind == K_ADD;
    if (p->in[1] == s && !second_is_ct) return false; // table / scalar operand: wanted on the host
    if (p->in[0] == s || (p->in[1] == s && second_is_ct)) used = true;
    if ((p->in[0] == s || p->in[1] == s) && p->ctx) ctx = p->ctx;
  }
  if (!used) return false;
  if (!ctx)
    for (Process *p : g->procs)
      if (p->ctx) ctx = p->ctx;
  if (!ctx) return false;
  const uint64_t G = (uint64_t)std::max(1, std::min<int>(g->num_gpus, (int)std::min<uint64_t>(rows, 64)));
  const uint32_t base = (uint32_t)ctx->device_base;
  uint64_t begin = 0;
  std::vector<cudaStream_t> used_streams;
  for (uint64_t d = 0; d < G; ++d) {
    const uint64_t cnt = rows / G + (d < rows % G ? 1 : 0);
    const uint32_t gpu = base + (uint32_t)d;
    S_CHECK(cudaSetDevice((int)gpu));
    GpuSlot &slot = ctx->slot(gpu);
    if (!slot.h2d_stream) slot.h2d_stream = cuda_create_stream(gpu);
    cudaStream_t h2d = (cudaStream_t)slot.h2d_stream;
    Shard sh;
    sh.gpu = gpu, sh.begin = begin, sh.count = cnt;
    sh.dev = pool_alloc(cnt * cols, h2d, (int)gpu);
    if (stride0 == cols) // contiguous rows: one linear copy (the 2-D path is row-granular)
      S_CHECK(cudaMemcpyAsync(sh.dev, src + begin * cols, cnt * cols * 8, cudaMemcpyHostToDevice, h2d));
    else
      S_CHECK(cudaMemcpy2DAsync(sh.dev, cols * 8, src + begin * stride0, stride0 * 8, cols * 8, cnt,
                                cudaMemcpyHostToDevice, h2d));
    s->shards.push_back(sh);
    used_streams.push_back(h2d);
    begin += cnt;
  }
  for (uint64_t d = 0; d < G; ++d) { // the caller's buffer is free again when PUT returns
    S_CHECK(cudaSetDevice((int)(base + d)));
    S_CHECK(cudaStreamSynchronize(used_streams[d]));
  }
  return true;
}

void put_host(Stream *s, uint64_t *allocated, uint64_t *aligned, uint64_t offset, uint64_t size0,
              uint64_t size1, uint64_t stride0, uint64_t stride1, uint64_t ownership) {
  if (stride1 != 1 && size1 > 1) S_PANIC("strided memrefs not supported");
  s->drop_host();
  s->drop_device();
  s->rows = size0;
  s->cols = size1;
  const uint64_t *src = aligned + offset;
  if (!ownership && upload_at_put(s, src, size0, size1, stride0)) {
    s->generation++;
    return;
  }
  if (ownership && (stride0 == size1 || size0 <= 1) && allocated == aligned && offset == 0) {
    s->host = allocated; // adopted: freed when the value is replaced (GPUDFG.cpp:1673-1685)
    s->host_owned = true;
  } else {
    s->host = static_cast<uint64_t *>(std::malloc(std::max<uint64_t>(1, size0 * size1) * 8));
    s->host_owned = true;
    for (uint64_t r = 0; r < size0; ++r)
      std::memcpy(s->host + r * size1, src + r * stride0, size1 * 8);
    if (ownership) std::free(allocated);
  }
  s->generation++;
}

// ---- evaluation ----------------------------------------------------------------------------------
// processes whose output is stale, in dependency order, needed to bring `s` up to date
void collect(Stream *s, std::vector<Process *> &plan, std::vector<Process *> &visiting) {
  Process *p = s->producer;
  if (!p) {
    if (s->generation == 0) S_PANIC("GET on stream '%s' before any PUT reached it", s->name.c_str());
    return;
  }
  if (std::find(plan.begin(), plan.end(), p) != plan.end()) return;
  if (std::find(visiting.begin(), visiting.end(), p) != visiting.end())
    S_PANIC("cycle through stream '%s'", s->name.c_str());
  visiting.push_back(p);
  bool stale = !p->ever_run;
  for (int i = 0; i < 2; ++i)
    if (p->in[i]) {
      const size_t before = plan.size();
      collect(p->in[i], plan, visiting);
      if (plan.size() != before || p->in[i]->generation != p->in_gen[i]) stale = true;
    }
  visiting.pop_back();
  if (stale) plan.push_back(p);
}

struct DevPtr { // where the rows [begin, begin+count) of a stream live on this device
  uint64_t *ptr = nullptr;
  bool from_host = false; // staged by this evaluation (freed at its end unless it is an output)
};

uint64_t *find_shard(Stream *s, uint32_t gpu, uint64_t begin, uint64_t count) {
  for (Shard &sh : s->shards)
    if (sh.gpu == gpu && sh.begin == begin && sh.count == count) return sh.dev;
  return nullptr;
}

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

__global__ void k_iota64(uint64_t *p, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    p[i] = i;
}

// Runs `plan` on the rows [begin, begin+cnt) on device `gpu`; downloads the rows of `want` (if
// not null) into want_out.  One host thread per device calls this.
void run_shard(const std::vector<Process *> &plan, uint32_t gpu, uint64_t begin, uint64_t cnt,
               uint64_t total_rows, Stream *want, uint64_t *want_out, uint64_t want_stride) {
  RuntimeContext *ctx = nullptr;
  for (Process *p : plan)
    if (p->ctx) ctx = p->ctx;
  S_CHECK(cudaSetDevice((int)gpu));
  // per-device stream: the context's slot when there is one (keys and PBS scratch live there)
  static std::mutex g_mu;
  static std::map<uint32_t, void *> g_streams;
  void *stream = nullptr;
  GpuSlot *slot = nullptr;
  std::unique_lock<std::mutex> slot_lock;
  if (ctx) {
    slot = &ctx->slot(gpu);
    slot_lock = std::unique_lock<std::mutex>(slot->mutex);
    stream = slot->stream;
  } else {
    std::lock_guard<std::mutex> g(g_mu);
    void *&st = g_streams[gpu];
    if (!st) st = cuda_create_stream(gpu);
    stream = st;
  }
  cudaStream_t cs = (cudaStream_t)stream;
  cudaStream_t h2d = cs, d2h = cs;
  if (slot) {
    if (!slot->h2d_stream) slot->h2d_stream = cuda_create_stream(gpu);
    if (!slot->d2h_stream) slot->d2h_stream = cuda_create_stream(gpu);
    h2d = (cudaStream_t)slot->h2d_stream;
    d2h = (cudaStream_t)slot->d2h_stream;
  }

  // ---- device residence of every stream the plan touches
  std::map<Stream *, DevPtr> where;
  std::vector<std::pair<Stream *, bool>> uploads; // (stream, is a per-row operand) in plan order
  std::vector<uint64_t *> temporaries;
  auto alloc = [&](uint64_t words) {
    return pool_alloc(words, cs, (int)gpu);
  };
  std::vector<uint64_t> staging; // expanded constants; must outlive the async copies
  staging.reserve(4 * cnt + 16);
  for (Process *p : plan) {
    for (int i = 0; i < 2; ++i) {
      Stream *s = p->in[i];
      if (!s || where.count(s)) continue;
      const bool table = p->kind == K_BOOTSTRAP && i == 1;
      const bool scalar_operand = (p->kind == K_ADD_PLAIN || p->kind == K_MUL_CLEAR) && i == 1;
      if (uint64_t *d = (!table && !scalar_operand) ? find_shard(s, gpu, begin, cnt) : nullptr) {
        where[s] = {d, false};
        continue;
      }
      if (s->producer && std::find(plan.begin(), plan.end(), s->producer) != plan.end())
        continue; // produced inside this plan
      if (!s->host) S_PANIC("stream '%s' has no value on the host or on GPU %u", s->name.c_str(), gpu);
      DevPtr dp;
      dp.from_host = true;
      if (table) {
        // one table {1, N} -> ONE trivial GLWE; one table per sample {num, N} -> this shard's rows
        const uint64_t N = p->poly_size, glwe = (uint64_t)(p->glwe_dim + 1) * N;
        if (s->cols != N) S_PANIC("table of %lu words for polynomial size %lu", s->cols, N);
        const uint64_t luts = s->rows == 1 ? 1 : cnt;
        if (s->rows != 1 && s->rows != total_rows) S_PANIC("number of tables does not match batch");
        dp.ptr = alloc(luts * glwe);
        cuda_memset_async(dp.ptr, 0, luts * glwe * 8, stream, gpu);
        const uint64_t *src = s->host + (s->rows == 1 ? 0 : begin * N);
        S_CHECK(cudaMemcpy2DAsync(dp.ptr + (uint64_t)p->glwe_dim * N, glwe * 8, src, N * 8, N * 8,
                                  luts, cudaMemcpyHostToDevice, cs));
      } else if (scalar_operand) {
        // {1, num} per-sample scalars or {1, 1} constant (GPUDFG.cpp:1306-1308,1364-1366)
        const uint64_t num = s->rows * s->cols;
        dp.ptr = alloc(cnt);
        if (num == 1) {
          const size_t at = staging.size();
          staging.insert(staging.end(), cnt, s->host[0]);
          cuda_memcpy_async_to_gpu(dp.ptr, staging.data() + at, cnt * 8, stream, gpu);
        } else {
          if (num != total_rows) S_PANIC("%lu scalars for %lu ciphertexts", num, total_rows);
          cuda_memcpy_async_to_gpu(dp.ptr, s->host + begin, cnt * 8, stream, gpu);
        }
      } else {
        if (s->rows != total_rows) S_PANIC("stream '%s' has %lu rows, batch has %lu", s->name.c_str(), s->rows, total_rows);
        dp.ptr = alloc(cnt * s->cols); // ciphertext rows: uploaded chunk by chunk below
        uploads.push_back({s, true});
      }
      where[s] = dp;
      temporaries.push_back(dp.ptr);
    }
  }
  // outputs: one device shard per produced stream (kept as the stream's device value)
  for (Process *p : plan) {
    Stream *o = p->out;
    uint64_t cols = 0;
    switch (p->kind) {
    case K_KEYSWITCH: cols = (uint64_t)p->output_lwe_dim + 1; break;
    case K_BOOTSTRAP: cols = (uint64_t)p->glwe_dim * p->poly_size + 1; break;
    default: cols = p->in[0]->cols; break;
    }
    DevPtr dp;
    dp.ptr = alloc(cnt * cols);
    where[o] = dp;
    o->cols = cols; // every shard thread writes the same value
  }

  // ---- identity indexes + PBS scratch for the largest call
  uint64_t *iota = alloc(cnt);
  k_iota64<<<(unsigned)((cnt + 255) / 256), 256, 0, cs>>>(iota, (uint32_t)cnt);
  S_CHECK(cudaGetLastError());
  temporaries.push_back(iota);
  uint64_t *lut_zero = nullptr;
  std::map<Process *, int8_t *> own_scratch;
  for (Process *p : plan)
    if (p->kind == K_BOOTSTRAP) {
      if (!lut_zero && p->in[1]->rows == 1) {
        lut_zero = alloc(cnt);
        cuda_memset_async(lut_zero, 0, cnt * 8, stream, gpu);
        temporaries.push_back(lut_zero);
      }
      if (slot && slot->pbs_buffer && slot->pbs_k == p->glwe_dim && slot->pbs_N == p->poly_size &&
          slot->pbs_l == p->level && slot->pbs_count >= cnt)
        continue;
      if (slot && plan.size() >= 1) {
        if (slot->pbs_buffer) cleanup_cuda_programmable_bootstrap(stream, gpu, &slot->pbs_buffer);
        scratch_cuda_programmable_bootstrap_64(stream, gpu, &slot->pbs_buffer, p->glwe_dim,
                                               p->poly_size, p->level, (uint32_t)cnt, true);
        slot->pbs_k = p->glwe_dim, slot->pbs_N = p->poly_size, slot->pbs_l = p->level;
        slot->pbs_count = (uint32_t)cnt;
      } else {
        int8_t *b = nullptr;
        scratch_cuda_programmable_bootstrap_64(stream, gpu, &b, p->glwe_dim, p->poly_size,
                                               p->level, (uint32_t)cnt, true);
        own_scratch[p] = b;
      }
    }
  // keys (converted / uploaded once per (context, device), context.h:86-145)
  std::map<Process *, void *> keys;
  for (Process *p : plan) {
    if (p->kind == K_KEYSWITCH)
      keys[p] = p->ctx->get_ksk_gpu(p->level, p->input_lwe_dim, p->output_lwe_dim, gpu, stream,
                                    p->key_index);
    if (p->kind == K_BOOTSTRAP)
      keys[p] = p->ctx->get_bsk_gpu(p->input_lwe_dim, p->poly_size, p->level, p->glwe_dim, gpu,
                                    stream, p->key_index);
  }

  // ---- the chunk pipeline: upload (h2d) | whole chain (compute) | download (d2h)
  bool has_pbs = false;
  for (Process *p : plan) has_pbs |= p->kind == K_BOOTSTRAP;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, (int)gpu);
  const std::vector<uint64_t> bounds = chunk_bounds(cnt, has_pbs ? (uint64_t)sms * 8 : 1024);
  const size_t C = bounds.size() - 1;
  std::vector<cudaEvent_t> ev_in(C), ev_done(C);
  cudaEvent_t ev_setup;
  S_CHECK(cudaEventCreateWithFlags(&ev_setup, cudaEventDisableTiming));
  S_CHECK(cudaEventRecord(ev_setup, cs)); // allocations, tables, keys precede the pipeline
  if (h2d != cs) S_CHECK(cudaStreamWaitEvent(h2d, ev_setup, 0));
  for (size_t c = 0; c < C; ++c) {
    S_CHECK(cudaEventCreateWithFlags(&ev_in[c], cudaEventDisableTiming));
    S_CHECK(cudaEventCreateWithFlags(&ev_done[c], cudaEventDisableTiming));
    const uint64_t off = bounds[c], n = bounds[c + 1] - off;
    for (auto &u : uploads) {
      Stream *s = u.first;
      cuda_memcpy_async_to_gpu(where[s].ptr + off * s->cols, s->host + (begin + off) * s->cols,
                               n * s->cols * 8, h2d, gpu);
    }
    S_CHECK(cudaEventRecord(ev_in[c], h2d));
  }
  for (size_t c = 0; c < C; ++c) {
    const uint64_t off = bounds[c], n = bounds[c + 1] - off;
    if (h2d != cs) S_CHECK(cudaStreamWaitEvent(cs, ev_in[c], 0));
    for (Process *p : plan) {
      uint64_t *o = where[p->out].ptr + off * p->out->cols;
      uint64_t *a = where[p->in[0]].ptr + off * p->in[0]->cols;
      switch (p->kind) {
      case K_ADD:
        cuda_add_lwe_ciphertext_vector_64(stream, gpu, o, a, where[p->in[1]].ptr + off * p->in[1]->cols,
                                          (uint32_t)p->in[0]->cols - 1, (uint32_t)n);
        break;
      case K_ADD_PLAIN:
        cuda_add_lwe_ciphertext_vector_plaintext_vector_64(stream, gpu, o, a, where[p->in[1]].ptr + off,
                                                           (uint32_t)p->in[0]->cols - 1, (uint32_t)n);
        break;
      case K_MUL_CLEAR:
        cuda_mult_lwe_ciphertext_vector_cleartext_vector_64(stream, gpu, o, a, where[p->in[1]].ptr + off,
                                                            (uint32_t)p->in[0]->cols - 1, (uint32_t)n);
        break;
      case K_NEGATE:
        cuda_negate_lwe_ciphertext_vector_64(stream, gpu, o, a, (uint32_t)p->in[0]->cols - 1, (uint32_t)n);
        break;
      case K_KEYSWITCH:
        if (p->in[0]->cols != (uint64_t)p->input_lwe_dim + 1) S_PANIC("keyswitch input width");
        cuda_keyswitch_lwe_ciphertext_vector_64(stream, gpu, o, iota, a, iota, keys[p],
                                                p->input_lwe_dim, p->output_lwe_dim, p->base_log,
                                                p->level, (uint32_t)n);
        break;
      case K_BOOTSTRAP: {
        if (p->in[0]->cols != (uint64_t)p->input_lwe_dim + 1) S_PANIC("bootstrap input width");
        const bool one = p->in[1]->rows == 1;
        int8_t *scratch = own_scratch.count(p) ? own_scratch[p] : slot->pbs_buffer;
        cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
            stream, gpu, o, iota, where[p->in[1]].ptr, one ? lut_zero : iota + off, a, iota, keys[p],
            scratch, p->input_lwe_dim, p->glwe_dim, p->poly_size, p->base_log, p->level,
            (uint32_t)n, 1, 1);
        break;
      }
      }
    }
    S_CHECK(cudaEventRecord(ev_done[c], cs));
    if (want) {
      if (d2h != cs) S_CHECK(cudaStreamWaitEvent(d2h, ev_done[c], 0));
      const uint64_t cols = want->cols;
      if (want_stride == cols)
        S_CHECK(cudaMemcpyAsync(want_out + (begin + off) * cols, where[want].ptr + off * cols,
                                n * cols * 8, cudaMemcpyDeviceToHost, d2h));
      else
        S_CHECK(cudaMemcpy2DAsync(want_out + (begin + off) * want_stride, want_stride * 8,
                                  where[want].ptr + off * cols, cols * 8, cols * 8, n,
                                  cudaMemcpyDeviceToHost, d2h));
    }
  }
  S_CHECK(cudaStreamSynchronize(d2h));
  S_CHECK(cudaStreamSynchronize(cs));
  for (size_t c = 0; c < C; ++c) {
    cudaEventDestroy(ev_in[c]);
    cudaEventDestroy(ev_done[c]);
  }
  cudaEventDestroy(ev_setup);
  for (auto &kv : own_scratch) cleanup_cuda_programmable_bootstrap(stream, gpu, &kv.second);
  for (uint64_t *t : temporaries) cuda_drop_async(t, stream, gpu);
  // produced streams keep their device shard (a later GET or a later plan reads it there)
  static std::mutex shard_mu;
  {
    std::lock_guard<std::mutex> g(shard_mu);
    for (Process *p : plan) {
      Shard sh;
      sh.gpu = gpu, sh.begin = begin, sh.count = cnt, sh.dev = where[p->out].ptr;
      p->out->shards.push_back(sh);
    }
  }
  S_CHECK(cudaStreamSynchronize(cs));
}

void evaluate(Stream *s, uint64_t *out, uint64_t out_stride) {
  std::vector<Process *> plan, visiting;
  collect(s, plan, visiting);
  Dfg *g = s->dfg;
  if (plan.empty()) return;
  // batch size: rows of the ciphertext inputs (every ciphertext stream of a plan has the same)
  uint64_t rows = 0;
  for (Process *p : plan) {
    Stream *c0 = p->in[0];
    if (!(c0->producer && std::find(plan.begin(), plan.end(), c0->producer) != plan.end())) {
      if (rows && c0->rows != rows) S_PANIC("ciphertext streams of one graph disagree on the batch size");
      rows = c0->rows;
    }
  }
  for (Process *p : plan) { // new generation of every produced stream
    p->out->drop_host();
    p->out->drop_device();
    p->out->rows = rows;
    p->out->generation++;
    for (int i = 0; i < 2; ++i)
      if (p->in[i]) p->in_gen[i] = p->in[i]->generation;
    p->ever_run = true;
  }
  // in_gen of consumers of streams produced in this plan must see the bumped generation
  for (Process *p : plan)
    for (int i = 0; i < 2; ++i)
      if (p->in[i]) p->in_gen[i] = p->in[i]->generation;
  if (rows == 0) return;
  const uint64_t G = (uint64_t)std::max(1, std::min<int>(g ? g->num_gpus : 1, (int)std::min<uint64_t>(rows, 64)));
  RuntimeContext *ctx = nullptr;
  for (Process *p : plan)
    if (p->ctx) ctx = p->ctx;
  const uint32_t base = ctx ? (uint32_t)ctx->device_base : 0;
  // An input that lives on devices only (uploaded at PUT, or computed by an earlier plan) must be
  // split exactly as this evaluation splits the batch; if it is not (the device count changed in
  // between), bring it back to the host once and let the shards upload their rows from there.
  for (Process *p : plan)
    for (int i = 0; i < 2; ++i) {
      Stream *in = p->in[i];
      if (!in || in->host || in->shards.empty()) continue;
      if (in->producer && std::find(plan.begin(), plan.end(), in->producer) != plan.end()) continue;
      bool match = in->shards.size() == G;
      uint64_t begin = 0;
      for (uint64_t d = 0; match && d < G; ++d) {
        const uint64_t cnt = rows / G + (d < rows % G ? 1 : 0);
        match = find_shard(in, base + (uint32_t)d, begin, cnt) != nullptr;
        begin += cnt;
      }
      if (match) continue;
      in->host = static_cast<uint64_t *>(std::malloc(std::max<uint64_t>(1, in->rows * in->cols) * 8));
      in->host_owned = true;
      for (Shard &sh : in->shards) {
        S_CHECK(cudaSetDevice((int)sh.gpu));
        S_CHECK(cudaMemcpy(in->host + sh.begin * in->cols, sh.dev, sh.count * in->cols * 8,
                           cudaMemcpyDeviceToHost));
      }
      in->drop_device();
    }
  if (G == 1) {
    run_shard(plan, base, 0, rows, rows, out ? s : nullptr, out, out_stride);
    return;
  }
  std::vector<std::thread> threads;
  uint64_t begin = 0;
  for (uint64_t d = 0; d < G; ++d) {
    const uint64_t cnt = rows / G + (d < rows % G ? 1 : 0);
    threads.emplace_back(run_shard, std::cref(plan), base + (uint32_t)d, begin, cnt, rows,
                         out ? s : nullptr, out, out_stride);
    begin += cnt;
  }
  for (auto &t : threads) t.join();
}

void get_host(Stream *s, uint64_t *out, uint64_t size0, uint64_t size1, uint64_t stride0) {
  std::vector<Process *> plan, visiting;
  collect(s, plan, visiting);
  if (!plan.empty()) {
    // the producer chain runs now; `s` itself is downloaded straight into the caller's memref
    evaluate(s, out, stride0);
    if (s->rows != size0 || s->cols != size1)
      S_PANIC("GET of %lu x %lu from stream '%s' holding %lu x %lu", size0, size1, s->name.c_str(), s->rows, s->cols);
    return;
  }
  if (s->rows != size0 || s->cols != size1)
    S_PANIC("GET of %lu x %lu from stream '%s' holding %lu x %lu", size0, size1, s->name.c_str(), s->rows, s->cols);
  if (s->host) {
    for (uint64_t r = 0; r < size0; ++r) std::memcpy(out + r * stride0, s->host + r * size1, size1 * 8);
    return;
  }
  for (Shard &sh : s->shards) { // computed earlier as an intermediate: download its shards
    S_CHECK(cudaSetDevice((int)sh.gpu));
    S_CHECK(cudaMemcpy2D(out + sh.begin * stride0, stride0 * 8, sh.dev, size1 * 8, size1 * 8,
                         sh.count, cudaMemcpyDeviceToHost));
  }
}

} // namespace

extern "C" {

void *stream_emulator_init() {
  Dfg *g = new Dfg();
  int num = cuda_get_number_of_gpus();
  if (num <= 0) S_PANIC("no GPUs available on system");
  g->num_gpus = 1;
  if (const char *env = std::getenv("SDFG_NUM_GPUS")) { // GPUDFG.cpp:1745-1758
    long v = std::strtol(env, nullptr, 10);
    if (v >= 1) g->num_gpus = (int)std::min<long>(v, num);
  } else {
    g->num_gpus = num; // the reference uses every device unless told otherwise
  }
  return g;
}
void stream_emulator_run(void *dfg) { (void)dfg; } // evaluation is demand-driven (GET)
void stream_emulator_delete(void *dfg) { delete static_cast<Dfg *>(dfg); }

// test / embedding hook: override SDFG_NUM_GPUS for one graph
void concrete_b200_sdfg_set_num_gpus(void *dfg, int n) {
  int num = cuda_get_number_of_gpus();
  if (n >= 1) static_cast<Dfg *>(dfg)->num_gpus = std::min(n, num);
}

void stream_emulator_make_memref_add_lwe_ciphertexts_u64_process(void *dfg, void *sin1, void *sin2,
                                                                 void *sout) {
  make_process(dfg, K_ADD, sin1, sin2, sout);
}
void stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process(void *dfg, void *sin1,
                                                                          void *sin2, void *sout) {
  make_process(dfg, K_ADD_PLAIN, sin1, sin2, sout);
}
void stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process(void *dfg, void *sin1,
                                                                          void *sin2, void *sout) {
  make_process(dfg, K_MUL_CLEAR, sin1, sin2, sout);
}
void stream_emulator_make_memref_negate_lwe_ciphertext_u64_process(void *dfg, void *sin1,
                                                                   void *sout) {
  make_process(dfg, K_NEGATE, sin1, nullptr, sout);
}
void stream_emulator_make_memref_keyswitch_lwe_u64_process(void *dfg, void *sin1, void *sout,
                                                           uint32_t level, uint32_t base_log,
                                                           uint32_t input_lwe_dim,
                                                           uint32_t output_lwe_dim,
                                                           uint32_t output_size, uint32_t ksk_index,
                                                           void *context) {
  Process *p = make_process(dfg, K_KEYSWITCH, sin1, nullptr, sout);
  p->level = level, p->base_log = base_log, p->input_lwe_dim = input_lwe_dim;
  p->output_lwe_dim = output_lwe_dim, p->output_size = output_size, p->key_index = ksk_index;
  p->ctx = static_cast<RuntimeContext *>(context);
}
void stream_emulator_make_memref_bootstrap_lwe_u64_process(void *dfg, void *sin1, void *sin2,
                                                           void *sout, uint32_t input_lwe_dim,
                                                           uint32_t poly_size, uint32_t level,
                                                           uint32_t base_log, uint32_t glwe_dim,
                                                           uint32_t output_size, uint32_t bsk_index,
                                                           void *context) {
  Process *p = make_process(dfg, K_BOOTSTRAP, sin1, sin2, sout);
  p->input_lwe_dim = input_lwe_dim, p->poly_size = poly_size, p->level = level;
  p->base_log = base_log, p->glwe_dim = glwe_dim, p->output_size = output_size;
  p->key_index = bsk_index;
  p->ctx = static_cast<RuntimeContext *>(context);
  if (output_size != glwe_dim * poly_size + 1) S_PANIC("bootstrap output size"); // GPUDFG.cpp:1116
}

// batched forms: same processes, streams marked batched (GPUDFG.cpp:1539-1648)
static void mark(void *s, bool batched, bool is_const = false) {
  Stream *st = static_cast<Stream *>(s);
  st->ct = true;
  st->batched = batched;
  st->is_const = is_const;
}
void stream_emulator_make_memref_batched_add_lwe_ciphertexts_u64_process(void *dfg, void *sin1,
                                                                         void *sin2, void *sout) {
  mark(sin1, true), mark(sin2, true);
  stream_emulator_make_memref_add_lwe_ciphertexts_u64_process(dfg, sin1, sin2, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_add_plaintext_lwe_ciphertext_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout) {
  mark(sin1, true), mark(sin2, true);
  stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process(dfg, sin1, sin2, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_add_plaintext_cst_lwe_ciphertext_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout) {
  mark(sin1, true), mark(sin2, false, true);
  stream_emulator_make_memref_add_plaintext_lwe_ciphertext_u64_process(dfg, sin1, sin2, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_mul_cleartext_lwe_ciphertext_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout) {
  mark(sin1, true), mark(sin2, true);
  stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process(dfg, sin1, sin2, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_mul_cleartext_cst_lwe_ciphertext_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout) {
  mark(sin1, true), mark(sin2, false, true);
  stream_emulator_make_memref_mul_cleartext_lwe_ciphertext_u64_process(dfg, sin1, sin2, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_negate_lwe_ciphertext_u64_process(void *dfg, void *sin1,
                                                                           void *sout) {
  mark(sin1, true);
  stream_emulator_make_memref_negate_lwe_ciphertext_u64_process(dfg, sin1, sout);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_keyswitch_lwe_u64_process(
    void *dfg, void *sin1, void *sout, uint32_t level, uint32_t base_log, uint32_t input_lwe_dim,
    uint32_t output_lwe_dim, uint32_t output_size, uint32_t ksk_index, void *context) {
  mark(sin1, true);
  stream_emulator_make_memref_keyswitch_lwe_u64_process(dfg, sin1, sout, level, base_log,
                                                        input_lwe_dim, output_lwe_dim, output_size,
                                                        ksk_index, context);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_bootstrap_lwe_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout, uint32_t input_lwe_dim, uint32_t poly_size,
    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t output_size, uint32_t bsk_index,
    void *context) {
  mark(sin1, true), mark(sin2, false, true);
  stream_emulator_make_memref_bootstrap_lwe_u64_process(dfg, sin1, sin2, sout, input_lwe_dim,
                                                        poly_size, level, base_log, glwe_dim,
                                                        output_size, bsk_index, context);
  mark(sout, true);
}
void stream_emulator_make_memref_batched_mapped_bootstrap_lwe_u64_process(
    void *dfg, void *sin1, void *sin2, void *sout, uint32_t input_lwe_dim, uint32_t poly_size,
    uint32_t level, uint32_t base_log, uint32_t glwe_dim, uint32_t output_size, uint32_t bsk_index,
    void *context) {
  mark(sin1, true), mark(sin2, true);
  stream_emulator_make_memref_bootstrap_lwe_u64_process(dfg, sin1, sin2, sout, input_lwe_dim,
                                                        poly_size, level, base_log, glwe_dim,
                                                        output_size, bsk_index, context);
  mark(sout, true);
}

// ---- streams -------------------------------------------------------------------------------------
static void *new_stream(const char *name, stream_type stype) {
  Stream *s = new Stream();
  s->name = name ? name : "";
  s->type = stype;
  return s;
}
void *stream_emulator_make_uint64_stream(const char *name, stream_type stype) {
  return new_stream(name, stype);
}
void stream_emulator_put_uint64(void *stream, uint64_t e) {
  put_host(static_cast<Stream *>(stream), &e, &e, 0, 1, 1, 1, 1, 0);
}
uint64_t stream_emulator_get_uint64(void *stream) {
  uint64_t res = 0;
  get_host(static_cast<Stream *>(stream), &res, 1, 1, 1);
  return res;
}
void *stream_emulator_make_memref_stream(const char *name, stream_type stype) {
  return new_stream(name, stype);
}
void stream_emulator_put_memref(void *stream, uint64_t *allocated, uint64_t *aligned,
                                uint64_t offset, uint64_t size, uint64_t stride,
                                uint64_t data_ownership) {
  assert(stride == 1 && "Strided memrefs not supported");
  put_host(static_cast<Stream *>(stream), allocated, aligned, offset, 1, size, size, stride,
           data_ownership);
}
void stream_emulator_get_memref(void *stream, uint64_t *out_allocated, uint64_t *out_aligned,
                                uint64_t out_offset, uint64_t out_size, uint64_t out_stride) {
  (void)out_allocated;
  assert(out_stride == 1 && "Strided memrefs not supported");
  (void)out_stride;
  get_host(static_cast<Stream *>(stream), out_aligned + out_offset, 1, out_size, out_size);
}
void *stream_emulator_make_memref_batch_stream(const char *name, stream_type stype) {
  return new_stream(name, stype);
}
void stream_emulator_put_memref_batch(void *stream, uint64_t *allocated, uint64_t *aligned,
                                      uint64_t offset, uint64_t size0, uint64_t size1,
                                      uint64_t stride0, uint64_t stride1,
                                      uint64_t data_ownership) {
  put_host(static_cast<Stream *>(stream), allocated, aligned, offset, size0, size1, stride0,
           stride1, data_ownership);
}
void stream_emulator_get_memref_batch(void *stream, uint64_t *out_allocated, uint64_t *out_aligned,
                                      uint64_t out_offset, uint64_t out_size0, uint64_t out_size1,
                                      uint64_t out_stride0, uint64_t out_stride1) {
  (void)out_allocated;
  assert(out_stride1 == 1 && "Strided memrefs not supported");
  (void)out_stride1;
  get_host(static_cast<Stream *>(stream), out_aligned + out_offset, out_size0, out_size1,
           out_stride0);
}
// streams that were never attached to a process are owned by whoever made them
void concrete_b200_sdfg_delete_stream(void *stream) {
  Stream *s = static_cast<Stream *>(stream);
  if (!s->dfg) delete s;
}

} // extern "C"
