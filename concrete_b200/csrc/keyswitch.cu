// keyswitch.cu — K1: batched LWE keyswitch, bit-exact in wrapping u64.
//
// Replaces third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda/src/crypto/keyswitch.cuh:42-124
// (one block row per sample, KSK re-read per sample) and computes exactly what
// tfhe/src/core_crypto/algorithms/lwe_keyswitch.rs:138-228 computes:
//   out = (0,..,0,b) - sum_i sum_j digit_j(a_i) * KSK[i][j][.]
// with closest_representable (commons/math/decomposition/decomposer.rs:23-47) and the balanced
// signed digits of commons/math/decomposition/iter.rs:121-133 (level l first).
//
// Design (B200): the sum is an integer GEMM  C[s][o] = sum_kappa D[s][kappa] * KSK[kappa][o],
// kappa = i*l + j, tiled BS samples x BO outputs per CTA so that each KSK tile is read once per
// 128 samples instead of once per sample.  Digits are produced on the fly while staging the
// ciphertext tile into shared memory (no digit buffer in HBM).  Digits are stored with an offset
// d' = d + B/2 >= 0 so the 64x32 multiply-accumulate is two IMADs (IMAD.WIDE.U32 + IMAD); the
// offset is removed at the end with the per-column sum of the KSK tile rows, accumulated on the
// otherwise idle integer-add pipe.  Wrapping u64 addition is associative and commutative, so any
// summation order (incl. the split-K atomics used for small batches) is bit-exact.
#include "common.cuh"
#include "../../include/concrete_b200.h"
#include <algorithm>
#include <mutex>
#include <string>

namespace {

constexpr int BS = 128; // samples per CTA tile
constexpr int BO = 64;  // outputs per CTA tile
constexpr int TS = 8;   // samples per thread
constexpr int TO = 4;   // outputs per thread
constexpr int NT = (BS / TS) * (BO / TO); // 256 threads
constexpr int MAX_KT = 48;                // max kappa rows per tile

__device__ __forceinline__ u64 closest_representable(u64 x, u32 base_log, u32 level) {
  const u32 shift = 64 - level * base_log - 1;
  u64 r = x >> shift;
  r += 1;
  r &= ~(u64)1;
  return r << shift;
}

__device__ __forceinline__ u64 decompose_one_level(u64 &state, u64 mask, u32 base_log) {
  u64 res = state & mask;
  state >>= base_log;
  u64 carry = ((res - 1) | state) & res;
  carry >>= (base_log - 1);
  state += carry;
  return res - (carry << base_log);
}

__global__ void ks_init_out(u64 *__restrict__ out, const u64 *__restrict__ out_idx,
                            const u64 *__restrict__ in, const u64 *__restrict__ in_idx,
                            u32 n_in, u32 n_out, u32 num_samples) {
  const u32 s = blockIdx.y;
  if (s >= num_samples) return;
  u64 *row = out + out_idx[s] * (u64)(n_out + 1);
  for (u32 o = blockIdx.x * blockDim.x + threadIdx.x; o <= n_out; o += gridDim.x * blockDim.x)
    row[o] = (o == n_out) ? in[in_idx[s] * (u64)(n_in + 1) + n_in] : 0;
}

// grid: (ceil((n_out+1)/BO), ceil(num_samples/BS), ksplit)
__global__ void __launch_bounds__(NT, 2)
ks_kernel(u64 *__restrict__ out, const u64 *__restrict__ out_idx, const u64 *__restrict__ in,
          const u64 *__restrict__ in_idx, const u64 *__restrict__ ksk, u32 n_in, u32 n_out,
          u32 base_log, u32 level, u32 num_samples, u32 KI /* inputs per tile */,
          u32 tiles_per_split, bool use_atomics) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const u32 KT = KI * level;                                   // kappa rows per tile
  u32 *sD = reinterpret_cast<u32 *>(smem_raw);                 // [KT][BS]
  u64 *sK = reinterpret_cast<u64 *>(smem_raw + (size_t)KT * BS * 4); // [KT][BO]
  __shared__ u64 s_in_row[BS];                                 // row offsets of the CTA's samples

  const u32 tid = threadIdx.x;
  const u32 tx = tid % (BO / TO), ty = tid / (BO / TO);
  const u32 o0 = blockIdx.x * BO, s0 = blockIdx.y * BS;
  const u32 out_size = n_out + 1;

  for (u32 x = tid; x < BS; x += NT) {
    u32 s = s0 + x;
    s_in_row[x] = (s < num_samples) ? in_idx[s] * (u64)(n_in + 1) : ~(u64)0;
  }
  __syncthreads();

  const u32 num_tiles = (n_in + KI - 1) / KI;
  const u32 tile_begin = blockIdx.z * tiles_per_split;
  const u32 tile_end = min(num_tiles, tile_begin + tiles_per_split);

  const u64 mask = ((u64)1 << base_log) - 1;
  const u32 half_b = 1u << (base_log - 1);
  const u32 state_shift = 64 - base_log * level;

  // register staging (prefetch of the next tile)
  constexpr int D_PER_T = 4;   // BS*KI/NT  with KI <= 8
  constexpr int K_PER_T = MAX_KT * BO / NT; // 16
  u64 rA[D_PER_T];
  u64 rK[K_PER_T];

  auto prefetch = [&](u32 tile) {
    const u32 i0 = tile * KI;
#pragma unroll
    for (int e = 0; e < D_PER_T; ++e) {
      u32 lin = tid + e * NT; // over BS*KI, ii fastest
      u32 ii = lin % KI, sl = lin / KI;
      u64 v = 0;
      if (lin < BS * KI && i0 + ii < n_in && s_in_row[sl] != ~(u64)0)
        v = in[s_in_row[sl] + i0 + ii];
      rA[e] = v;
    }
#pragma unroll
    for (int e = 0; e < K_PER_T; ++e) {
      u32 lin = tid + e * NT; // over KT*BO, o fastest
      u32 oo = lin % BO, kk = lin / BO;
      u64 v = 0;
      u32 ii = kk / level;
      if (kk < KT && i0 + ii < n_in && o0 + oo < out_size)
        v = ksk[((u64)i0 * level + kk) * out_size + o0 + oo];
      rK[e] = v;
    }
  };

  auto stage = [&]() {
#pragma unroll
    for (int e = 0; e < D_PER_T; ++e) {
      u32 lin = tid + e * NT;
      if (lin < BS * KI) {
        u32 ii = lin % KI, sl = lin / KI;
        u64 state = closest_representable(rA[e], base_log, level) >> state_shift;
        for (u32 j = 0; j < level; ++j) {
          u64 d = decompose_one_level(state, mask, base_log);
          sD[(ii * level + j) * BS + sl] = (u32)d + half_b; // d' = d + B/2 in [0, B]
        }
      }
    }
#pragma unroll
    for (int e = 0; e < K_PER_T; ++e) {
      u32 lin = tid + e * NT;
      if (lin < KT * BO) sK[lin] = rK[e];
    }
  };

  u64 acc[TS][TO];
  u64 colsum[TO];
#pragma unroll
  for (int a = 0; a < TS; ++a)
#pragma unroll
    for (int b = 0; b < TO; ++b) acc[a][b] = 0;
#pragma unroll
  for (int b = 0; b < TO; ++b) colsum[b] = 0;

  if (tile_begin < tile_end) prefetch(tile_begin);
  for (u32 tile = tile_begin; tile < tile_end; ++tile) {
    stage();
    __syncthreads();
    if (tile + 1 < tile_end) prefetch(tile + 1);
#pragma unroll 4
    for (u32 kk = 0; kk < KT; ++kk) {
      // thread outputs: o = 2*tx + {0,1} + 32*c, c in {0,1}; samples ty*TS .. +7
      const uint4 d0 = *reinterpret_cast<const uint4 *>(&sD[kk * BS + ty * TS]);
      const uint4 d1 = *reinterpret_cast<const uint4 *>(&sD[kk * BS + ty * TS + 4]);
      const ulonglong2 k0 = *reinterpret_cast<const ulonglong2 *>(&sK[kk * BO + 2 * tx]);
      const ulonglong2 k1 = *reinterpret_cast<const ulonglong2 *>(&sK[kk * BO + 2 * tx + 32]);
      const u32 d[TS] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
      const u64 kv[TO] = {k0.x, k0.y, k1.x, k1.y};
#pragma unroll
      for (int b = 0; b < TO; ++b) {
        colsum[b] += kv[b];
        const u32 klo = (u32)kv[b], khi = (u32)(kv[b] >> 32);
#pragma unroll
        for (int a = 0; a < TS; ++a)
          acc[a][b] += (u64)klo * d[a] + ((u64)(khi * d[a]) << 32);
      }
    }
    __syncthreads();
  }

  // epilogue: out = body*[o==n_out] - acc + (B/2)*colsum
#pragma unroll
  for (int a = 0; a < TS; ++a) {
    const u32 sl = ty * TS + a, s = s0 + sl;
    if (s >= num_samples) continue;
    u64 *row = out + out_idx[s] * (u64)out_size;
#pragma unroll
    for (int b = 0; b < TO; ++b) {
      const u32 o = o0 + 2 * tx + (b & 1) + 32 * (b >> 1);
      if (o >= out_size) continue;
      u64 v = (u64)half_b * colsum[b] - acc[a][b];
      if (use_atomics) {
        atomicAdd(reinterpret_cast<unsigned long long *>(&row[o]), (unsigned long long)v);
      } else {
        if (o == n_out) v += in[s_in_row[sl] + n_in];
        row[o] = v;
      }
    }
  }
}

} // namespace
#include "keyswitch_umma.cuh"
#include <cstdlib>

namespace {

// Tensor-core path (keyswitch_umma.cuh).  Chosen when the batch fills the GEMM tiles and the
// exactness conditions hold: digits fit s8 (|d| <= B/2 <= 64) and the s32 accumulators cannot
// wrap (B/2 * 255 * K < 2^31).  Otherwise the IMAD kernel below runs; both are bit-exact.
// concrete_b200_set_ks_plan(): TEST hook.  0 = automatic, 1 = integer-pipe kernel, 2 = INT8
// tensor-core kernel whenever it is exact (also for batches below 512).
std::atomic<int> g_ks_plan{0};
bool ks_tensor_eligible(u32 n_in, u32 base_log, u32 level, u32 num_samples) {
  const int plan = g_ks_plan.load(std::memory_order_relaxed);
  if (plan == 1) return false;
  const bool forced = plan == 2;
  if (base_log > 7) return false;
  const u64 K = (u64)n_in * level;
  if ((K << (base_log - 1)) * 255 >= ((u64)1 << 31)) return false;
  return forced || num_samples >= 512;
}

cudaMemPool_t ks_scratch_pool() {
  static std::mutex mu;
  static cudaMemPool_t pools[64] = {};
  int dev = 0;
  CB_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> g(mu);
  cudaMemPool_t &p = pools[dev & 63];
  if (!p) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    CB_CHECK(cudaMemPoolCreate(&p, &props));
    unsigned long long keep = ~0ull;
    CB_CHECK(cudaMemPoolSetAttribute(p, cudaMemPoolAttrReleaseThreshold, &keep));
  }
  return p;
}

// tcgen05 path (keyswitch_umma.cuh): same exactness conditions, same pre-pass arithmetic, operands
// written as tile images, GEMM on the 5th-generation tensor cores with TMEM accumulators.
void ks_umma_launch(cudaStream_t stream, u64 *out, const u64 *out_idx, const u64 *in,
                    const u64 *in_idx, const u64 *ksk, u32 n_in, u32 n_out, u32 base_log,
                    u32 level, u32 num_samples) {
  const u32 K = n_in * level, out_size = n_out + 1;
  const u32 Kp = (K + umma::UKT - 1) / umma::UKT * umma::UKT;
  const u32 Np = (out_size * 8 + umma::UN - 1) / umma::UN * umma::UN;
  u32 chunk = (u32)std::min<u64>(((u64)256 << 20) / Kp, (u64)num_samples + umma::UM - 1);
  chunk = std::max<u32>(umma::UM, chunk / umma::UM * umma::UM);
  unsigned char *Bimg = nullptr;
  signed char *Dimg = nullptr;
  cudaMemPool_t pool = ks_scratch_pool();
  CB_CHECK(cudaMallocFromPoolAsync((void **)&Bimg, (size_t)Np * Kp, pool, stream));
  CB_CHECK(cudaMallocFromPoolAsync((void **)&Dimg, (size_t)chunk * Kp, pool, stream));
  {
    dim3 grid((Np / 8 + 31) / 32, Kp / 64);
    umma::ks_planes_img_kernel<<<grid, 256, 0, stream>>>(Bimg, ksk, K, out_size, Kp, Np / 8);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  int dev = 0, sms = 148;
  CB_CHECK(cudaGetDevice(&dev));
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    CB_CHECK(cudaFuncSetAttribute(umma::ks_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)umma::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  for (u32 s0 = 0; s0 < num_samples; s0 += chunk) {
    const u32 count = std::min(chunk, num_samples - s0);
    const u32 rows = (count + umma::UM - 1) / umma::UM * umma::UM;
    dim3 gd(rows, std::min<u32>((Kp / 16 + 255) / 256, 8));
    umma::ks_digits_img_kernel<<<gd, 256, 0, stream>>>(Dimg, in, in_idx, n_in, base_log, level, s0,
                                                       count, Kp);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    const u32 m_tiles = rows / umma::UM, n_tiles = Np / umma::UN;
    const u32 grid = std::min<u32>((u32)sms, m_tiles * n_tiles);
    umma::ks_umma_kernel<<<grid, umma::NT, umma::SMEM_BYTES, stream>>>(
        out, out_idx, in, in_idx, Dimg, Bimg, n_in, n_out, s0, count, Kp, m_tiles, n_tiles);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  CB_CHECK(cudaFreeAsync(Dimg, stream));
  CB_CHECK(cudaFreeAsync(Bimg, stream));
}

} // namespace

void cb_keyswitch_launch(cudaStream_t stream, u64 *out, const u64 *out_idx, const u64 *in,
                         const u64 *in_idx, const u64 *ksk, u32 n_in, u32 n_out, u32 base_log,
                         u32 level, u32 num_samples) {
  if (num_samples == 0) return;
  if (base_log == 0 || level == 0 || base_log * level >= 64 || base_log > 31)
    CB_PANIC("Cuda error (keyswitch): unsupported decomposition base_log=%u level=%u", base_log,
             level);
  if (level > MAX_KT)
    CB_PANIC("Cuda error (keyswitch): level_count %u > %d unsupported", level, MAX_KT);
  if (ks_tensor_eligible(n_in, base_log, level, num_samples)) {
    ks_umma_launch(stream, out, out_idx, in, in_idx, ksk, n_in, n_out, base_log, level,
                   num_samples);
    return;
  }
  u32 KI = MAX_KT / level;
  if (KI > 8) KI = 8;
  if (KI < 1) KI = 1;
  const u32 KT = KI * level;
  const u32 num_tiles = (n_in + KI - 1) / KI;
  dim3 grid((n_out + 1 + BO - 1) / BO, (num_samples + BS - 1) / BS, 1);
  // split-K so that small batches still fill the 148 SMs (2 CTAs/SM resident)
  u32 ksplit = 1;
  const u32 base_ctas = grid.x * grid.y;
  if (base_ctas < 296) {
    ksplit = (296 + base_ctas - 1) / base_ctas;
    if (ksplit > num_tiles) ksplit = num_tiles;
    if (ksplit < 1) ksplit = 1;
  }
  const u32 tiles_per_split = (num_tiles + ksplit - 1) / ksplit;
  ksplit = (num_tiles + tiles_per_split - 1) / tiles_per_split;
  grid.z = ksplit;
  const bool use_atomics = ksplit > 1;
  if (use_atomics) {
    dim3 g((n_out + 1 + 255) / 256, num_samples);
    ks_init_out<<<g, 256, 0, stream>>>(out, out_idx, in, in_idx, n_in, n_out, num_samples);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  const size_t smem = (size_t)KT * BS * 4 + (size_t)KT * BO * 8;
  // KT = 48 (e.g. l_ks = 6, 12, 16, 24) needs 48 KiB dynamic + 1 KiB static: above the default
  // 48 KiB per-CTA limit, so the kernel always opts in to its maximum (MAX_KT rows)
  CB_CHECK(cudaFuncSetAttribute(ks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)((size_t)MAX_KT * BS * 4 + (size_t)MAX_KT * BO * 8)));
  ks_kernel<<<grid, NT, smem, stream>>>(out, out_idx, in, in_idx, ksk, n_in, n_out, base_log,
                                        level, num_samples, KI, tiles_per_split, use_atomics);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

extern "C" void concrete_b200_set_ks_plan(int plan) {
  g_ks_plan.store(plan, std::memory_order_relaxed);
}

extern "C" void cuda_keyswitch_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lwe_array_in, void *lwe_input_indexes, void *ksk, uint32_t lwe_dimension_in,
    uint32_t lwe_dimension_out, uint32_t base_log, uint32_t level_count,
    uint32_t num_samples) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_keyswitch_launch(as_stream(stream), (u64 *)lwe_array_out, (const u64 *)lwe_output_indexes,
                      (const u64 *)lwe_array_in, (const u64 *)lwe_input_indexes,
                      (const u64 *)ksk, lwe_dimension_in, lwe_dimension_out, base_log,
                      level_count, num_samples);
}

// ---- no-allocation variant ----------------------------------------------------------------------
// concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc: the tcgen05 keyswitch above with its
// two operand images (key byte planes + digits) in a fixed device array instead of the stream-
// ordered pool, for callers that may not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
// Nothing on this path calls cudaMalloc*/cudaFree* or an allocating library routine.
//
// The array is held by every GPU that loads the library until the process exits.  72 MiB is the
// smallest round size that holds the headline parameter set (P2048: key image 6144 x 10240 B =
// 60 MiB, the rest is the digit image, >= 128 rows per chunk).  Every call shares it: the pre-
// passes rewrite every byte of the region a call uses (padding rows and columns included), so
// nothing has to start at zero, but two calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) unsigned char g_ks_static_scratch[CONCRETE_B200_KS_NO_ALLOC_SCRATCH_BYTES];

namespace {

// In host code `g_ks_static_scratch` is only a handle for the runtime; the device address comes
// from cudaGetSymbolAddress, once per device, with that device current.
int ks_static_scratch(int dev, unsigned char **ptr, char *err, size_t err_cap) {
  static std::mutex mu;
  static unsigned char *cached[64] = {};
  if (dev < 0 || dev >= 64) {
    std::snprintf(err, err_cap, "keyswitch_no_alloc: gpu_index %d out of range [0, 64)", dev);
    return CONCRETE_B200_NO_ALLOC_BAD_ARGUMENT;
  }
  std::lock_guard<std::mutex> g(mu);
  if (!cached[dev]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_ks_static_scratch);
    if (e != cudaSuccess || !p) {
      std::snprintf(err, err_cap, "keyswitch_no_alloc: cudaGetSymbolAddress failed on gpu %d: %s",
                    dev, cudaGetErrorString(e));
      return CONCRETE_B200_NO_ALLOC_CUDA_ERROR;
    }
    e = cudaFuncSetAttribute(umma::ks_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)umma::SMEM_BYTES);
    if (e != cudaSuccess) {
      std::snprintf(err, err_cap, "keyswitch_no_alloc: cudaFuncSetAttribute failed on gpu %d: %s",
                    dev, cudaGetErrorString(e));
      return CONCRETE_B200_NO_ALLOC_CUDA_ERROR;
    }
    cached[dev] = static_cast<unsigned char *>(p);
  }
  *ptr = cached[dev];
  return 0;
}

} // namespace

extern "C" int concrete_b200_keyswitch_lwe_ciphertext_vector_64_no_alloc(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lwe_array_in, void *lwe_input_indexes, void *ksk, uint32_t lwe_dimension_in,
    uint32_t lwe_dimension_out, uint32_t base_log, uint32_t level_count, uint32_t num_samples,
    char *error_message, uint64_t error_capacity) {
  char local[8];
  char *err = error_message && error_capacity ? error_message : local;
  const size_t cap = error_message && error_capacity ? (size_t)error_capacity : sizeof(local);
  err[0] = 0;
  if (num_samples == 0) return 0;
  if (!lwe_array_out || !lwe_output_indexes || !lwe_array_in || !lwe_input_indexes || !ksk) {
    std::snprintf(err, cap, "keyswitch_no_alloc: null device pointer argument");
    return CONCRETE_B200_NO_ALLOC_BAD_ARGUMENT;
  }
  const u32 n_in = lwe_dimension_in, n_out = lwe_dimension_out, level = level_count;
  if (n_in == 0 || base_log == 0 || level == 0 || base_log * level >= 64 || base_log > 7 ||
      (((u64)n_in * level) << (base_log - 1)) * 255 >= ((u64)1 << 31)) {
    std::snprintf(err, cap,
                  "keyswitch_no_alloc: decomposition base_log=%u level=%u over n_in=%u is outside "
                  "the exact INT8 range of the tensor-core kernel (base_log <= 7 and "
                  "n_in*level*2^(base_log-1)*255 < 2^31); use "
                  "cuda_keyswitch_lwe_ciphertext_vector_64",
                  base_log, level, n_in);
    return CONCRETE_B200_NO_ALLOC_UNSUPPORTED;
  }
  const u64 K = (u64)n_in * level, out_size = (u64)n_out + 1;
  const u64 Kp = (K + umma::UKT - 1) / umma::UKT * umma::UKT;
  const u64 Np = (out_size * 8 + umma::UN - 1) / umma::UN * umma::UN;
  const u64 key_img = (Np * Kp + 255) / 256 * 256, total = CONCRETE_B200_KS_NO_ALLOC_SCRATCH_BYTES;
  if (key_img + (u64)umma::UM * Kp > total) {
    std::snprintf(err, cap,
                  "keyswitch_no_alloc: n_in=%u level=%u n_out=%u needs %llu bytes of scratch (key "
                  "image %llu x %llu + one 128-row digit tile), the fixed array holds %llu",
                  n_in, level, n_out, (unsigned long long)(key_img + (u64)umma::UM * Kp),
                  (unsigned long long)Np, (unsigned long long)Kp, (unsigned long long)total);
    return CONCRETE_B200_NO_ALLOC_TOO_LARGE;
  }
  cudaError_t e = cudaSetDevice((int)gpu_index);
  if (e != cudaSuccess) {
    std::snprintf(err, cap, "keyswitch_no_alloc: cudaSetDevice(%u): %s", gpu_index,
                  cudaGetErrorString(e));
    return CONCRETE_B200_NO_ALLOC_CUDA_ERROR;
  }
  unsigned char *scratch = nullptr;
  if (int rc = ks_static_scratch((int)gpu_index, &scratch, err, cap)) return rc;
  int sms = 148;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, (int)gpu_index) != cudaSuccess)
    sms = 148;

  cudaStream_t st = as_stream(stream);
  unsigned char *Bimg = scratch;
  signed char *Dimg = reinterpret_cast<signed char *>(scratch + key_img);
  u32 chunk = (u32)std::min<u64>((total - key_img) / Kp, (u64)num_samples + umma::UM - 1);
  chunk = chunk / umma::UM * umma::UM; // >= UM by the size check above
  const u64 *in = (const u64 *)lwe_array_in, *in_idx = (const u64 *)lwe_input_indexes;
  u64 *out = (u64 *)lwe_array_out;
  const u64 *out_idx = (const u64 *)lwe_output_indexes;
  auto launched = [&](const char *what) -> int {
    cudaError_t le = cudaGetLastError();
    if (le == cudaSuccess) {
      CB_LAUNCHED();
      return 0;
    }
    std::snprintf(err, cap, "keyswitch_no_alloc: launch of %s failed: %s", what,
                  cudaGetErrorString(le));
    return CONCRETE_B200_NO_ALLOC_CUDA_ERROR;
  };
  {
    dim3 grid((u32)((Np / 8 + 31) / 32), (u32)(Kp / 64));
    umma::ks_planes_img_kernel<<<grid, 256, 0, st>>>(Bimg, (const u64 *)ksk, (u32)K, (u32)out_size,
                                                     (u32)Kp, (u32)(Np / 8));
    if (int rc = launched("ks_planes_img_kernel")) return rc;
  }
  for (u32 s0 = 0; s0 < num_samples; s0 += chunk) {
    const u32 count = std::min(chunk, num_samples - s0);
    const u32 rows = (count + umma::UM - 1) / umma::UM * umma::UM;
    dim3 gd(rows, std::min<u32>((u32)((Kp / 16 + 255) / 256), 8));
    umma::ks_digits_img_kernel<<<gd, 256, 0, st>>>(Dimg, in, in_idx, n_in, base_log, level, s0,
                                                   count, (u32)Kp);
    if (int rc = launched("ks_digits_img_kernel")) return rc;
    const u32 m_tiles = rows / umma::UM, n_tiles = (u32)(Np / umma::UN);
    const u32 grid = std::min<u32>((u32)sms, m_tiles * n_tiles);
    umma::ks_umma_kernel<<<grid, umma::NT, umma::SMEM_BYTES, st>>>(
        out, out_idx, in, in_idx, Dimg, Bimg, n_in, n_out, s0, count, (u32)Kp, m_tiles, n_tiles);
    if (int rc = launched("ks_umma_kernel")) return rc;
  }
  return 0;
}
