// seeded.cu — K6: seeded-key decompression on the device.
//
// Concrete ships bootstrapping and keyswitching keys compressed: only the bodies plus a 128-bit
// seed; the masks are the output of the reference's AES-128-CTR generator
// (compilers/concrete-compiler/compiler/lib/Common/Keys.cpp:193-213 decompresses on the host with
// concrete_cpu_decompress_seeded_lwe_{bootstrap,keyswitch}_key_u64,
// backends/concrete-cpu/implementation/src/c_api/bootstrap.rs:169-237, then the runtime uploads
// the FULL key, context.h:101-109,134-139).  Here the masks are regenerated where they are
// consumed: the host sends n*l*(k+1)*N body words instead of n*l*(k+1)^2*N, every thread
// encrypts its own counters and writes the mask words straight into the key, and the existing
// in-place Fourier conversion (K4) follows on the same stream.
//
// Stream definition restated from third_party/tfhe-rs/concrete-csprng/src/generators:
//   table byte (a, b) = AES128_seed(a)[b]  (implem/soft/block_cipher.rs:16-36: key and counter
//   are u128::to_ne_bytes, little endian); a fresh generator starts at table byte (0, 1)
//   (aes_ctr/generic.rs:60-72, index.rs:28-31); forks hand out consecutive ranges
//   (generic.rs:124-170) sized by the children's mask bytes, so mask word w of a key (key order:
//   GGSW, level, row, polynomial, coefficient / keyswitch block, level, element) is table bytes
//   1 + 8w .. 8 + 8w, little endian (tfhe core_crypto/commons/math/random/uniform.rs:10-18).
// AES-128 is FIPS-197; one 1 KiB T-table (SubBytes + MixColumns of one byte) in shared memory,
// the other three by byte rotation.  Bit-exact against oracle/csprng_oracle.cpp.
#include "common.cuh"
#include "../../include/concrete_b200.h"
#include <cstring>
#include <mutex>
#include <set>

void cb_bsk_convert(cudaStream_t stream, int gpu, void *dest, const void *src_host, u32 n, u32 k,
                    u32 level, u32 N, bool src_on_device);

namespace {

struct AesTables {
  u32 te0[256]; // (2*S[x], S[x], S[x], 3*S[x]) from the most significant byte down
  u32 sbox[256];
};
__device__ AesTables g_aes_tables;

struct RoundKeys {
  u32 w[44]; // big-endian words of the 11 round keys
};

inline uint8_t xtime(uint8_t x) { return (uint8_t)((x << 1) ^ ((x & 0x80) ? 0x1B : 0)); }

// S-box from its definition (inverse in GF(2^8), then the affine map), not from a typed-in table
void host_sbox(uint8_t (&sbox)[256]) {
  uint8_t p = 1, q = 1;
  do {
    p = (uint8_t)(p ^ (p << 1) ^ ((p & 0x80) ? 0x1B : 0));
    q ^= (uint8_t)(q << 1);
    q ^= (uint8_t)(q << 2);
    q ^= (uint8_t)(q << 4);
    if (q & 0x80) q ^= 0x09;
    const uint8_t x = (uint8_t)(q ^ ((q << 1) | (q >> 7)) ^ ((q << 2) | (q >> 6)) ^
                                ((q << 3) | (q >> 5)) ^ ((q << 4) | (q >> 4)));
    sbox[p] = (uint8_t)(x ^ 0x63);
  } while (p != 1);
  sbox[0] = 0x63;
}

void ensure_tables(int gpu) {
  static std::mutex mu;
  static std::set<int> done;
  std::lock_guard<std::mutex> g(mu);
  if (done.count(gpu)) return;
  uint8_t sbox[256];
  host_sbox(sbox);
  AesTables t;
  for (int x = 0; x < 256; ++x) {
    const uint8_t s = sbox[x], s2 = xtime(s), s3 = (uint8_t)(s2 ^ s);
    t.te0[x] = ((u32)s2 << 24) | ((u32)s << 16) | ((u32)s << 8) | (u32)s3;
    t.sbox[x] = s;
  }
  CB_CHECK(cudaMemcpyToSymbol(g_aes_tables, &t, sizeof(t)));
  done.insert(gpu);
}

RoundKeys expand_key(u64 seed_lo, u64 seed_hi) {
  uint8_t sbox[256];
  host_sbox(sbox);
  uint8_t rk[11][16];
  for (int i = 0; i < 8; ++i) {
    rk[0][i] = (uint8_t)(seed_lo >> (8 * i));
    rk[0][8 + i] = (uint8_t)(seed_hi >> (8 * i));
  }
  uint8_t rcon = 1;
  for (int r = 1; r <= 10; ++r) {
    const uint8_t *prev = rk[r - 1];
    uint8_t t[4] = {sbox[prev[13]], sbox[prev[14]], sbox[prev[15]], sbox[prev[12]]};
    t[0] ^= rcon;
    rcon = xtime(rcon);
    for (int c = 0; c < 4; ++c)
      for (int b = 0; b < 4; ++b)
        rk[r][4 * c + b] = (uint8_t)(prev[4 * c + b] ^ (c == 0 ? t[b] : rk[r][4 * (c - 1) + b]));
  }
  RoundKeys out;
  for (int r = 0; r < 11; ++r)
    for (int c = 0; c < 4; ++c)
      out.w[4 * r + c] = ((u32)rk[r][4 * c] << 24) | ((u32)rk[r][4 * c + 1] << 16) |
                         ((u32)rk[r][4 * c + 2] << 8) | (u32)rk[r][4 * c + 3];
  return out;
}

constexpr int ST = 256;      // threads per CTA
constexpr int WPT = 8;       // mask words per thread = 4 blocks + the first byte of a fifth

__device__ __forceinline__ u32 ror8(u32 x) { return __funnelshift_r(x, x, 8); }
__device__ __forceinline__ u32 ror16(u32 x) { return __funnelshift_r(x, x, 16); }
__device__ __forceinline__ u32 ror24(u32 x) { return __funnelshift_r(x, x, 24); }

// AES128(counter), counter < 2^64 as a little-endian 16-byte block; output as four big-endian
// column words (byte b of the block = byte 3 - (b & 3) of word b >> 2)
__device__ __forceinline__ void aes_block(u64 ctr, const RoundKeys &rk, const u32 *te0,
                                          const u32 *sb, u32 (&o)[4]) {
  // plaintext bytes 0..7 = ctr little endian, 8..15 = 0
  u32 s0 = __byte_perm((u32)ctr, 0, 0x0123) ^ rk.w[0];
  u32 s1 = __byte_perm((u32)(ctr >> 32), 0, 0x0123) ^ rk.w[1];
  u32 s2 = rk.w[2], s3 = rk.w[3];
#pragma unroll
  for (int r = 1; r < 10; ++r) {
    const u32 t0 = te0[s0 >> 24] ^ ror8(te0[(s1 >> 16) & 0xff]) ^ ror16(te0[(s2 >> 8) & 0xff]) ^
                   ror24(te0[s3 & 0xff]) ^ rk.w[4 * r + 0];
    const u32 t1 = te0[s1 >> 24] ^ ror8(te0[(s2 >> 16) & 0xff]) ^ ror16(te0[(s3 >> 8) & 0xff]) ^
                   ror24(te0[s0 & 0xff]) ^ rk.w[4 * r + 1];
    const u32 t2 = te0[s2 >> 24] ^ ror8(te0[(s3 >> 16) & 0xff]) ^ ror16(te0[(s0 >> 8) & 0xff]) ^
                   ror24(te0[s1 & 0xff]) ^ rk.w[4 * r + 2];
    const u32 t3 = te0[s3 >> 24] ^ ror8(te0[(s0 >> 16) & 0xff]) ^ ror16(te0[(s1 >> 8) & 0xff]) ^
                   ror24(te0[s2 & 0xff]) ^ rk.w[4 * r + 3];
    s0 = t0, s1 = t1, s2 = t2, s3 = t3;
  }
  o[0] = (sb[s0 >> 24] << 24) ^ (sb[(s1 >> 16) & 0xff] << 16) ^ (sb[(s2 >> 8) & 0xff] << 8) ^
         sb[s3 & 0xff] ^ rk.w[40];
  o[1] = (sb[s1 >> 24] << 24) ^ (sb[(s2 >> 16) & 0xff] << 16) ^ (sb[(s3 >> 8) & 0xff] << 8) ^
         sb[s0 & 0xff] ^ rk.w[41];
  o[2] = (sb[s2 >> 24] << 24) ^ (sb[(s3 >> 16) & 0xff] << 16) ^ (sb[(s0 >> 8) & 0xff] << 8) ^
         sb[s1 & 0xff] ^ rk.w[42];
  o[3] = (sb[s3 >> 24] << 24) ^ (sb[(s0 >> 16) & 0xff] << 16) ^ (sb[(s1 >> 8) & 0xff] << 8) ^
         sb[s2 & 0xff] ^ rk.w[43];
}

// Mask word w of the key -> out[(w / mask_len) * row_stride + w % mask_len].  A thread owns WPT
// consecutive words = table bytes 64g + 1 .. 64g + 64: blocks 4g .. 4g + 3 and byte 0 of 4g + 4.
__global__ void __launch_bounds__(ST) mask_fill_kernel(u64 *__restrict__ out, u64 total_words,
                                                       u64 mask_len, u64 row_stride,
                                                       const __grid_constant__ RoundKeys rk) {
  __shared__ u32 te0[256], sb[256];
  for (int i = threadIdx.x; i < 256; i += ST) {
    te0[i] = g_aes_tables.te0[i];
    sb[i] = g_aes_tables.sbox[i];
  }
  __syncthreads();
  const u64 groups = (total_words + WPT - 1) / WPT;
  for (u64 g = blockIdx.x * (u64)ST + threadIdx.x; g < groups; g += (u64)gridDim.x * ST) {
    // bytes[0..79] = blocks 4g .. 4g + 4 in table order; as little-endian u32 words le[0..19]
    u32 le[20];
#pragma unroll
    for (int b = 0; b < 5; ++b) {
      u32 o[4];
      aes_block(4 * g + b, rk, te0, sb, o);
#pragma unroll
      for (int c = 0; c < 4; ++c) le[4 * b + c] = __byte_perm(o[c], 0, 0x0123);
    }
    const u64 w0 = g * WPT;
    u64 row = w0 / mask_len, pos = w0 - row * mask_len;
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
      if (w0 + j < total_words) {
        // word j = table bytes 8j + 1 .. 8j + 8: a 64-bit window at byte offset 1 of le[2j..2j+2]
        const u32 lo = __funnelshift_r(le[2 * j], le[2 * j + 1], 8);
        const u32 hi = __funnelshift_r(le[2 * j + 1], le[2 * j + 2], 8);
        out[row * row_stride + pos] = ((u64)hi << 32) | lo;
        if (++pos == mask_len) {
          pos = 0;
          ++row;
        }
      }
    }
  }
}

// bodies: out[row * row_stride + mask_len + c] = seeded[row * body_len + c]
__global__ void body_scatter_kernel(u64 *__restrict__ out, const u64 *__restrict__ seeded,
                                    u64 total, u64 body_len, u64 mask_len, u64 row_stride) {
  for (u64 i = blockIdx.x * (u64)blockDim.x + threadIdx.x; i < total;
       i += (u64)gridDim.x * blockDim.x) {
    const u64 row = i / body_len, c = i - row * body_len;
    out[row * row_stride + mask_len + c] = seeded[i];
  }
}

bool on_device(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) return attr.type == cudaMemoryTypeDevice;
  (void)cudaGetLastError();
  return false;
}

// rows x (mask_len mask words from the generator + body_len body words from `seeded`)
void decompress_rows(cudaStream_t stream, int gpu, u64 *dest, const u64 *seeded, u64 rows,
                     u64 mask_len, u64 body_len, u64 seed_lo, u64 seed_hi) {
  if (rows == 0) return;
  ensure_tables(gpu);
  const u64 row_stride = mask_len + body_len, total_words = rows * mask_len;
  int sms = 148;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  if (total_words) {
    const RoundKeys rk = expand_key(seed_lo, seed_hi);
    const u64 groups = (total_words + WPT - 1) / WPT;
    const u32 grid = (u32)std::min<u64>((groups + ST - 1) / ST, (u64)sms * 8);
    mask_fill_kernel<<<grid, ST, 0, stream>>>(dest, total_words, mask_len, row_stride, rk);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  const u64 bodies = rows * body_len;
  const u64 *src = seeded;
  u64 *staged = nullptr;
  if (!on_device(seeded)) { // host bodies: one contiguous upload, scattered on the device
    CB_CHECK(cudaMallocAsync((void **)&staged, bodies * sizeof(u64), stream));
    CB_CHECK(cudaMemcpyAsync(staged, seeded, bodies * sizeof(u64), cudaMemcpyHostToDevice, stream));
    src = staged;
  }
  const u32 grid = (u32)std::min<u64>((bodies + 255) / 256, (u64)sms * 16);
  body_scatter_kernel<<<grid, 256, 0, stream>>>(dest, src, bodies, body_len, mask_len, row_stride);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
  if (staged) CB_CHECK(cudaFreeAsync(staged, stream));
}

} // namespace

extern "C" {

void cuda_decompress_seeded_lwe_bootstrap_key_64(void *stream, uint32_t gpu_index, void *dest,
                                                 const void *seeded_bsk, uint32_t input_lwe_dim,
                                                 uint32_t glwe_dim, uint32_t level_count,
                                                 uint32_t polynomial_size, uint64_t seed_lo,
                                                 uint64_t seed_hi) {
  CB_CHECK(cudaSetDevice(gpu_index));
  const u64 rows = (u64)input_lwe_dim * level_count * (glwe_dim + 1);
  decompress_rows(as_stream(stream), gpu_index, (u64 *)dest, (const u64 *)seeded_bsk, rows,
                  (u64)glwe_dim * polynomial_size, polynomial_size, seed_lo, seed_hi);
}

void cuda_decompress_seeded_lwe_keyswitch_key_64(void *stream, uint32_t gpu_index, void *dest,
                                                 const void *seeded_ksk, uint32_t lwe_dim_in,
                                                 uint32_t lwe_dim_out, uint32_t level_count,
                                                 uint64_t seed_lo, uint64_t seed_hi) {
  CB_CHECK(cudaSetDevice(gpu_index));
  decompress_rows(as_stream(stream), gpu_index, (u64 *)dest, (const u64 *)seeded_ksk,
                  (u64)lwe_dim_in * level_count, lwe_dim_out, 1, seed_lo, seed_hi);
}

// decompress into `dest`, then the in-place Fourier conversion: `dest` ends up exactly as after
// cuda_convert_lwe_programmable_bootstrap_key_64 on the decompressed key
void cuda_convert_seeded_lwe_programmable_bootstrap_key_64(
    void *stream, uint32_t gpu_index, void *dest, const void *seeded_bsk, uint32_t input_lwe_dim,
    uint32_t glwe_dim, uint32_t level_count, uint32_t polynomial_size, uint64_t seed_lo,
    uint64_t seed_hi) {
  cuda_decompress_seeded_lwe_bootstrap_key_64(stream, gpu_index, dest, seeded_bsk, input_lwe_dim,
                                              glwe_dim, level_count, polynomial_size, seed_lo,
                                              seed_hi);
  cb_bsk_convert(as_stream(stream), gpu_index, dest, dest, input_lwe_dim, glwe_dim, level_count,
                 polynomial_size, true);
}

} // extern "C"
