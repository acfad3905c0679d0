// csprng_oracle.cpp — CPU restatement of the reference's seeded-key decompression.
//
// TEST INFRASTRUCTURE ONLY (linked into liboracle.so; see the header of tfhe_oracle.cpp for the
// rule: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may call it).
//
// What it restates (paths relative to /root/reference):
//   * AES-128 block encryption, FIPS-197 — the reference delegates to the `aes` crate
//     (third_party/tfhe-rs/concrete-csprng/src/generators/implem/soft/block_cipher.rs:16-36:
//     key = seed.to_ne_bytes(), plaintext = counter(u128).to_ne_bytes(), little endian on x86);
//     pinned by the FIPS-197 appendix C.1 vector the reference's own test holds
//     (block_cipher.rs:90-112) and by appendix B.
//   * the byte table of the AES-CTR generator: byte (a, b) = AES_k(a)[b]; a fresh generator
//     starts at TableIndex::SECOND = (0, 1) (aes_ctr/generic.rs:60-72, index.rs:28-31,
//     states.rs:35-50), so stream byte q is table byte 1 + q; forking hands out CONSECUTIVE
//     ranges of the parent's table (generic.rs:124-170), sized exactly by the mask bytes of the
//     children, so the masks of a whole key are one contiguous run of the stream in key order;
//   * a uniform u64 mask element = 8 consecutive stream bytes, little endian
//     (third_party/tfhe-rs/tfhe/src/core_crypto/commons/math/random/uniform.rs:10-18);
//   * decompress_seeded_lwe_bootstrap_key -> ggsw list -> ggsw (levels) -> rows -> glwe:
//     mask from the generator, body copied (core_crypto/algorithms/
//     seeded_lwe_bootstrap_key_decompression.rs:36-62, seeded_ggsw_ciphertext_list_decompression.rs:
//     8-41, seeded_ggsw_ciphertext_decompression.rs:8-52, seeded_glwe_ciphertext_decompression.rs:
//     6-45); C entry backends/concrete-cpu/implementation/src/c_api/bootstrap.rs:169-237;
//   * decompress_seeded_lwe_keyswitch_key -> lwe ciphertext list (seeded_lwe_keyswitch_key_
//     decompression.rs:6-25, seeded_lwe_ciphertext_list_decompression.rs:9-60); C entry
//     c_api/keyswitch.rs (concrete_cpu_decompress_seeded_lwe_keyswitch_key_u64).
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

typedef uint8_t u8;
typedef uint64_t u64;

namespace {

// FIPS-197 S-box, generated (multiplicative inverse in GF(2^8) + affine map) rather than typed in
struct Tables {
  u8 sbox[256];
  Tables() {
    u8 p = 1, q = 1;
    do { // p runs over all non-zero elements (generator 3), q = p^-1
      p = (u8)(p ^ (p << 1) ^ ((p & 0x80) ? 0x1B : 0));
      q ^= (u8)(q << 1);
      q ^= (u8)(q << 2);
      q ^= (u8)(q << 4);
      if (q & 0x80) q ^= 0x09;
      u8 x = (u8)(q ^ ((q << 1) | (q >> 7)) ^ ((q << 2) | (q >> 6)) ^ ((q << 3) | (q >> 5)) ^
                  ((q << 4) | (q >> 4)));
      sbox[p] = (u8)(x ^ 0x63);
    } while (p != 1);
    sbox[0] = 0x63;
  }
};
const Tables T;

inline u8 xtime(u8 x) { return (u8)((x << 1) ^ ((x & 0x80) ? 0x1B : 0)); }

struct Aes128 {
  u8 rk[11][16];
  explicit Aes128(const u8 key[16]) {
    std::memcpy(rk[0], key, 16);
    u8 rcon = 1;
    for (int r = 1; r <= 10; ++r) {
      const u8 *prev = rk[r - 1];
      u8 t[4] = {T.sbox[prev[13]], T.sbox[prev[14]], T.sbox[prev[15]], T.sbox[prev[12]]};
      t[0] ^= rcon;
      rcon = xtime(rcon);
      for (int c = 0; c < 4; ++c) {
        for (int b = 0; b < 4; ++b) {
          const u8 left = (c == 0) ? t[b] : rk[r][4 * (c - 1) + b];
          rk[r][4 * c + b] = (u8)(prev[4 * c + b] ^ left);
        }
      }
    }
  }
  void encrypt(const u8 in[16], u8 out[16]) const {
    u8 s[16];
    for (int i = 0; i < 16; ++i) s[i] = (u8)(in[i] ^ rk[0][i]);
    for (int r = 1; r <= 10; ++r) {
      u8 t[16];
      for (int c = 0; c < 4; ++c) // SubBytes + ShiftRows (state is column-major: s[4c + row])
        for (int row = 0; row < 4; ++row) t[4 * c + row] = T.sbox[s[4 * ((c + row) & 3) + row]];
      if (r < 10) {
        for (int c = 0; c < 4; ++c) { // MixColumns
          const u8 a0 = t[4 * c], a1 = t[4 * c + 1], a2 = t[4 * c + 2], a3 = t[4 * c + 3];
          s[4 * c + 0] = (u8)(xtime(a0) ^ (xtime(a1) ^ a1) ^ a2 ^ a3);
          s[4 * c + 1] = (u8)(a0 ^ xtime(a1) ^ (xtime(a2) ^ a2) ^ a3);
          s[4 * c + 2] = (u8)(a0 ^ a1 ^ xtime(a2) ^ (xtime(a3) ^ a3));
          s[4 * c + 3] = (u8)((xtime(a0) ^ a0) ^ a1 ^ a2 ^ xtime(a3));
        }
      } else {
        std::memcpy(s, t, 16);
      }
      for (int i = 0; i < 16; ++i) s[i] ^= rk[r][i];
    }
    std::memcpy(out, s, 16);
  }
};

// The generator of one seed: random access into its byte table.
struct Stream {
  Aes128 aes;
  static void seed_bytes(u8 key[16], u64 lo, u64 hi) {
    for (int i = 0; i < 8; ++i) {
      key[i] = (u8)(lo >> (8 * i));     // u128::to_ne_bytes, little endian
      key[8 + i] = (u8)(hi >> (8 * i));
    }
  }
  static Aes128 make(u64 lo, u64 hi) {
    u8 key[16];
    seed_bytes(key, lo, hi);
    return Aes128(key);
  }
  Stream(u64 lo, u64 hi) : aes(make(lo, hi)) {}
  void block(u64 aes_index, u8 out[16]) const { // counters beyond 2^64 are never reached here
    u8 ctr[16] = {0};
    for (int i = 0; i < 8; ++i) ctr[i] = (u8)(aes_index >> (8 * i));
    aes.encrypt(ctr, out);
  }
  // stream bytes [first, first + count): table bytes 1 + first ...
  void bytes(u8 *out, u64 first, u64 count) const {
    u64 idx = first + 1;
    u8 blk[16];
    u64 have = ~(u64)0;
    for (u64 i = 0; i < count; ++i, ++idx) {
      if ((idx >> 4) != have) {
        have = idx >> 4;
        block(have, blk);
      }
      out[i] = blk[idx & 15];
    }
  }
  void words(u64 *out, u64 first_word, u64 count) const {
    std::vector<u8> b(8 * 1024);
    for (u64 done = 0; done < count;) {
      const u64 n = std::min<u64>(1024, count - done);
      bytes(b.data(), 8 * (first_word + done), 8 * n);
      for (u64 i = 0; i < n; ++i) {
        u64 v = 0;
        for (int j = 0; j < 8; ++j) v |= (u64)b[8 * i + j] << (8 * j); // from_le_bytes
        out[done + i] = v;
      }
      done += n;
    }
  }
};

} // namespace

extern "C" {

void orc_aes128_encrypt(const u8 *key, const u8 *in, u8 *out) { Aes128(key).encrypt(in, out); }

void orc_csprng_bytes(u8 *out, u64 first_byte, u64 count, u64 seed_lo, u64 seed_hi) {
  Stream(seed_lo, seed_hi).bytes(out, first_byte, count);
}

void orc_csprng_fill_u64(u64 *out, u64 first_word, u64 count, u64 seed_lo, u64 seed_hi) {
  const Stream s(seed_lo, seed_hi);
  const long chunks = (long)((count + 4095) / 4096);
#pragma omp parallel for schedule(static)
  for (long c = 0; c < chunks; ++c) {
    const u64 b = (u64)c * 4096, n = std::min<u64>(4096, count - b);
    s.words(out + b, first_word + b, n);
  }
}

// seeded BSK [n][l][k+1 rows][N] (bodies) -> standard BSK [n][l][k+1 rows][k+1 polys][N]
void orc_decompress_seeded_bsk(u64 *out, const u64 *seeded, unsigned n, unsigned k, size_t N,
                               unsigned level_count, u64 seed_lo, u64 seed_hi) {
  const Stream s(seed_lo, seed_hi);
  const size_t gs = (size_t)k + 1, mask = (size_t)k * N;
  const long rows = (long)n * level_count * (long)gs; // GGSW i, level matrix, row: generator order
#pragma omp parallel for schedule(static)
  for (long row = 0; row < rows; ++row) {
    u64 *glwe = out + (size_t)row * gs * N;
    s.words(glwe, (u64)row * mask, mask);                        // mask: k polynomials
    std::memcpy(glwe + mask, seeded + (size_t)row * N, N * sizeof(u64)); // body copied
  }
}

// seeded KSK [n_in][l][1] (bodies) -> KSK [n_in][l][n_out + 1]
void orc_decompress_seeded_ksk(u64 *out, const u64 *seeded, unsigned n_in, unsigned n_out,
                               unsigned level_count, u64 seed_lo, u64 seed_hi) {
  const Stream s(seed_lo, seed_hi);
  const long cts = (long)n_in * level_count;
#pragma omp parallel for schedule(static)
  for (long c = 0; c < cts; ++c) {
    u64 *lwe = out + (size_t)c * (n_out + 1);
    s.words(lwe, (u64)c * n_out, n_out);
    lwe[n_out] = seeded[c];
  }
}

} // extern "C"
