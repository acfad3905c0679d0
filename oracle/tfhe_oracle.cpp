// =============================================================================
// oracle/tfhe_oracle.cpp — CPU restatement of the reference's TFHE hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under concrete_b200/ may include, link or
// call this file.  It is used by tests/, __graft_entry__.smoke() and by
// bench.py's cpu_baseline / `--impl reference` legs as the checker / baseline.
//
// What it restates (file:line are under /root/reference/third_party/tfhe-rs/
// tfhe/src/core_crypto unless another root is given):
//   closest_representable      commons/math/decomposition/decomposer.rs:23-47
//   decompose (signed digits)  commons/math/decomposition/iter.rs:38-51,121-133
//   keyswitch                  algorithms/lwe_keyswitch.rs:138-228
//   modulus switch             fft_impl/common.rs:10-25
//   monomial div / mul-sub     algorithms/polynomial_algorithms.rs:500-539,610-690
//   negacyclic f64 transform   fft_impl/fft64/math/fft/mod.rs:62-73,201-308,491-561
//   from_torus                 commons/math/torus/mod.rs:72-78
//   external product           fft_impl/fft64/crypto/ggsw.rs:231-245,481-602
//   blind rotate / bootstrap   fft_impl/fft64/crypto/bootstrap.rs:250-368
//   sample extraction          algorithms/glwe_sample_extraction.rs:90-165
//   BSK -> Fourier             fft_impl/fft64/crypto/{bootstrap.rs:187-201,ggsw.rs:253-276}
//   key layouts (harness)      algorithms/lwe_keyswitch_key_generation.rs:155-180,
//                              algorithms/ggsw_encryption.rs:20-44,130-160,320-372
//
// PARITY PIN STATUS
//   * integer pieces (decomposer, keyswitch, monomial ops, sample extraction,
//     modulus switch) are pinned against the reference's own doc-test vectors
//     (tests/test_oracle_golden.py) and restated property tests.
//   * PBS ciphertext words: "parity unpinned".  The f64 FFT core of the
//     reference is concrete-fft 0.5.1 (crates.io, not vendored; plan chosen at
//     run time by timing), so the reference itself is not bit-stable there.
//     Only decryptions and a torus tolerance are pinned.
// =============================================================================
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef std::complex<double> c64;
typedef uint64_t u64;
typedef int64_t i64;

// ----------------------------------------------------------------------------
// Decomposition (generic over the unsigned scalar so the u8/u32 doc-test
// vectors of the reference can be replayed).
// ----------------------------------------------------------------------------
template <class T>
static inline T closest_representable(T input, unsigned level_count,
                                      unsigned base_log) {
  // decomposer.rs:23-47
  const unsigned bits = sizeof(T) * 8;
  const unsigned non_rep_bit_count = bits - level_count * base_log;
  const unsigned shift = non_rep_bit_count - 1;
  T res = (T)(input >> shift);
  res = (T)(res + (T)1);
  res = (T)(res & (T)(~(T)1));
  return (T)(res << shift);
}

template <class T>
static inline T decompose_one_level(unsigned base_log, T &state, T mod_b_mask) {
  // iter.rs:121-133
  T res = (T)(state & mod_b_mask);
  state = (T)(state >> base_log);
  T carry = (T)(((T)(res - (T)1) | state) & res);
  carry = (T)(carry >> (base_log - 1));
  state = (T)(state + carry);
  return (T)(res - (T)(carry << base_log));
}

// Yields level l first (least significant), level 1 last — iter.rs:38-51.
template <class T>
static inline void decompose(T input, unsigned level_count, unsigned base_log,
                             T *terms /* [level_count], index 0 = level l */) {
  const unsigned bits = sizeof(T) * 8;
  T rounded = closest_representable<T>(input, level_count, base_log);
  T state = (T)(rounded >> (bits - base_log * level_count));
  T mask = (T)(((T)1 << base_log) - (T)1);
  for (unsigned j = 0; j < level_count; ++j)
    terms[j] = decompose_one_level<T>(base_log, state, mask);
}

static inline u64 modulus_switch(u64 input, unsigned log_modulus) {
  // fft_impl/common.rs:17-25
  u64 t = input + ((u64)1 << (64 - log_modulus - 1));
  return t >> (64 - log_modulus);
}

// ----------------------------------------------------------------------------
// Polynomial helpers
// ----------------------------------------------------------------------------
template <class T>
static void monomial_div(T *out, const T *in, size_t N, size_t degree) {
  // polynomial_algorithms.rs:500-539
  size_t rem = degree % N;
  size_t cycles = degree / N;
  bool neg_first = (cycles % 2) != 0;
  for (size_t i = 0; i < N - rem; ++i)
    out[i] = neg_first ? (T)(0 - in[rem + i]) : in[rem + i];
  for (size_t i = 0; i < rem; ++i)
    out[N - rem + i] = neg_first ? in[i] : (T)(0 - in[i]);
}

template <class T>
static void monomial_mul(T *out, const T *in, size_t N, size_t degree) {
  // polynomial_algorithms.rs:560-600
  size_t rem = degree % N;
  size_t cycles = degree / N;
  bool flip = (cycles % 2) != 0;
  for (size_t i = 0; i < rem; ++i)
    out[i] = flip ? in[N - rem + i] : (T)(0 - in[N - rem + i]);
  for (size_t i = rem; i < N; ++i)
    out[i] = flip ? (T)(0 - in[i - rem]) : in[i - rem];
}

template <class T>
static void monomial_mul_and_subtract(T *out, const T *in, size_t N,
                                      size_t degree) {
  // polynomial_algorithms.rs:610-690 : out = in * X^degree - in
  size_t rem = degree % N;
  size_t cycles = degree / N;
  bool flip = (cycles % 2) != 0;
  for (size_t i = 0; i < rem; ++i) {
    T s = in[N - rem + i];
    out[i] = (T)((flip ? s : (T)(0 - s)) - in[i]);
  }
  for (size_t i = rem; i < N; ++i) {
    T s = in[i - rem];
    out[i] = (T)((flip ? (T)(0 - s) : s) - in[i]);
  }
}

// glwe_sample_extraction.rs:90-165 (native modulus branch), any nth.
template <class T>
static void sample_extract(T *lwe_out, const T *glwe, size_t k, size_t N,
                           size_t nth) {
  lwe_out[k * N] = glwe[k * N + nth];
  size_t opposite_count = N - nth - 1;
  for (size_t p = 0; p < k; ++p) {
    T *m = lwe_out + p * N;
    const T *g = glwe + p * N;
    // reverse
    for (size_t i = 0; i < N; ++i) m[i] = g[N - 1 - i];
    // negate the first opposite_count
    for (size_t i = 0; i < opposite_count; ++i) m[i] = (T)(0 - m[i]);
    // rotate left by opposite_count
    std::rotate(m, m + opposite_count, m + N);
  }
}

// ----------------------------------------------------------------------------
// Keyswitch  (lwe_keyswitch.rs:138-228; KSK layout [n_in][l][n_out+1], memory
// level index 0 == level l, lwe_keyswitch_key_generation.rs:161-172)
// ----------------------------------------------------------------------------
static void keyswitch_one(u64 *out, const u64 *in, const u64 *ksk,
                          unsigned n_in, unsigned n_out, unsigned base_log,
                          unsigned level_count) {
  const size_t out_size = (size_t)n_out + 1;
  std::fill(out, out + out_size, (u64)0);
  out[n_out] = in[n_in];
  u64 terms[64];
  for (unsigned i = 0; i < n_in; ++i) {
    decompose<u64>(in[i], level_count, base_log, terms);
    const u64 *block = ksk + (size_t)i * level_count * out_size;
    for (unsigned j = 0; j < level_count; ++j) {
      const u64 d = terms[j];
      if (d == 0) continue; // x - 0*k == x ; pure speed-up, bit-identical
      const u64 *row = block + (size_t)j * out_size;
      // slice_wrapping_sub_scalar_mul_assign (slice_algorithms.rs:363-)
      for (size_t o = 0; o < out_size; ++o) out[o] -= row[o] * d;
    }
  }
}

// ----------------------------------------------------------------------------
// Negacyclic f64 transform of size N (N/2 complex points).
// fold z_j = p_j + i p_{j+N/2}; twist by exp(i*pi*j/N)  (Twisties::new,
// fft/mod.rs:62-73); N/2-point complex DFT; "unordered" output.  Our plan:
// radix-2 DIF natural->bit-reversed forward, DIT bit-reversed->natural inverse.
// ----------------------------------------------------------------------------
struct NegFft {
  size_t N, h; // h = N/2
  unsigned logh;
  std::vector<double> tw_re, tw_im;  // twisties, size h
  std::vector<c64> w;                // w[k] = exp(-2*pi*i*k/h), k < h/2

  explicit NegFft(size_t n) : N(n), h(n / 2) {
    logh = 0;
    while (((size_t)1 << logh) < h) ++logh;
    tw_re.resize(h);
    tw_im.resize(h);
    const long double unit = 3.14159265358979323846264338327950288L / (long double)N;
    for (size_t i = 0; i < h; ++i) {
      tw_re[i] = (double)cosl(unit * (long double)i);
      tw_im[i] = (double)sinl(unit * (long double)i);
    }
    w.resize(std::max<size_t>(h / 2, 1));
    for (size_t k = 0; k < h / 2; ++k) {
      long double a = -2.0L * 3.14159265358979323846264338327950288L *
                      (long double)k / (long double)h;
      w[k] = c64((double)cosl(a), (double)sinl(a));
    }
  }

  // natural -> bit-reversed, forward sign
  void fwd(c64 *x) const {
    for (size_t len = h; len >= 2; len >>= 1) {
      size_t half = len >> 1, step = h / len;
      for (size_t base = 0; base < h; base += len) {
        c64 *a = x + base, *b = x + base + half;
        for (size_t j = 0; j < half; ++j) {
          c64 u = a[j], v = b[j];
          a[j] = u + v;
          c64 d = u - v;
          const c64 ww = w[j * step];
          b[j] = c64(d.real() * ww.real() - d.imag() * ww.imag(),
                     d.real() * ww.imag() + d.imag() * ww.real());
        }
      }
    }
  }
  // bit-reversed -> natural, inverse sign, unnormalised
  void inv(c64 *x) const {
    for (size_t len = 2; len <= h; len <<= 1) {
      size_t half = len >> 1, step = h / len;
      for (size_t base = 0; base < h; base += len) {
        c64 *a = x + base, *b = x + base + half;
        for (size_t j = 0; j < half; ++j) {
          const c64 ww = w[j * step]; // conj used below
          c64 v = b[j];
          c64 t(v.real() * ww.real() + v.imag() * ww.imag(),
                v.imag() * ww.real() - v.real() * ww.imag());
          c64 u = a[j];
          a[j] = u + t;
          b[j] = u - t;
        }
      }
    }
  }

  // convert_forward_integer (fft/mod.rs:224-243) + plan.fwd
  void forward_as_integer(c64 *out, const u64 *poly) const {
    for (size_t j = 0; j < h; ++j) {
      double re = (double)(i64)poly[j], im = (double)(i64)poly[j + h];
      out[j] = c64(re * tw_re[j] - im * tw_im[j], re * tw_im[j] + im * tw_re[j]);
    }
    fwd(out);
  }
  // convert_forward_torus (fft/mod.rs:201-222) + plan.fwd
  void forward_as_torus(c64 *out, const u64 *poly) const {
    const double norm = std::ldexp(1.0, -64);
    for (size_t j = 0; j < h; ++j) {
      double re = (double)(i64)poly[j] * norm, im = (double)(i64)poly[j + h] * norm;
      out[j] = c64(re * tw_re[j] - im * tw_im[j], re * tw_im[j] + im * tw_re[j]);
    }
    fwd(out);
  }
  static inline u64 from_torus(double x) {
    // torus/mod.rs:72-78
    double fract = x - std::round(x);
    fract *= 18446744073709551616.0; // 2^64
    fract = std::round(fract);
    if (fract >= 9223372036854775808.0) return (u64)INT64_MAX; // Rust `as` saturates
    if (fract < -9223372036854775808.0) return (u64)INT64_MIN;
    return (u64)(i64)fract;
  }
  // plan.inv + convert_add_backward_torus_scalar (fft/mod.rs:289-308); `f` is clobbered
  void add_backward_as_torus(u64 *poly, c64 *f) const {
    inv(f);
    const double norm = 1.0 / (double)h;
    for (size_t j = 0; j < h; ++j) {
      // tmp = inp * (conj(w) * normalization)
      double wr = tw_re[j] * norm, wi = -tw_im[j] * norm;
      double re = f[j].real() * wr - f[j].imag() * wi;
      double im = f[j].real() * wi + f[j].imag() * wr;
      poly[j] += from_torus(re);
      poly[j + h] += from_torus(im);
    }
  }
  void backward_as_torus(u64 *poly, c64 *f) const {
    std::fill(poly, poly + N, (u64)0);
    add_backward_as_torus(poly, f);
  }
};

static std::mutex g_plan_mutex;
static std::map<size_t, std::shared_ptr<NegFft>> g_plans;
static const NegFft &get_plan(size_t N) {
  std::lock_guard<std::mutex> g(g_plan_mutex);
  auto it = g_plans.find(N);
  if (it == g_plans.end())
    it = g_plans.emplace(N, std::make_shared<NegFft>(N)).first;
  return *it->second;
}

// ----------------------------------------------------------------------------
// External product / blind rotation / bootstrap
// Fourier BSK layout (== FourierLweBootstrapKey, bootstrap.rs:30-69):
//   [n][l (memory index m <-> level l-m)][k+1 rows][k+1 cols][N/2] c64
// ----------------------------------------------------------------------------
struct PbsScratch {
  std::vector<u64> ct1, states, tmp;
  std::vector<c64> fourier, out_fft;
};

static void add_external_product_assign(u64 *out, const c64 *ggsw, const u64 *glwe,
                                        unsigned k, size_t N, unsigned base_log,
                                        unsigned level_count, const NegFft &fft,
                                        PbsScratch &s) {
  // ggsw.rs:481-602
  const size_t h = N / 2, gs = k + 1, total = gs * N;
  s.states.resize(total);
  s.tmp.resize(total);
  s.fourier.resize(h);
  s.out_fft.assign(gs * h, c64(0, 0));
  const unsigned shift = 64 - base_log * level_count;
  for (size_t i = 0; i < total; ++i)
    s.states[i] = closest_representable<u64>(glwe[i], level_count, base_log) >> shift;
  const u64 mask = ((u64)1 << base_log) - 1;
  bool uninit = true;
  for (unsigned m = 0; m < level_count; ++m) { // memory index m <-> level l-m
    for (size_t i = 0; i < total; ++i)
      s.tmp[i] = decompose_one_level<u64>(base_log, s.states[i], mask);
    const c64 *level_mat = ggsw + (size_t)m * gs * gs * h;
    for (size_t r = 0; r < gs; ++r) {
      fft.forward_as_integer(s.fourier.data(), s.tmp.data() + r * N);
      const c64 *row = level_mat + r * gs * h;
      for (size_t c = 0; c < gs; ++c) {
        c64 *o = s.out_fft.data() + c * h;
        const c64 *g = row + c * h;
        const c64 *f = s.fourier.data();
        if (uninit) {
          for (size_t j = 0; j < h; ++j)
            o[j] = c64(g[j].real() * f[j].real() - g[j].imag() * f[j].imag(),
                       g[j].real() * f[j].imag() + g[j].imag() * f[j].real());
        } else {
          for (size_t j = 0; j < h; ++j)
            o[j] += c64(g[j].real() * f[j].real() - g[j].imag() * f[j].imag(),
                        g[j].real() * f[j].imag() + g[j].imag() * f[j].real());
        }
      }
      uninit = false;
    }
  }
  for (size_t c = 0; c < gs; ++c)
    fft.add_backward_as_torus(out + c * N, s.out_fft.data() + c * h);
}

static void blind_rotate_assign(u64 *acc /*(k+1)N, in: LUT*/, const u64 *lwe,
                                const c64 *fbsk, unsigned n, unsigned k, size_t N,
                                unsigned base_log, unsigned level_count,
                                const NegFft &fft, PbsScratch &s) {
  // bootstrap.rs:250-333
  const size_t gs = k + 1, total = gs * N, h = N / 2;
  unsigned log2N = 0;
  while (((size_t)1 << log2N) < N) ++log2N;
  const unsigned log_mod = log2N + 1;
  s.ct1.resize(total);
  {
    size_t deg = (size_t)modulus_switch(lwe[n], log_mod);
    for (size_t p = 0; p < gs; ++p) {
      std::copy(acc + p * N, acc + (p + 1) * N, s.ct1.begin() + p * N);
      monomial_div<u64>(acc + p * N, s.ct1.data() + p * N, N, deg);
    }
  }
  const size_t ggsw_stride = (size_t)level_count * gs * gs * h;
  for (unsigned i = 0; i < n; ++i) {
    if (lwe[i] == 0) continue; // bootstrap.rs:287
    size_t deg = (size_t)modulus_switch(lwe[i], log_mod);
    for (size_t p = 0; p < gs; ++p)
      monomial_mul_and_subtract<u64>(s.ct1.data() + p * N, acc + p * N, N, deg);
    add_external_product_assign(acc, fbsk + (size_t)i * ggsw_stride, s.ct1.data(), k,
                                N, base_log, level_count, fft, s);
  }
}

// ----------------------------------------------------------------------------
// Harness RNG (free design; conventions are the reference's)
// ----------------------------------------------------------------------------
struct Rng {
  u64 s;
  explicit Rng(u64 seed, u64 stream = 0) {
    s = seed * 0x9E3779B97F4A7C15ULL + stream * 0xD1342543DE82EF95ULL + 0x2545F4914F6CDD1DULL;
    next(); next();
  }
  inline u64 next() { // splitmix64
    u64 z = (s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
  }
  inline double uniform01() { return ((double)(next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
  inline double gauss() {
    double u1 = uniform01(), u2 = uniform01();
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586476925 * u2);
  }
  // torus gaussian noise with std `sigma` (fraction of the torus), as u64
  inline u64 noise(double sigma) {
    if (sigma <= 0.0) return 0;
    double v = gauss() * sigma; // |v| << 0.5
    return (u64)(i64)std::llround(v * 18446744073709551616.0);
  }
};

// Exact negacyclic product of a u64 polynomial by a binary polynomial mod 2^64
// (harness keygen only).  16-bit limbs of `a` times bits of `s`: every exact
// coefficient is < 2^16 * N <= 2^30, far inside f64-FFT exactness.
static void negacyclic_mul_binary(u64 *out, const u64 *a, const u64 *sbits, size_t N,
                                  const NegFft &fft) {
  const size_t h = N / 2;
  std::vector<c64> fs(h), fa(h);
  std::vector<u64> limb(N);
  fft.forward_as_integer(fs.data(), sbits);
  std::fill(out, out + N, (u64)0);
  for (unsigned L = 0; L < 4; ++L) {
    for (size_t i = 0; i < N; ++i) limb[i] = (a[i] >> (16 * L)) & 0xFFFF;
    fft.forward_as_integer(fa.data(), limb.data());
    for (size_t j = 0; j < h; ++j) fa[j] *= fs[j];
    fft.inv(fa.data());
    const double norm = 1.0 / (double)h;
    for (size_t j = 0; j < h; ++j) {
      double wr = fft.tw_re[j] * norm, wi = -fft.tw_im[j] * norm;
      double re = fa[j].real() * wr - fa[j].imag() * wi;
      double im = fa[j].real() * wi + fa[j].imag() * wr;
      out[j] += (u64)(i64)std::llround(re) << (16 * L);
      out[j + h] += (u64)(i64)std::llround(im) << (16 * L);
    }
  }
}

static void negacyclic_mul_schoolbook(u64 *out, const u64 *a, const u64 *b, size_t N) {
  std::fill(out, out + N, (u64)0);
  for (size_t i = 0; i < N; ++i) {
    if (b[i] == 0) continue;
    for (size_t j = 0; j < N; ++j) {
      size_t d = i + j;
      u64 v = a[j] * b[i];
      if (d < N) out[d] += v; else out[d - N] -= v;
    }
  }
}

// =============================================================================
// C API
// =============================================================================
extern "C" {

int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_num_threads(int t) {
#ifdef _OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}

// ---- scalar pieces exposed for golden-vector tests --------------------------
uint8_t orc_closest_representable_u8(uint8_t x, unsigned l, unsigned b) { return closest_representable<uint8_t>(x, l, b); }
uint32_t orc_closest_representable_u32(uint32_t x, unsigned l, unsigned b) { return closest_representable<uint32_t>(x, l, b); }
u64 orc_closest_representable_u64(u64 x, unsigned l, unsigned b) { return closest_representable<u64>(x, l, b); }
void orc_decompose_u32(uint32_t x, unsigned l, unsigned b, uint32_t *terms) { decompose<uint32_t>(x, l, b, terms); }
void orc_decompose_u64(u64 x, unsigned l, unsigned b, u64 *terms) { decompose<u64>(x, l, b, terms); }
u64 orc_modulus_switch(u64 x, unsigned log_modulus) { return modulus_switch(x, log_modulus); }
u64 orc_from_torus(double x) { return NegFft::from_torus(x); }
void orc_monomial_div_u8(uint8_t *o, const uint8_t *i, size_t N, size_t d) { monomial_div<uint8_t>(o, i, N, d); }
void orc_monomial_mul_u8(uint8_t *o, const uint8_t *i, size_t N, size_t d) { monomial_mul<uint8_t>(o, i, N, d); }
void orc_monomial_div_u64(u64 *o, const u64 *i, size_t N, size_t d) { monomial_div<u64>(o, i, N, d); }
void orc_monomial_mul_u64(u64 *o, const u64 *i, size_t N, size_t d) { monomial_mul<u64>(o, i, N, d); }
void orc_monomial_mul_and_subtract_u64(u64 *o, const u64 *i, size_t N, size_t d) { monomial_mul_and_subtract<u64>(o, i, N, d); }
void orc_sample_extract_u64(u64 *lwe, const u64 *glwe, size_t k, size_t N, size_t nth) { sample_extract<u64>(lwe, glwe, k, N, nth); }
void orc_negacyclic_mul_schoolbook(u64 *o, const u64 *a, const u64 *b, size_t N) { negacyclic_mul_schoolbook(o, a, b, N); }
void orc_negacyclic_mul_binary(u64 *o, const u64 *a, const u64 *s, size_t N) { negacyclic_mul_binary(o, a, s, N, get_plan(N)); }

// ---- FFT pieces (fft/tests.rs restatements) ---------------------------------
void orc_fft_forward_as_torus(double *out /*N doubles: N/2 c64*/, const u64 *poly, size_t N) {
  get_plan(N).forward_as_torus(reinterpret_cast<c64 *>(out), poly);
}
void orc_fft_forward_as_integer(double *out, const u64 *poly, size_t N) {
  get_plan(N).forward_as_integer(reinterpret_cast<c64 *>(out), poly);
}
void orc_fft_add_backward_as_torus(u64 *poly, const double *f, size_t N) {
  std::vector<c64> tmp(N / 2);
  std::memcpy(tmp.data(), f, N / 2 * sizeof(c64));
  get_plan(N).add_backward_as_torus(poly, tmp.data());
}

// ---- keyswitch (batched; concrete_cpu_keyswitch_lwe_ciphertext_u64,
//      /root/reference/backends/concrete-cpu/implementation/src/c_api/keyswitch.rs:185-223,
//      looped like memref_batched_keyswitch_lwe_u64, wrappers.cpp:745-760) ------
void orc_keyswitch_batch(u64 *out, const u64 *in, const u64 *ksk, unsigned n_in,
                         unsigned n_out, unsigned base_log, unsigned level_count,
                         size_t num) {
#pragma omp parallel for schedule(dynamic, 1)
  for (long s = 0; s < (long)num; ++s)
    keyswitch_one(out + (size_t)s * (n_out + 1), in + (size_t)s * (n_in + 1), ksk, n_in,
                  n_out, base_log, level_count);
}

// ---- BSK standard -> Fourier (concrete_cpu_bootstrap_key_convert_u64_to_fourier,
//      c_api/bootstrap.rs:261-318) ----------------------------------------------
void orc_bsk_to_fourier(double *fourier, const u64 *bsk, unsigned n, unsigned k,
                        unsigned level_count, size_t N) {
  const NegFft &fft = get_plan(N);
  const size_t polys = (size_t)n * level_count * (k + 1) * (k + 1);
  c64 *f = reinterpret_cast<c64 *>(fourier);
#pragma omp parallel for schedule(static)
  for (long p = 0; p < (long)polys; ++p)
    fft.forward_as_torus(f + (size_t)p * (N / 2), bsk + (size_t)p * N);
}

// ---- bootstrap (batched; concrete_cpu_bootstrap_lwe_ciphertext_u64,
//      c_api/bootstrap.rs:347-414, looped like memref_batched_bootstrap_lwe_u64,
//      wrappers.cpp:807-826).  luts: [num_luts][(k+1)N] trivial GLWEs; lut_idx
//      per sample (NULL => 0).  nth = sample extraction index (0 in Concrete).
void orc_bootstrap_batch(u64 *out, const u64 *in, const u64 *luts, const u64 *lut_idx,
                         const double *fourier_bsk, unsigned n, unsigned k, size_t N,
                         unsigned base_log, unsigned level_count, size_t num,
                         size_t nth) {
  const NegFft &fft = get_plan(N);
  const c64 *fbsk = reinterpret_cast<const c64 *>(fourier_bsk);
  const size_t gs = k + 1;
#pragma omp parallel
  {
    PbsScratch s;
    std::vector<u64> acc(gs * N);
#pragma omp for schedule(dynamic, 1)
    for (long i = 0; i < (long)num; ++i) {
      const u64 *lut = luts + (lut_idx ? lut_idx[i] : 0) * gs * N;
      std::copy(lut, lut + gs * N, acc.begin());
      blind_rotate_assign(acc.data(), in + (size_t)i * (n + 1), fbsk, n, k, N, base_log,
                          level_count, fft, s);
      sample_extract<u64>(out + (size_t)i * (k * N + 1), acc.data(), k, N, nth);
    }
  }
}

// blind rotation only (returns the accumulator) — used to test the GPU kernel's
// intermediate state.
void orc_blind_rotate_batch(u64 *acc_out, const u64 *in, const u64 *luts,
                            const u64 *lut_idx, const double *fourier_bsk, unsigned n,
                            unsigned k, size_t N, unsigned base_log,
                            unsigned level_count, size_t num) {
  const NegFft &fft = get_plan(N);
  const c64 *fbsk = reinterpret_cast<const c64 *>(fourier_bsk);
  const size_t gs = k + 1;
#pragma omp parallel
  {
    PbsScratch s;
#pragma omp for schedule(dynamic, 1)
    for (long i = 0; i < (long)num; ++i) {
      u64 *acc = acc_out + (size_t)i * gs * N;
      const u64 *lut = luts + (lut_idx ? lut_idx[i] : 0) * gs * N;
      std::copy(lut, lut + gs * N, acc);
      blind_rotate_assign(acc, in + (size_t)i * (n + 1), fbsk, n, k, N, base_log,
                          level_count, fft, s);
    }
  }
}

// ---- harness: keys, encryption, decryption -----------------------------------
void orc_gen_binary_key(u64 *key, size_t len, u64 seed) {
  Rng r(seed, 0x5EC2E7);
  for (size_t i = 0; i < len; ++i) key[i] = r.next() >> 63;
}

// body = <mask, s> + e + plaintext
void orc_lwe_encrypt_batch(u64 *ct, const u64 *key, const u64 *plaintexts, unsigned n,
                           size_t num, double sigma, u64 seed) {
#pragma omp parallel for schedule(static)
  for (long s = 0; s < (long)num; ++s) {
    Rng r(seed, 0x1000000ULL + (u64)s);
    u64 *c = ct + (size_t)s * (n + 1);
    u64 body = plaintexts[s] + r.noise(sigma);
    for (unsigned i = 0; i < n; ++i) {
      c[i] = r.next();
      body += c[i] * key[i];
    }
    c[n] = body;
  }
}

void orc_lwe_decrypt_batch(u64 *plaintexts, const u64 *ct, const u64 *key, unsigned n,
                           size_t num) {
#pragma omp parallel for schedule(static)
  for (long s = 0; s < (long)num; ++s) {
    const u64 *c = ct + (size_t)s * (n + 1);
    u64 v = c[n];
    for (unsigned i = 0; i < n; ++i) v -= c[i] * key[i];
    plaintexts[s] = v;
  }
}

// KSK: block i, memory level index j <-> level l-j, plaintext
// s_in[i] << (64 - b*level)   (lwe_keyswitch_key_generation.rs:161-172)
// `masks` (optional, [n_in][l][n_out]): mask elements to use instead of the harness RNG's — the
// seeded-key tests pass the compression generator's stream (oracle/csprng_oracle.cpp) here, as
// generate_seeded_lwe_keyswitch_key does (lwe_keyswitch_key_generation.rs, seeded variant).
void orc_gen_ksk_masks(u64 *ksk, const u64 *key_in, const u64 *key_out, unsigned n_in,
                       unsigned n_out, unsigned base_log, unsigned level_count, double sigma,
                       u64 seed, const u64 *masks);
void orc_gen_ksk(u64 *ksk, const u64 *key_in, const u64 *key_out, unsigned n_in,
                 unsigned n_out, unsigned base_log, unsigned level_count, double sigma,
                 u64 seed) {
  orc_gen_ksk_masks(ksk, key_in, key_out, n_in, n_out, base_log, level_count, sigma, seed, nullptr);
}
void orc_gen_ksk_masks(u64 *ksk, const u64 *key_in, const u64 *key_out, unsigned n_in,
                       unsigned n_out, unsigned base_log, unsigned level_count, double sigma,
                       u64 seed, const u64 *masks) {
  const size_t out_size = (size_t)n_out + 1;
#pragma omp parallel for schedule(static)
  for (long i = 0; i < (long)n_in; ++i) {
    Rng r(seed, 0x2000000ULL + (u64)i);
    for (unsigned j = 0; j < level_count; ++j) {
      unsigned level = level_count - j;
      u64 pt = key_in[i] << (64 - base_log * level);
      u64 *c = ksk + ((size_t)i * level_count + j) * out_size;
      u64 body = pt + r.noise(sigma);
      const u64 *given = masks ? masks + ((size_t)i * level_count + j) * n_out : nullptr;
      for (unsigned o = 0; o < n_out; ++o) {
        const u64 drawn = r.next(); // drawn either way: the noise stream does not depend on `masks`
        c[o] = given ? given[o] : drawn;
        body += c[o] * key_out[o];
      }
      c[n_out] = body;
    }
  }
}

// BSK: GGSW(s_i) for every bit of the small key, standard domain.
// layout [n][l (memory m <-> level l-m)][k+1 rows][k+1 polys][N]
// factor = (-s_i) << (64 - b*level); row r<k encrypts S_r*factor, last row encrypts
// -factor in coefficient 0  (ggsw_encryption.rs:20-44,320-372).
// `masks` (optional, [n][l][k+1 rows][k][N]): as for orc_gen_ksk_masks
// (generate_seeded_lwe_bootstrap_key, c_api/bootstrap.rs:120-165).
void orc_gen_bsk_masks(u64 *bsk, const u64 *lwe_key, const u64 *glwe_key, unsigned n, unsigned k,
                       size_t N, unsigned base_log, unsigned level_count, double sigma, u64 seed,
                       const u64 *masks);
void orc_gen_bsk(u64 *bsk, const u64 *lwe_key /*n*/, const u64 *glwe_key /*k*N*/,
                 unsigned n, unsigned k, size_t N, unsigned base_log,
                 unsigned level_count, double sigma, u64 seed) {
  orc_gen_bsk_masks(bsk, lwe_key, glwe_key, n, k, N, base_log, level_count, sigma, seed, nullptr);
}
void orc_gen_bsk_masks(u64 *bsk, const u64 *lwe_key /*n*/, const u64 *glwe_key /*k*N*/,
                       unsigned n, unsigned k, size_t N, unsigned base_log,
                       unsigned level_count, double sigma, u64 seed, const u64 *masks) {
  const NegFft &fft = get_plan(N);
  const size_t gs = k + 1;
  const long rows = (long)n * level_count * gs;
#pragma omp parallel
  {
    std::vector<u64> prod(N);
#pragma omp for schedule(dynamic, 4)
    for (long idx = 0; idx < rows; ++idx) {
      size_t i = idx / (level_count * gs);
      unsigned m = (idx / gs) % level_count;
      size_t r = idx % gs;
      unsigned level = level_count - m;
      Rng rng(seed, 0x3000000ULL + (u64)idx);
      u64 factor = (u64)(0 - lwe_key[i]) << (64 - base_log * level);
      u64 *row = bsk + (size_t)idx * gs * N;
      u64 *body = row + k * N;
      // plaintext + noise
      for (size_t c = 0; c < N; ++c) body[c] = rng.noise(sigma);
      if (r < k) {
        for (size_t c = 0; c < N; ++c) body[c] += glwe_key[r * N + c] * factor;
      } else {
        body[0] += (u64)(0 - factor);
      }
      // mask + <mask, key>
      for (size_t p = 0; p < k; ++p) {
        u64 *a = row + p * N;
        const u64 *given = masks ? masks + ((size_t)idx * k + p) * N : nullptr;
        for (size_t c = 0; c < N; ++c) {
          const u64 drawn = rng.next();
          a[c] = given ? given[c] : drawn;
        }
        negacyclic_mul_binary(prod.data(), a, glwe_key + p * N, N, fft);
        for (size_t c = 0; c < N; ++c) body[c] += prod[c];
      }
    }
  }
}

} // extern "C"

#include "wop_oracle.inl"
