// oracle/tfhe_fast.cpp — the TIMED CPU baseline of bench.py: keyswitch + programmable bootstrap,
// vectorised.  TEST/BENCH INFRASTRUCTURE, not product code and not the checker: the scalar
// restatement in tfhe_oracle.cpp stays the oracle; this file exists so that the CPU number next
// to the GPU number is not handicapped by a scalar radix-2 FFT ("696x overstates the
// distance to the real reference CPU path by ~10x").  concrete-cpu itself (Rust, concrete-fft
// 0.5.1 with AVX-512 kernels) cannot be built on this box.
//
// Same algorithm as the oracle, i.e. as the reference:
//   keyswitch      tfhe/src/core_crypto/algorithms/lwe_keyswitch.rs:138-228
//   blind rotate   tfhe/src/core_crypto/fft_impl/fft64/crypto/bootstrap.rs:250-333
//   ext. product   tfhe/src/core_crypto/fft_impl/fft64/crypto/ggsw.rs:481-602
//   transforms     tfhe/src/core_crypto/fft_impl/fft64/math/fft/mod.rs:201-308
// What is different is the schedule: W = 8 ciphertexts travel together through every transform,
// one per SIMD lane (GCC vector extensions: AVX-512 where the host has it, 2 x AVX2 otherwise), so
// every butterfly is a full-width vector operation with scalar (broadcast) twiddles and no
// shuffles; two radix-2 stages are fused per pass over the data to halve the cache traffic.  The
// Fourier key is the oracle's (same fold, twist, radix-2 DIF ordering), so both consume the same
// key bytes.  Results are validated against the scalar oracle by tests/test_oracle_golden.py
// (decryptions equal, phases within the PBS tolerance).
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#if defined(__AVX512F__)
#include <immintrin.h>
#endif

typedef uint64_t u64;
typedef int64_t i64;
constexpr int W = 8;
typedef double vd __attribute__((vector_size(W * 8)));
typedef u64 vu __attribute__((vector_size(W * 8)));
typedef i64 vi __attribute__((vector_size(W * 8)));

namespace {

struct Plan {
  size_t N, h;
  std::vector<double> tw_re, tw_im; // twisties exp(i*pi*j/N), j < h
  std::vector<double> w_re, w_im;   // exp(-2*pi*i*k/h), k < h/2
  explicit Plan(size_t n) : N(n), h(n / 2) {
    const long double pi = 3.14159265358979323846264338327950288L;
    tw_re.resize(h), tw_im.resize(h);
    for (size_t j = 0; j < h; ++j) {
      tw_re[j] = (double)cosl(pi * (long double)j / (long double)N);
      tw_im[j] = (double)sinl(pi * (long double)j / (long double)N);
    }
    w_re.resize(std::max<size_t>(h / 2, 1)), w_im.resize(std::max<size_t>(h / 2, 1));
    for (size_t k = 0; k < h / 2; ++k) {
      w_re[k] = (double)cosl(-2.0L * pi * (long double)k / (long double)h);
      w_im[k] = (double)sinl(-2.0L * pi * (long double)k / (long double)h);
    }
  }
  // radix-2 DIF stage of block length `len` on (a, b) = (x[j], x[j + len/2]), twiddle index j*step
  inline void dif(vd &ar, vd &ai, vd &br, vd &bi, size_t widx) const {
    const double wr = w_re[widx], wi = w_im[widx];
    const vd dr = ar - br, di = ai - bi;
    ar += br;
    ai += bi;
    br = dr * wr - di * wi;
    bi = dr * wi + di * wr;
  }
  inline void dit(vd &ar, vd &ai, vd &br, vd &bi, size_t widx) const {
    const double wr = w_re[widx], wi = w_im[widx]; // conj(w) applied to b
    const vd tr = br * wr + bi * wi, ti = bi * wr - br * wi;
    br = ar - tr;
    bi = ai - ti;
    ar += tr;
    ai += ti;
  }
  // natural -> bit-reversed (forward sign), same ordering as NegFft::fwd of the oracle
  void fwd(vd *re, vd *im) const {
    size_t len = h;
    for (; len >= 4; len >>= 2) { // stages `len` and `len/2` in one pass
      const size_t q = len >> 2, s1 = h / len, s2 = 2 * s1;
      for (size_t base = 0; base < h; base += len)
        for (size_t j = 0; j < q; ++j) {
          vd ar = re[base + j], ai = im[base + j], br = re[base + j + q], bi = im[base + j + q];
          vd cr = re[base + j + 2 * q], ci = im[base + j + 2 * q];
          vd dr = re[base + j + 3 * q], di = im[base + j + 3 * q];
          dif(ar, ai, cr, ci, j * s1);
          dif(br, bi, dr, di, (j + q) * s1);
          dif(ar, ai, br, bi, j * s2);
          dif(cr, ci, dr, di, j * s2);
          re[base + j] = ar, im[base + j] = ai, re[base + j + q] = br, im[base + j + q] = bi;
          re[base + j + 2 * q] = cr, im[base + j + 2 * q] = ci;
          re[base + j + 3 * q] = dr, im[base + j + 3 * q] = di;
        }
    }
    if (len == 2)
      for (size_t base = 0; base < h; base += 2) dif(re[base], im[base], re[base + 1], im[base + 1], 0);
  }
  // bit-reversed -> natural (inverse sign), unnormalised: mirror of fwd
  void inv(vd *re, vd *im) const {
    size_t len = 2;
    unsigned logh = 0;
    while (((size_t)1 << logh) < h) ++logh;
    if (logh & 1) {
      for (size_t base = 0; base < h; base += 2) dit(re[base], im[base], re[base + 1], im[base + 1], 0);
      len = 4;
    }
    for (; len * 2 <= h; len <<= 2) { // stages `len` and `2*len` in one pass
      const size_t L2 = len * 2, q = len >> 1, s1 = h / len, s2 = h / L2;
      for (size_t base = 0; base < h; base += L2)
        for (size_t j = 0; j < q; ++j) {
          vd ar = re[base + j], ai = im[base + j], br = re[base + j + q], bi = im[base + j + q];
          vd cr = re[base + j + 2 * q], ci = im[base + j + 2 * q];
          vd dr = re[base + j + 3 * q], di = im[base + j + 3 * q];
          dit(ar, ai, br, bi, j * s1);
          dit(cr, ci, dr, di, j * s1);
          dit(ar, ai, cr, ci, j * s2);
          dit(br, bi, dr, di, (j + q) * s2);
          re[base + j] = ar, im[base + j] = ai, re[base + j + q] = br, im[base + j + q] = bi;
          re[base + j + 2 * q] = cr, im[base + j + 2 * q] = ci;
          re[base + j + 3 * q] = dr, im[base + j + 3 * q] = di;
        }
    }
  }
};

inline u64 closest_representable(u64 x, unsigned level, unsigned base_log) {
  const unsigned shift = 64 - level * base_log - 1;
  u64 r = x >> shift;
  r += 1;
  r &= ~(u64)1;
  return r << shift;
}
inline u64 modulus_switch(u64 x, unsigned log_mod) {
  return (x + ((u64)1 << (63 - log_mod))) >> (64 - log_mod);
}

struct Scratch {
  std::vector<vu> acc, state;        // [(k+1)*N], lanes = ciphertexts
  std::vector<vd> f_re, f_im;        // [h]
  std::vector<vd> o_re, o_im;        // [(k+1)*h]
};

// lane w of the result = src[idx[w]] lane w (per-ciphertext rotation amounts)
inline vu gather_lanes(const vu *src, const vu idx) {
#if defined(__AVX512F__)
  const __m512i off = _mm512_add_epi64(_mm512_slli_epi64((__m512i)idx, 3),
                                       _mm512_set_epi64(7, 6, 5, 4, 3, 2, 1, 0));
  return (vu)_mm512_i64gather_epi64(off, (const void *)src, 8);
#else
  vu r;
  for (int w = 0; w < W; ++w) r[w] = src[idx[w]][w];
  return r;
#endif
}

// round(x) for |x| < 2^51 without leaving the vector unit
inline vd vround(vd x) {
  const double C = 6755399441055744.0; // 1.5 * 2^52
  return (x + C) - C;
}
// (x - round(x)) * 2^64 as wrapping u64 (torus/mod.rs:72-78), lanes in parallel
inline vu from_torus(vd x) {
  vd v = (x - vround(x)) * 18446744073709551616.0; // |v| <= 2^63
  // values of magnitude >= 2^52 are integers already; below, round through the magic constant
  const vd av = v < 0.0 ? -v : v;
  const vd rv = av < 4503599627370496.0 ? vround(v) : v;
  return (vu) __builtin_convertvector(rv, vi);
}

// One batch of up to W ciphertexts through blind rotation + sample extraction.
void bootstrap_batch(u64 *out, const u64 *cts, const u64 *luts, const u64 *lut_idx,
                     const std::complex<double> *fbsk, unsigned n, unsigned k, size_t N,
                     unsigned base_log, unsigned level, size_t first, size_t count,
                     const Plan &pl, Scratch &s) {
  const size_t gs = k + 1, total = gs * N, h = N / 2;
  unsigned log2N = 0;
  while (((size_t)1 << log2N) < N) ++log2N;
  const unsigned log_mod = log2N + 1;
  s.acc.resize(total);
  s.state.resize(total);
  s.f_re.resize(h), s.f_im.resize(h), s.o_re.resize(gs * h), s.o_im.resize(gs * h);
  const u64 *lwe[W];
  for (int w = 0; w < W; ++w) {
    const size_t smp = first + std::min<size_t>(w, count - 1); // idle lanes repeat the last one
    lwe[w] = cts + smp * (n + 1);
    const u64 *lut = luts + (lut_idx ? lut_idx[smp] : 0) * total;
    const size_t deg = (size_t)modulus_switch(lwe[w][n], log_mod);
    for (size_t p = 0; p < gs; ++p) // ACC = LUT * X^(-deg)
      for (size_t j = 0; j < N; ++j) {
        const size_t u = (j + deg) & (2 * N - 1);
        const u64 v = lut[p * N + (u & (N - 1))];
        s.acc[p * N + j][w] = (u < N) ? v : (u64)0 - v;
      }
  }
  const unsigned shift = 64 - base_log * level;
  const unsigned rshift = shift - 1;
  const u64 mask = ((u64)1 << base_log) - 1;
  const size_t ggsw_stride = (size_t)level * gs * gs * h;
  const double inv_h = 1.0 / (double)h;
  for (unsigned i = 0; i < n; ++i) {
    // ct1 = ACC * X^a - ACC, rounded, as decomposition states (lane w = ciphertext w); a lane whose
    // mask element is 0 gets states 0 and therefore adds exactly 0 (bootstrap.rs:287)
    vu deg;
    for (int w = 0; w < W; ++w)
      deg[w] = lwe[w][i] == 0 ? 0 : modulus_switch(lwe[w][i], log_mod);
    for (size_t p = 0; p < gs; ++p) {
      const vu *ap = s.acc.data() + p * N;
      vu *sp = s.state.data() + p * N;
      for (size_t j = 0; j < N; ++j) {
        const vu u = ((u64)(j + 2 * N) - deg) & (u64)(2 * N - 1);
        const vu v = gather_lanes(ap, u & (u64)(N - 1));
        const vu rot = u < (u64)N ? v : (u64)0 - v;
        vu r = (rot - ap[j]) >> rshift; // closest_representable, then the state
        r += 1;
        sp[j] = (r >> 1) & (~(u64)0 >> shift);
      }
    }
    std::fill(s.o_re.begin(), s.o_re.end(), vd{});
    std::fill(s.o_im.begin(), s.o_im.end(), vd{});
    const std::complex<double> *ggsw = fbsk + (size_t)i * ggsw_stride;
    for (unsigned m = 0; m < level; ++m) { // memory level m <-> level l-m
      for (size_t r = 0; r < gs; ++r) {
        vu *sp = s.state.data() + r * N;
        for (size_t j = 0; j < h; ++j) { // digits, fold, twist
          vd d[2];
          for (int half = 0; half < 2; ++half) {
            vu &state = sp[j + half * h];
            vu res = state & mask;
            state >>= base_log;
            vu carry = ((res - 1) | state) & res;
            carry >>= (base_log - 1);
            state += carry;
            res -= carry << base_log;
            d[half] = __builtin_convertvector((vi)res, vd);
          }
          const double tr = pl.tw_re[j], ti = pl.tw_im[j];
          s.f_re[j] = d[0] * tr - d[1] * ti;
          s.f_im[j] = d[0] * ti + d[1] * tr;
        }
        pl.fwd(s.f_re.data(), s.f_im.data());
        const std::complex<double> *row = ggsw + ((size_t)m * gs + r) * gs * h;
        for (size_t c = 0; c < gs; ++c) {
          const std::complex<double> *g = row + c * h;
          vd *ore = s.o_re.data() + c * h, *oim = s.o_im.data() + c * h;
          for (size_t j = 0; j < h; ++j) {
            const double gr = g[j].real(), gi = g[j].imag();
            ore[j] += s.f_re[j] * gr - s.f_im[j] * gi;
            oim[j] += s.f_re[j] * gi + s.f_im[j] * gr;
          }
        }
      }
    }
    for (size_t c = 0; c < gs; ++c) { // inverse, untwist, torus rounding, accumulate
      vd *ore = s.o_re.data() + c * h, *oim = s.o_im.data() + c * h;
      pl.inv(ore, oim);
      vu *a = s.acc.data() + c * N;
      for (size_t j = 0; j < h; ++j) {
        const double wr = pl.tw_re[j] * inv_h, wi = -pl.tw_im[j] * inv_h;
        a[j] += from_torus(ore[j] * wr - oim[j] * wi);
        a[j + h] += from_torus(ore[j] * wi + oim[j] * wr);
      }
    }
  }
  for (size_t w = 0; w < count; ++w) { // sample extraction, nth = 0
    u64 *o = out + (first + w) * (k * N + 1);
    for (size_t p = 0; p < k; ++p) {
      o[p * N] = s.acc[p * N][w];
      for (size_t j = 1; j < N; ++j) o[p * N + j] = (u64)0 - s.acc[p * N + N - j][w];
    }
    o[k * N] = s.acc[k * N][w];
  }
}

} // namespace

extern "C" {

// out[num][k*N+1] <- PBS(cts[num][n+1]); fbsk in the oracle's Fourier layout
void fast_bootstrap_batch(u64 *out, const u64 *cts, const u64 *luts, const u64 *lut_idx,
                          const std::complex<double> *fbsk, unsigned n, unsigned k, size_t N,
                          unsigned base_log, unsigned level, size_t num) {
  const Plan pl(N);
  const size_t batches = (num + W - 1) / W;
#pragma omp parallel
  {
    Scratch s;
#pragma omp for schedule(dynamic, 1)
    for (size_t b = 0; b < batches; ++b)
      bootstrap_batch(out, cts, luts, lut_idx, fbsk, n, k, N, base_log, level, b * W,
                      std::min<size_t>(W, num - b * W), pl, s);
  }
}

// out[num][n_out+1]; KSK rows streamed once per 8 ciphertexts
void fast_keyswitch_batch(u64 *out, const u64 *cts, const u64 *ksk, unsigned n_in, unsigned n_out,
                          unsigned base_log, unsigned level, size_t num) {
  const size_t out_size = (size_t)n_out + 1;
  const size_t batches = (num + W - 1) / W;
  const unsigned shift = 64 - base_log * level;
  const u64 mask = ((u64)1 << base_log) - 1;
#pragma omp parallel for schedule(dynamic, 1)
  for (size_t b = 0; b < batches; ++b) {
    const size_t first = b * W, count = std::min<size_t>(W, num - first);
    for (size_t w = 0; w < count; ++w) {
      u64 *o = out + (first + w) * out_size;
      std::fill(o, o + out_size, (u64)0);
      o[n_out] = cts[(first + w) * (n_in + 1) + n_in];
    }
    for (unsigned i = 0; i < n_in; ++i) {
      u64 state[W];
      for (size_t w = 0; w < count; ++w)
        state[w] = closest_representable(cts[(first + w) * (n_in + 1) + i], level, base_log) >> shift;
      for (unsigned j = 0; j < level; ++j) {
        const u64 *row = ksk + ((size_t)i * level + j) * out_size;
        for (size_t w = 0; w < count; ++w) {
          u64 res = state[w] & mask;
          state[w] >>= base_log;
          u64 carry = ((res - 1) | state[w]) & res;
          carry >>= (base_log - 1);
          state[w] += carry;
          res -= carry << base_log;
          if (res == 0) continue;
          u64 *o = out + (first + w) * out_size;
          for (size_t c = 0; c < out_size; ++c) o[c] -= row[c] * res;
        }
      }
    }
  }
}

int fast_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void fast_set_num_threads(int t) {
#ifdef _OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}
int fast_simd_bits() {
#if defined(__AVX512F__)
  return 512;
#elif defined(__AVX2__)
  return 256;
#else
  return 128;
#endif
}

} // extern "C"
