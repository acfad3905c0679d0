// wop_oracle.inl — CPU restatement of the reference's WoP-PBS path, included at
// the end of tfhe_oracle.cpp (it uses that file's static helpers).  TEST INFRASTRUCTURE ONLY.
//
// Restates (paths relative to /root/reference; TFHE = third_party/tfhe-rs/tfhe/src/core_crypto):
//   * extract_bits                      TFHE/fft_impl/fft64/crypto/wop_pbs/mod.rs:63-222
//   * homomorphic_shift_boolean         ...:366-441
//   * circuit_bootstrap_boolean         ...:240-341
//   * cmux_tree_memory_optimized        ...:465-587   (evaluated layer by layer: the same cmuxes on
//                                        the same operands, only the schedule differs)
//   * blind_rotate_assign (GGSW list)   ...:854-880,  cmux = ggsw.rs `cmux` (ct1 -= ct0; ct0 += G x ct1)
//   * vertical_packing                  ...:776-841
//   * circuit_bootstrap_boolean_vertical_packing ...:642-737
//   * private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext
//                                       TFHE/algorithms/lwe_private_functional_packing_keyswitch.rs:20-88
//   * generate_circuit_bootstrap_lwe_pfpksk_list / generate_lwe_private_functional_packing_keyswitch_key
//                                       TFHE/algorithms/lwe_wopbs.rs:81-160,
//                                       lwe_private_functional_packing_keyswitch_key_generation.rs:19-131
//   * the C entries and the runtime wrapper that chains them:
//       backends/concrete-cpu/implementation/src/c_api/wop_pbs.rs:103-412
//       compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:855-997 (memref_wop_pbs_crt_buffer)

namespace {

// The private functional packing keyswitch is the LWE keyswitch formula with the BODY decomposed
// like a mask element and a GLWE as output (no body pass-through): keyswitch_one() on the input
// extended by a zero "body" does exactly that (n_in + 1 inputs, (k+1)N - 1 + 1 output words).
void pfpks_one(u64 *glwe_out, const u64 *lwe_in, const u64 *pfpksk, unsigned n_in, size_t glwe_words,
               unsigned base_log, unsigned level_count) {
  std::vector<u64> ext(lwe_in, lwe_in + n_in + 1);
  ext.push_back(0);
  keyswitch_one(glwe_out, ext.data(), pfpksk, n_in + 1, (unsigned)(glwe_words - 1), base_log,
                level_count);
}

// PBS with an explicit accumulator (FourierLweBootstrapKey::bootstrap, bootstrap.rs:336-368)
void bootstrap_one(u64 *lwe_out, const u64 *lwe_in, const u64 *accumulator, const c64 *fbsk,
                   unsigned n, unsigned k, size_t N, unsigned base_log, unsigned level_count,
                   const NegFft &fft, PbsScratch &s) {
  std::vector<u64> acc(accumulator, accumulator + (size_t)(k + 1) * N);
  blind_rotate_assign(acc.data(), lwe_in, fbsk, n, k, N, base_log, level_count, fft, s);
  sample_extract<u64>(lwe_out, acc.data(), k, N, 0);
}

void extract_bits_one(u64 *out_list, const u64 *lwe_in, const c64 *fbsk, const u64 *ksk,
                      unsigned n_small, unsigned k, size_t N, unsigned bsk_base_log,
                      unsigned bsk_level, unsigned ks_base_log, unsigned ks_level,
                      unsigned delta_log, unsigned nbits) {
  const NegFft &fft = get_plan(N);
  PbsScratch s;
  const unsigned big = (unsigned)(k * N);
  std::vector<u64> in_buf(lwe_in, lwe_in + big + 1), shifted(big + 1), ks_out(n_small + 1),
      acc((size_t)(k + 1) * N), pbs_out(big + 1);
  for (unsigned bit_idx = 0; bit_idx < nbits; ++bit_idx) { // output index nbits-1-bit_idx: MSB at 0
    const unsigned sh = 64 - delta_log - bit_idx - 1;
    for (unsigned i = 0; i <= big; ++i) shifted[i] = in_buf[i] << sh;
    keyswitch_one(ks_out.data(), shifted.data(), ksk, big, n_small, ks_base_log, ks_level);
    std::copy(ks_out.begin(), ks_out.end(), out_list + (size_t)(nbits - 1 - bit_idx) * (n_small + 1));
    if (bit_idx == nbits - 1) break;
    ks_out[n_small] += (u64)1 << 62; // q/4: centre the error of the negacyclic LUT
    std::fill(acc.begin(), acc.end(), (u64)0);
    for (size_t c = 0; c < N; ++c) acc[(size_t)k * N + c] = (u64)0 - ((u64)1 << (delta_log - 1 + bit_idx));
    bootstrap_one(pbs_out.data(), ks_out.data(), acc.data(), fbsk, n_small, k, N, bsk_base_log,
                  bsk_level, fft, s);
    pbs_out[big] += (u64)1 << (delta_log + bit_idx - 1);
    for (unsigned i = 0; i <= big; ++i) in_buf[i] -= pbs_out[i];
  }
}

void circuit_bootstrap_one(u64 *ggsw_out /*[l_cbs][k+1][(k+1)N]*/, const u64 *lwe_in, const c64 *fbsk,
                           const u64 *pfpksk_list, unsigned n_small, unsigned k, size_t N,
                           unsigned bsk_base_log, unsigned bsk_level, unsigned pks_base_log,
                           unsigned pks_level, unsigned cbs_base_log, unsigned cbs_level,
                           unsigned delta_log) {
  const NegFft &fft = get_plan(N);
  PbsScratch s;
  const unsigned big = (unsigned)(k * N);
  const size_t gs = k + 1, glwe_words = gs * N;
  const size_t pfpksk_words = (size_t)(big + 1) * pks_level * glwe_words;
  std::vector<u64> shifted(n_small + 1), acc(glwe_words), bs_out(big + 1);
  for (unsigned out_idx = 0; out_idx < cbs_level; ++out_idx) {
    const unsigned level = cbs_level - out_idx;
    // homomorphic_shift_boolean
    for (unsigned i = 0; i <= n_small; ++i) shifted[i] = lwe_in[i] << (64 - delta_log - 1);
    shifted[n_small] += (u64)1 << 62;
    std::fill(acc.begin(), acc.end(), (u64)0);
    const u64 alpha = (u64)1 << (64 - 1 - cbs_base_log * level);
    for (size_t c = 0; c < N; ++c) acc[(size_t)k * N + c] = (u64)0 - alpha;
    bootstrap_one(bs_out.data(), shifted.data(), acc.data(), fbsk, n_small, k, N, bsk_base_log,
                  bsk_level, fft, s);
    bs_out[big] += alpha;
    for (size_t r = 0; r < gs; ++r)
      pfpks_one(ggsw_out + ((size_t)out_idx * gs + r) * glwe_words, bs_out.data(),
                pfpksk_list + r * pfpksk_words, big, glwe_words, pks_base_log, pks_level);
  }
}

// ct0 <- ct0 + ggsw (x) (ct1 - ct0)
void cmux_assign(u64 *ct0, const u64 *ct1, const c64 *ggsw, unsigned k, size_t N, unsigned base_log,
                 unsigned level, const NegFft &fft, PbsScratch &s) {
  const size_t words = (size_t)(k + 1) * N;
  std::vector<u64> diff(words);
  for (size_t i = 0; i < words; ++i) diff[i] = ct1[i] - ct0[i];
  add_external_product_assign(ct0, ggsw, diff.data(), k, N, base_log, level, fft, s);
}

// lut: `polys` polynomials of N coefficients; ggsw_list: `nggsw` Fourier GGSWs, MSB first
void vertical_packing_one(u64 *lwe_out, const u64 *lut, size_t polys, const c64 *ggsw_list,
                          unsigned nggsw, unsigned k, size_t N, unsigned base_log, unsigned level) {
  const NegFft &fft = get_plan(N);
  PbsScratch s;
  const size_t gs = k + 1, words = gs * N, ggsw_stride = (size_t)level * gs * gs * (N / 2);
  unsigned log_polys = 0;
  while (((size_t)2 << log_polys) <= polys) ++log_polys; // floor(log2(polys))
  const unsigned tree = log_polys > nggsw ? 0 : log_polys;
  // cmux tree over the first `tree` GGSWs (used last to first), trivial GLWEs at the leaves
  std::vector<u64> result(words, 0);
  if (tree == 0) {
    std::copy(lut, lut + N, result.begin() + (size_t)k * N);
  } else {
    size_t count = (size_t)1 << tree;
    std::vector<u64> layer(count * words, 0);
    for (size_t p = 0; p < count; ++p) std::copy(lut + p * N, lut + (p + 1) * N, layer.begin() + p * words + (size_t)k * N);
    for (unsigned j = 0; j < tree; ++j) {
      const c64 *g = ggsw_list + (size_t)(tree - 1 - j) * ggsw_stride;
      for (size_t p = 0; p < count / 2; ++p) {
        u64 *t0 = layer.data() + (2 * p) * words;
        cmux_assign(t0, layer.data() + (2 * p + 1) * words, g, k, N, base_log, level, fft, s);
        if (p) std::copy(t0, t0 + words, layer.data() + p * words);
      }
      count /= 2;
    }
    std::copy(layer.begin(), layer.begin() + words, result.begin());
  }
  // blind rotation by the remaining GGSWs, last one first, monomial degrees 1, 2, 4, ...
  std::vector<u64> ct1(words);
  size_t degree = 1;
  for (int g = (int)nggsw - 1; g >= (int)tree; --g, degree <<= 1) {
    for (size_t p = 0; p < gs; ++p) monomial_div<u64>(ct1.data() + p * N, result.data() + p * N, N, degree);
    cmux_assign(result.data(), ct1.data(), ggsw_list + (size_t)g * ggsw_stride, k, N, base_log, level, fft, s);
  }
  sample_extract<u64>(lwe_out, result.data(), k, N, 0);
}

} // namespace

extern "C" {

// (k+1) private functional packing keyswitch keys, each [n_in + 1][level][(k+1)N]; key r < k
// packs x -> -x * S_r, the last one x -> x (polynomial -1, function -x)
void orc_gen_cbs_pfpksk_list(u64 *out, const u64 *lwe_key_in, unsigned n_in, const u64 *glwe_key,
                             unsigned k, size_t N, unsigned base_log, unsigned level_count,
                             double sigma, u64 seed) {
  const NegFft &fft = get_plan(N);
  const size_t gs = k + 1, glwe_words = gs * N;
  const long rows = (long)gs * (n_in + 1) * level_count;
#pragma omp parallel
  {
    std::vector<u64> prod(N);
#pragma omp for schedule(dynamic, 8)
    for (long idx = 0; idx < rows; ++idx) {
      const size_t key = idx / ((long)(n_in + 1) * level_count);
      const size_t i = (idx / level_count) % (n_in + 1);
      const unsigned j = idx % level_count, lvl = level_count - j; // memory level j <-> level l - j
      Rng rng(seed, 0x5000000ULL + (u64)idx);
      const u64 key_bit = i < n_in ? lwe_key_in[i] : ~(u64)0;
      const u64 factor = ((u64)0 - key_bit) << (64 - base_log * lvl); // f(1) * key_bit, f = -x
      u64 *glwe = out + (size_t)idx * glwe_words, *body = glwe + (size_t)k * N;
      for (size_t c = 0; c < N; ++c) body[c] = rng.noise(sigma);
      if (key < k) {
        for (size_t c = 0; c < N; ++c) body[c] += glwe_key[key * N + c] * factor;
      } else {
        body[0] += (~(u64)0) * factor; // the polynomial -1
      }
      for (size_t p = 0; p < k; ++p) {
        u64 *a = glwe + p * N;
        for (size_t c = 0; c < N; ++c) a[c] = rng.next();
        negacyclic_mul_binary(prod.data(), a, glwe_key + p * N, N, fft);
        for (size_t c = 0; c < N; ++c) body[c] += prod[c];
      }
    }
  }
}

void orc_extract_bits(u64 *out_list, const u64 *lwe_in, const double *fbsk, const u64 *ksk,
                      unsigned n_small, unsigned k, size_t N, unsigned bsk_base_log,
                      unsigned bsk_level, unsigned ks_base_log, unsigned ks_level,
                      unsigned delta_log, unsigned nbits) {
  extract_bits_one(out_list, lwe_in, reinterpret_cast<const c64 *>(fbsk), ksk, n_small, k, N,
                   bsk_base_log, bsk_level, ks_base_log, ks_level, delta_log, nbits);
}

void orc_circuit_bootstrap_boolean(u64 *ggsw_out, const u64 *lwe_in, const double *fbsk,
                                   const u64 *pfpksk_list, unsigned n_small, unsigned k, size_t N,
                                   unsigned bsk_base_log, unsigned bsk_level, unsigned pks_base_log,
                                   unsigned pks_level, unsigned cbs_base_log, unsigned cbs_level,
                                   unsigned delta_log) {
  circuit_bootstrap_one(ggsw_out, lwe_in, reinterpret_cast<const c64 *>(fbsk), pfpksk_list, n_small,
                        k, N, bsk_base_log, bsk_level, pks_base_log, pks_level, cbs_base_log,
                        cbs_level, delta_log);
}

// concrete_cpu_circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_u64 (wop_pbs.rs:239-412)
void orc_cbs_vertical_packing(u64 *out /*[lut_count][kN+1]*/, const u64 *bits_in /*[nbits][n_small+1]*/,
                              const u64 *luts /*[lut_count][lut_size]*/, const double *fbsk,
                              const u64 *pfpksk_list, unsigned n_small, unsigned k, size_t N,
                              unsigned bsk_base_log, unsigned bsk_level, unsigned pks_base_log,
                              unsigned pks_level, unsigned cbs_base_log, unsigned cbs_level,
                              unsigned nbits, size_t lut_count, size_t lut_size) {
  const NegFft &fft = get_plan(N);
  const size_t gs = k + 1, ggsw_words = (size_t)cbs_level * gs * gs * N;
  std::vector<c64> ggsw_list((size_t)nbits * ggsw_words / 2);
#pragma omp parallel for schedule(dynamic, 1)
  for (long b = 0; b < (long)nbits; ++b) {
    std::vector<u64> ggsw(ggsw_words);
    circuit_bootstrap_one(ggsw.data(), bits_in + (size_t)b * (n_small + 1),
                          reinterpret_cast<const c64 *>(fbsk), pfpksk_list, n_small, k, N,
                          bsk_base_log, bsk_level, pks_base_log, pks_level, cbs_base_log, cbs_level,
                          63);
    for (size_t p = 0; p < ggsw_words / N; ++p) // fill_with_forward_fourier (ggsw.rs:253-276)
      fft.forward_as_torus(ggsw_list.data() + ((size_t)b * ggsw_words + p * N) / 2, ggsw.data() + p * N);
  }
  // luts shorter than a polynomial are zero-extended to one polynomial each (wop_pbs.rs:300-312)
  const size_t polys = lut_size < N ? 1 : lut_size / N;
#pragma omp parallel for schedule(dynamic, 1)
  for (long l = 0; l < (long)lut_count; ++l) {
    std::vector<u64> lut(polys * N, 0);
    std::copy(luts + (size_t)l * lut_size, luts + (size_t)(l + 1) * lut_size, lut.begin());
    vertical_packing_one(out + (size_t)l * (k * N + 1), lut.data(), polys, ggsw_list.data(), nbits, k,
                         N, cbs_base_log, cbs_level);
  }
}

// memref_wop_pbs_crt_buffer (wrappers.cpp:855-997): `blocks` CRT blocks in, `blocks` out
void orc_wop_pbs_crt(u64 *out, const u64 *in, const u64 *lut_ct /*[blocks][2^total_bits]*/,
                     const u64 *crt_decomp, unsigned blocks, const double *fbsk, const u64 *ksk,
                     const u64 *pfpksk_list, unsigned n_small, unsigned k, size_t N,
                     unsigned bsk_base_log, unsigned bsk_level, unsigned ks_base_log,
                     unsigned ks_level, unsigned pks_base_log, unsigned pks_level,
                     unsigned cbs_base_log, unsigned cbs_level) {
  const size_t big_size = (size_t)k * N + 1, small_size = (size_t)n_small + 1;
  std::vector<unsigned> nbits(blocks);
  unsigned total = 0;
  for (unsigned i = 0; i < blocks; ++i) {
    nbits[i] = (unsigned)std::ceil(std::log2((double)crt_decomp[i]));
    total += nbits[i];
  }
  std::vector<u64> bits((size_t)total * small_size, 0), copy(in, in + (size_t)blocks * big_size);
  std::vector<unsigned> offset(blocks);
  {
    unsigned off = 0;
    for (int i = (int)blocks - 1; i >= 0; --i) {
      offset[i] = off;
      off += nbits[i];
    }
  }
#pragma omp parallel for schedule(dynamic, 1)
  for (long i = 0; i < (long)blocks; ++i) {
    const unsigned nb = nbits[i], delta_log = 64 - nb;
    u64 *blk = copy.data() + (size_t)i * big_size;
    blk[big_size - 1] -= ((u64)1 << (64 - nb - 1)) - ((u64)1 << (64 - nb - 5)); // ct - delta/2 + delta/2^4
    extract_bits_one(bits.data() + (size_t)offset[i] * small_size, blk,
                     reinterpret_cast<const c64 *>(fbsk), ksk, n_small, k, N, bsk_base_log, bsk_level,
                     ks_base_log, ks_level, delta_log, nb);
  }
  orc_cbs_vertical_packing(out, bits.data(), lut_ct, fbsk, pfpksk_list, n_small, k, N, bsk_base_log,
                           bsk_level, pks_base_log, pks_level, cbs_base_log, cbs_level, total, blocks,
                           (size_t)1 << total);
}

} // extern "C"
