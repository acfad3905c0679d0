// wop_pbs.cu — K7: WoP-PBS on the device: bit extraction, circuit bootstrap and
// vertical packing, built from K1 (keyswitch), K3 (fused PBS) and one new kernel (the CMUX of the
// vertical-packing tree, pbs.cu cmux_kernel).
//
// Replaces, for the GPU runtime, what the reference only has on the CPU:
//   compilers/concrete-compiler/compiler/lib/Runtime/wrappers.cpp:855-997 (memref_wop_pbs_crt_buffer)
//   -> concrete_cpu_extract_bit_lwe_ciphertext_u64 and
//      concrete_cpu_circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_u64
//      (backends/concrete-cpu/implementation/src/c_api/wop_pbs.rs:103-412)
//   -> tfhe-rs core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs:63-222 (extract_bits), :240-341
//      (circuit_bootstrap_boolean), :366-441 (homomorphic_shift_boolean), :465-587 (cmux tree),
//      :776-841 (vertical_packing), :854-880 (blind_rotate_assign) and
//      core_crypto/algorithms/lwe_private_functional_packing_keyswitch.rs:20-88.
//
// How the pieces map onto the existing kernels (all on one stream, nothing visits the host):
//   * extract_bits: the blocks of the CRT decomposition are independent, the bits of one block are
//     strictly sequential; step s extracts bit s of EVERY block that still has one: a row-wise
//     shift, ONE keyswitch launch whose output index array drops each result at its final row of
//     the bit list, ONE PBS launch (one constant LUT per block), a row-wise subtract;
//   * circuit bootstrap: the l_cbs homomorphic shifts of all extracted bits are ONE PBS launch
//     (bits x levels samples, LUT index = level); the private functional packing keyswitch is the
//     LWE keyswitch formula with the body decomposed like a mask element and a GLWE as output, so
//     it runs on K1 unchanged: input rows extended by a zero body, n_in = kN + 1,
//     n_out + 1 = (k+1)N, one launch per key r = 0..k, the output index array lays the rows out
//     as the GGSW list [bit][level][row];
//   * GGSW list -> Fourier in place: the selectors of the CMUX tree in the generic layout
//     (cmux_kernel), the selectors of the final blind rotation in the layout of the PBS kernel
//     family that serves (k, N, l_cbs);
//   * vertical packing: tree layers with cmux_kernel (all luts and all pairs of a layer in one
//     launch), then the blind rotation by the remaining bits as ONE launch of the fused PBS kernel:
//     a blind rotation by a GGSW list is a bootstrap whose "LWE mask" is the fixed sequence
//     X^(-1), X^(-2), X^(-4), ... (a_i = (2N - 2^i) * 2^64 / 2N mod-switches to exactly that, the
//     body 0 to no initial rotation), whose LUT is the tree result and whose "bootstrapping key"
//     is the GGSW list in the order the reference applies it (last bit first).
#include "common.cuh"
#include "../../include/concrete_b200.h"
#include <algorithm>
#include <cmath>
#include <vector>

void cb_keyswitch_launch(cudaStream_t stream, u64 *out, const u64 *out_idx, const u64 *in,
                         const u64 *in_idx, const u64 *ksk, u32 n_in, u32 n_out, u32 base_log,
                         u32 level, u32 num_samples);
void cb_pbs_scratch(cudaStream_t stream, int gpu, int8_t **buffer, u32 k, u32 N, u32 level,
                    u32 count, bool allocate);
void cb_pbs_cleanup(cudaStream_t stream, int gpu, int8_t **buffer);
void cb_pbs_launch(cudaStream_t stream, int gpu, u64 *lwe_out, const u64 *out_idx,
                   const u64 *lut_vector, const u64 *lut_idx, const u64 *lwe_in,
                   const u64 *in_idx, const double2 *bsk, int8_t *buffer, u32 n, u32 k, u32 N,
                   u32 base_log, u32 level, u32 num_samples, u32 lut_count, u32 lut_stride);
void cb_bsk_convert(cudaStream_t stream, int gpu, void *dest, const void *src_host, u32 n, u32 k,
                    u32 level, u32 N, bool src_on_device);
void cb_ggsw_convert_generic(cudaStream_t stream, int gpu, void *data, size_t polys, u32 k, u32 N);
void cb_cmux_launch(cudaStream_t stream, int gpu, u64 *out, const u64 *in, const double2 *ggsw,
                    u32 k, u32 N, u32 base_log, u32 level, u32 pairs);

namespace {

constexpr int WT = 256;
inline u32 grid_for(size_t total) {
  const size_t g = (total + WT - 1) / WT;
  return (u32)std::min<size_t>(std::max<size_t>(g, 1), 148 * 16);
}

// dst[r][i] = src[rows[r]][i] << shift[r]   (row length `width`)
__global__ void k_shift_rows(u64 *dst, const u64 *src, const u64 *rows, const u32 *shift, u32 width,
                             size_t total) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / width, c = i - r * width;
    dst[i] = src[rows[r] * width + c] << shift[r];
  }
}
// dst[r][i] = src[rows[r]][i] (row length src_width, dst rows dst_width >= src_width, the rest 0),
// body (element src_width - 1) += add[r]
__global__ void k_gather_rows_add_body(u64 *dst, const u64 *src, const u64 *rows, const u64 *add,
                                       u32 src_width, u32 dst_width, size_t total) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / dst_width, c = i - r * dst_width;
    u64 v = 0;
    if (c < src_width) {
      v = src[rows[r] * src_width + c];
      if (c == src_width - 1) v += add[r];
    }
    dst[i] = v;
  }
}
// acc[rows[r]][i] -= pbs[r][i], with pbs body += add[r] first (extract_bits: remove the bit)
__global__ void k_sub_rows(u64 *acc, const u64 *pbs, const u64 *rows, const u64 *add, u32 width,
                           size_t total) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / width, c = i - r * width;
    u64 v = pbs[i];
    if (c == width - 1) v += add[r];
    acc[rows[r] * width + c] -= v;
  }
}
// trivial GLWE accumulators with a constant body: lut[r] = (0, ..., 0, value[r] in every coefficient)
__global__ void k_const_luts(u64 *luts, const u64 *value, u32 k, u32 N, size_t total) {
  const size_t words = (size_t)(k + 1) * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / words, c = i - r * words;
    luts[i] = c >= (size_t)k * N ? value[r] : 0;
  }
}
// leaves of the CMUX tree: trivial GLWEs whose bodies are the polynomials of the luts
__global__ void k_tree_leaves(u64 *glwe, const u64 *luts, u32 k, u32 N, size_t lut_size,
                              size_t polys_per_lut, size_t total) {
  const size_t words = (size_t)(k + 1) * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t g = i / words, c = i - g * words; // g = lut * polys_per_lut + poly
    u64 v = 0;
    if (c >= (size_t)k * N) {
      const size_t lut = g / polys_per_lut, poly = g - lut * polys_per_lut;
      const size_t pos = poly * N + (c - (size_t)k * N);
      if (pos < lut_size) v = luts[lut * lut_size + pos];
    }
    glwe[i] = v;
  }
}

template <class T> struct DevArr { // small host array mirrored on the device, stream-ordered
  T *d = nullptr;
  size_t n = 0;
  void upload(cudaStream_t s, const std::vector<T> &h) {
    n = h.size();
    if (!n) return;
    CB_CHECK(cudaMallocAsync((void **)&d, n * sizeof(T), s));
    CB_CHECK(cudaMemcpyAsync(d, h.data(), n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void alloc(cudaStream_t s, size_t count) {
    n = count;
    if (n) CB_CHECK(cudaMallocAsync((void **)&d, n * sizeof(T), s));
  }
  void free(cudaStream_t s) {
    if (d) CB_CHECK(cudaFreeAsync(d, s));
    d = nullptr;
  }
};

struct WopShape {
  u32 n_small, k, N;
  u32 bsk_level, bsk_base_log, ks_level, ks_base_log, pks_level, pks_base_log, cbs_level,
      cbs_base_log;
};

// bits_out: device [total_bits][n_small + 1], MSB of the LAST block first (wrappers.cpp:903-909)
void extract_bits_blocks(cudaStream_t stream, int gpu, u64 *bits_out, const u64 *lwe_in,
                         const std::vector<u32> &nbits, const double2 *fbsk, const u64 *ksk,
                         const WopShape &p, bool rounding_trick) {
  const u32 blocks = (u32)nbits.size(), big = p.k * p.N, big_size = big + 1,
            small_size = p.n_small + 1;
  u32 max_bits = 0;
  std::vector<u32> offset(blocks);
  {
    u32 off = 0;
    for (int i = (int)blocks - 1; i >= 0; --i) {
      offset[i] = off;
      off += nbits[i];
      max_bits = std::max(max_bits, nbits[i]);
    }
  }
  // private copy of the input (optionally with the rounding trick on the body: ct - delta/2 + delta/2^4)
  DevArr<u64> acc, shifted, pbs_in, pbs_out, luts, iota_rows, sub0;
  acc.alloc(stream, (size_t)blocks * big_size);
  {
    std::vector<u64> rows(blocks), add(blocks);
    for (u32 i = 0; i < blocks; ++i) {
      rows[i] = i;
      add[i] = rounding_trick // wrappers.cpp:928-931
                   ? (u64)0 - (((u64)1 << (64 - nbits[i] - 1)) - ((u64)1 << (64 - nbits[i] - 5)))
                   : (u64)0;
    }
    iota_rows.upload(stream, rows);
    sub0.upload(stream, add);
    k_gather_rows_add_body<<<grid_for(acc.n), WT, 0, stream>>>(acc.d, lwe_in, iota_rows.d, sub0.d,
                                                               big_size, big_size, acc.n);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  shifted.alloc(stream, (size_t)blocks * big_size);
  pbs_in.alloc(stream, (size_t)blocks * small_size);
  pbs_out.alloc(stream, (size_t)blocks * big_size);
  luts.alloc(stream, (size_t)blocks * (p.k + 1) * p.N);
  int8_t *scratch = nullptr;
  cb_pbs_scratch(stream, gpu, &scratch, p.k, p.N, p.bsk_level, blocks, true);
  std::vector<DevArr<u64>> keep64;
  std::vector<DevArr<u32>> keep32;
  for (u32 bit = 0; bit < max_bits; ++bit) {
    std::vector<u64> rows, out_rows, pbs_rows, q4, lutv, add_back, lut_idx, seq;
    std::vector<u32> shift;
    for (u32 i = 0; i < blocks; ++i) {
      if (nbits[i] <= bit) continue;
      const u32 delta_log = 64 - nbits[i];
      rows.push_back(i);
      shift.push_back(64 - delta_log - bit - 1);
      out_rows.push_back(offset[i] + (nbits[i] - 1 - bit));
      if (bit + 1 < nbits[i]) { // not the last bit of this block: a PBS removes the bit
        pbs_rows.push_back(out_rows.back());
        q4.push_back((u64)1 << 62);
        lutv.push_back((u64)0 - ((u64)1 << (delta_log - 1 + bit)));
        add_back.push_back((u64)1 << (delta_log + bit - 1));
      }
    }
    const u32 act = (u32)rows.size();
    DevArr<u64> d_rows, d_out_rows, d_seq;
    DevArr<u32> d_shift;
    seq.resize(act);
    for (u32 r = 0; r < act; ++r) seq[r] = r;
    d_rows.upload(stream, rows);
    d_out_rows.upload(stream, out_rows);
    d_seq.upload(stream, seq);
    d_shift.upload(stream, shift);
    keep64.insert(keep64.end(), {d_rows, d_out_rows, d_seq});
    keep32.push_back(d_shift);
    k_shift_rows<<<grid_for((size_t)act * big_size), WT, 0, stream>>>(
        shifted.d, acc.d, d_rows.d, d_shift.d, big_size, (size_t)act * big_size);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    cb_keyswitch_launch(stream, bits_out, d_out_rows.d, shifted.d, d_seq.d, ksk, big, p.n_small,
                        p.ks_base_log, p.ks_level, act);
    const u32 np = (u32)pbs_rows.size();
    if (np == 0) continue;
    // blocks that go on are a prefix-compatible subset: rows_p[r] = block id of PBS sample r
    std::vector<u64> blk_of;
    for (u32 i = 0; i < blocks; ++i)
      if (nbits[i] > bit + 1) blk_of.push_back(i);
    DevArr<u64> d_pbs_rows, d_q4, d_lutv, d_add_back, d_blk_of, d_seqp;
    std::vector<u64> seqp(np);
    for (u32 r = 0; r < np; ++r) seqp[r] = r;
    d_pbs_rows.upload(stream, pbs_rows);
    d_q4.upload(stream, q4);
    d_lutv.upload(stream, lutv);
    d_add_back.upload(stream, add_back);
    d_blk_of.upload(stream, blk_of);
    d_seqp.upload(stream, seqp);
    keep64.insert(keep64.end(), {d_pbs_rows, d_q4, d_lutv, d_add_back, d_blk_of, d_seqp});
    k_gather_rows_add_body<<<grid_for((size_t)np * small_size), WT, 0, stream>>>(
        pbs_in.d, bits_out, d_pbs_rows.d, d_q4.d, small_size, small_size, (size_t)np * small_size);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    k_const_luts<<<grid_for((size_t)np * (p.k + 1) * p.N), WT, 0, stream>>>(
        luts.d, d_lutv.d, p.k, p.N, (size_t)np * (p.k + 1) * p.N);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    cb_pbs_launch(stream, gpu, pbs_out.d, d_seqp.d, luts.d, d_seqp.d, pbs_in.d, d_seqp.d, fbsk,
                  scratch, p.n_small, p.k, p.N, p.bsk_base_log, p.bsk_level, np, 1, 1);
    k_sub_rows<<<grid_for((size_t)np * big_size), WT, 0, stream>>>(
        acc.d, pbs_out.d, d_blk_of.d, d_add_back.d, big_size, (size_t)np * big_size);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
  }
  cb_pbs_cleanup(stream, gpu, &scratch);
  for (auto &a : keep64) a.free(stream);
  for (auto &a : keep32) a.free(stream);
  for (DevArr<u64> *a : {&acc, &shifted, &pbs_in, &pbs_out, &luts, &iota_rows, &sub0}) a->free(stream);
}

// bits_in: device [nbits][n_small + 1]; luts: device [lut_count][lut_size]; out [lut_count][kN + 1]
void cbs_vertical_packing(cudaStream_t stream, int gpu, u64 *out, const u64 *bits_in, u32 nbits,
                          const u64 *luts, u32 lut_count, size_t lut_size, const double2 *fbsk,
                          const u64 *pfpksk, const WopShape &p) {
  const u32 big = p.k * p.N, big_size = big + 1, small_size = p.n_small + 1, gs = p.k + 1;
  const size_t glwe_words = (size_t)gs * p.N, ggsw_words = (size_t)p.cbs_level * gs * glwe_words;
  const size_t pfpksk_words = (size_t)(big + 1) * p.pks_level * glwe_words;
  const size_t polys_per_lut = lut_size < p.N ? 1 : lut_size / p.N;
  u32 log_polys = 0;
  while (((size_t)2 << log_polys) <= polys_per_lut) ++log_polys;
  const u32 tree = log_polys > nbits ? 0 : log_polys, nbr = nbits - tree;
  if (tree == 0 && polys_per_lut != 1)
    CB_PANIC("Cuda error (WoP-PBS): %zu lut polynomials need %u selector bits, got %u",
             polys_per_lut, log_polys, nbits);
  if (nbr == 0) CB_PANIC("Cuda error (WoP-PBS): no bit left for the blind rotation");

  // ---- circuit bootstrap of every bit: l_cbs homomorphic shifts per bit in ONE PBS launch
  const u32 ns = nbits * p.cbs_level;
  DevArr<u64> pbs_in, pbs_out, ext, lut_cbs, ggsw, d_in_idx, d_lut_idx, d_seq, d_q4, d_bitrows,
      d_alpha_neg, d_alpha;
  {
    std::vector<u64> in_idx(ns), lut_idx(ns), seq(ns), q4(nbits, (u64)1 << 62), bitrows(nbits),
        alpha_neg(p.cbs_level), alpha(ns);
    for (u32 b = 0; b < nbits; ++b) {
      bitrows[b] = b;
      for (u32 o = 0; o < p.cbs_level; ++o) {
        const u32 level = p.cbs_level - o;
        in_idx[b * p.cbs_level + o] = b;
        lut_idx[b * p.cbs_level + o] = o;
        seq[b * p.cbs_level + o] = b * p.cbs_level + o;
        alpha[b * p.cbs_level + o] = (u64)1 << (63 - p.cbs_base_log * level);
        alpha_neg[o] = (u64)0 - ((u64)1 << (63 - p.cbs_base_log * level));
      }
    }
    d_in_idx.upload(stream, in_idx);
    d_lut_idx.upload(stream, lut_idx);
    d_seq.upload(stream, seq);
    d_q4.upload(stream, q4);
    d_bitrows.upload(stream, bitrows);
    d_alpha_neg.upload(stream, alpha_neg);
    d_alpha.upload(stream, alpha);
  }
  pbs_in.alloc(stream, (size_t)nbits * small_size);
  pbs_out.alloc(stream, (size_t)ns * big_size);
  ext.alloc(stream, (size_t)ns * (big_size + 1));
  lut_cbs.alloc(stream, (size_t)p.cbs_level * glwe_words);
  ggsw.alloc(stream, (size_t)nbits * ggsw_words);
  // the bit sits in the MSB already (delta_log = 63: shift by 0); + q/4 on the body
  k_gather_rows_add_body<<<grid_for(pbs_in.n), WT, 0, stream>>>(pbs_in.d, bits_in, d_bitrows.d, d_q4.d,
                                                              small_size, small_size, pbs_in.n);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
  k_const_luts<<<grid_for(lut_cbs.n), WT, 0, stream>>>(lut_cbs.d, d_alpha_neg.d, p.k, p.N, lut_cbs.n);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
  int8_t *scratch = nullptr;
  cb_pbs_scratch(stream, gpu, &scratch, p.k, p.N, p.bsk_level, ns, true);
  cb_pbs_launch(stream, gpu, pbs_out.d, d_seq.d, lut_cbs.d, d_lut_idx.d, pbs_in.d, d_in_idx.d, fbsk,
                scratch, p.n_small, p.k, p.N, p.bsk_base_log, p.bsk_level, ns, 1, 1);
  cb_pbs_cleanup(stream, gpu, &scratch);
  // + alpha on the body, rows extended by a zero "body" for the packing keyswitch
  k_gather_rows_add_body<<<grid_for(ext.n), WT, 0, stream>>>(ext.d, pbs_out.d, d_seq.d, d_alpha.d,
                                                           big_size, big_size + 1, ext.n);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
  // private functional packing keyswitch with key r -> row r of GGSW (bit, level).  GGSW slots:
  // tree selectors (bits 0..tree-1) keep their order; the blind-rotation selectors are stored
  // last-bit-first, the order in which blind_rotate_assign applies them
  std::vector<DevArr<u64>> keep;
  for (u32 r = 0; r < gs; ++r) {
    std::vector<u64> out_rows(ns);
    for (u32 b = 0; b < nbits; ++b) {
      const u32 slot = b < tree ? b : tree + (nbits - 1 - b);
      for (u32 o = 0; o < p.cbs_level; ++o)
        out_rows[b * p.cbs_level + o] = ((u64)slot * p.cbs_level + o) * gs + r;
    }
    DevArr<u64> d_out_rows;
    d_out_rows.upload(stream, out_rows);
    keep.push_back(d_out_rows);
    cb_keyswitch_launch(stream, ggsw.d, d_out_rows.d, ext.d, d_seq.d, pfpksk + (size_t)r * pfpksk_words,
                        big + 1, (u32)glwe_words - 1, p.pks_base_log, p.pks_level, ns);
  }
  // ---- Fourier: tree selectors in the generic layout, blind-rotation selectors in the layout of
  // the PBS kernel family of (k, N, l_cbs)
  if (tree)
    cb_ggsw_convert_generic(stream, gpu, ggsw.d, (size_t)tree * p.cbs_level * gs * gs, p.k, p.N);
  cb_bsk_convert(stream, gpu, ggsw.d + (size_t)tree * ggsw_words, ggsw.d + (size_t)tree * ggsw_words,
                 nbr, p.k, p.cbs_level, p.N, true);

  // ---- vertical packing: CMUX tree (all luts, all pairs of a layer per launch)
  DevArr<u64> layer_a, layer_b;
  size_t count = (size_t)1 << tree; // GLWEs per lut in the current layer
  layer_a.alloc(stream, (size_t)lut_count * count * glwe_words);
  layer_b.alloc(stream, std::max<size_t>((size_t)lut_count * (count / 2) * glwe_words, 1));
  k_tree_leaves<<<grid_for(layer_a.n), WT, 0, stream>>>(layer_a.d, luts, p.k, p.N, lut_size,
                                                       polys_per_lut, layer_a.n);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
  u64 *cur = layer_a.d, *nxt = layer_b.d;
  for (u32 j = 0; j < tree; ++j) { // layer j: selector = tree GGSW tree-1-j
    const double2 *g = reinterpret_cast<const double2 *>(ggsw.d + (size_t)(tree - 1 - j) * ggsw_words);
    cb_cmux_launch(stream, gpu, nxt, cur, g, p.k, p.N, p.cbs_base_log, p.cbs_level,
                   (u32)(lut_count * (count / 2)));
    std::swap(cur, nxt);
    count /= 2;
  }
  // ---- blind rotation by the remaining bits + sample extraction: one fused PBS launch
  DevArr<u64> d_mask, d_zero, d_lutsel, d_oseq;
  {
    std::vector<u64> mask(nbr + 1, 0), zero(lut_count, 0), sel(lut_count), oseq(lut_count);
    u32 log2n = 0;
    while ((1u << log2n) < 2 * p.N) ++log2n;
    for (u32 i = 0; i < nbr; ++i) mask[i] = (u64)(2 * p.N - (1u << i)) << (64 - log2n);
    for (u32 l = 0; l < lut_count; ++l) sel[l] = oseq[l] = l;
    d_mask.upload(stream, mask);
    d_zero.upload(stream, zero);
    d_lutsel.upload(stream, sel);
    d_oseq.upload(stream, oseq);
  }
  cb_pbs_scratch(stream, gpu, &scratch, p.k, p.N, p.cbs_level, lut_count, true);
  cb_pbs_launch(stream, gpu, out, d_oseq.d, cur, d_lutsel.d, d_mask.d, d_zero.d,
                reinterpret_cast<const double2 *>(ggsw.d + (size_t)tree * ggsw_words), scratch, nbr,
                p.k, p.N, p.cbs_base_log, p.cbs_level, lut_count, 1, 1);
  cb_pbs_cleanup(stream, gpu, &scratch);
  for (auto &a : keep) a.free(stream);
  for (DevArr<u64> *a : {&pbs_in, &pbs_out, &ext, &lut_cbs, &ggsw, &d_in_idx, &d_lut_idx, &d_seq, &d_q4,
                         &d_bitrows, &d_alpha_neg, &d_alpha, &layer_a, &layer_b, &d_mask, &d_zero,
                         &d_lutsel, &d_oseq})
    a->free(stream);
}

WopShape make_shape(uint32_t lwe_small_dim, uint32_t glwe_dim, uint32_t polynomial_size,
                    uint32_t bsk_level, uint32_t bsk_base_log, uint32_t ksk_level,
                    uint32_t ksk_base_log, uint32_t pksk_level, uint32_t pksk_base_log,
                    uint32_t cbs_level, uint32_t cbs_base_log) {
  WopShape p;
  p.n_small = lwe_small_dim;
  p.k = glwe_dim;
  p.N = polynomial_size;
  p.bsk_level = bsk_level;
  p.bsk_base_log = bsk_base_log;
  p.ks_level = ksk_level;
  p.ks_base_log = ksk_base_log;
  p.pks_level = pksk_level;
  p.pks_base_log = pksk_base_log;
  p.cbs_level = cbs_level;
  p.cbs_base_log = cbs_base_log;
  if (cbs_level == 0 || cbs_base_log == 0 || cbs_level * cbs_base_log > 63)
    CB_PANIC("Cuda error (WoP-PBS): unsupported circuit-bootstrap decomposition base_log=%u level=%u",
             cbs_base_log, cbs_level);
  return p;
}

} // namespace

extern "C" {

void cuda_extract_bits_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *bits_out, const void *lwe_array_in,
    const uint32_t *number_of_bits, uint32_t blocks, const void *fourier_bsk, const void *ksk,
    uint32_t lwe_small_dim, uint32_t glwe_dim, uint32_t polynomial_size, uint32_t bsk_level,
    uint32_t bsk_base_log, uint32_t ksk_level, uint32_t ksk_base_log) {
  CB_CHECK(cudaSetDevice(gpu_index));
  if (blocks == 0) return;
  const WopShape p = make_shape(lwe_small_dim, glwe_dim, polynomial_size, bsk_level, bsk_base_log,
                                ksk_level, ksk_base_log, 1, 1, 1, 1);
  std::vector<u32> nbits(number_of_bits, number_of_bits + blocks);
  for (u32 b : nbits)
    if (b == 0 || b > 58) CB_PANIC("Cuda error (extract bits): %u bits per block unsupported", b);
  extract_bits_blocks(as_stream(stream), gpu_index, (u64 *)bits_out, (const u64 *)lwe_array_in, nbits,
                      (const double2 *)fourier_bsk, (const u64 *)ksk, p, false);
}

void cuda_circuit_bootstrap_vertical_packing_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, const void *bits_in, const void *luts,
    const void *fourier_bsk, const void *pfpksk, uint32_t number_of_bits, uint32_t lut_count,
    uint64_t lut_size, uint32_t lwe_small_dim, uint32_t glwe_dim, uint32_t polynomial_size,
    uint32_t bsk_level, uint32_t bsk_base_log, uint32_t pksk_level, uint32_t pksk_base_log,
    uint32_t cbs_level, uint32_t cbs_base_log) {
  CB_CHECK(cudaSetDevice(gpu_index));
  if (lut_count == 0) return;
  if (lut_size != ((uint64_t)1 << number_of_bits))
    CB_PANIC("Cuda error (vertical packing): lut_size %llu != 2^%u", (unsigned long long)lut_size,
             number_of_bits);
  const WopShape p = make_shape(lwe_small_dim, glwe_dim, polynomial_size, bsk_level, bsk_base_log, 1,
                                1, pksk_level, pksk_base_log, cbs_level, cbs_base_log);
  cbs_vertical_packing(as_stream(stream), gpu_index, (u64 *)lwe_array_out, (const u64 *)bits_in,
                       number_of_bits, (const u64 *)luts, lut_count, lut_size,
                       (const double2 *)fourier_bsk, (const u64 *)pfpksk, p);
}

void cuda_wop_pbs_crt_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, const void *lwe_array_in,
    const void *lut_ct, const uint64_t *crt_decomposition, uint32_t blocks, const void *fourier_bsk,
    const void *ksk, const void *pfpksk, uint32_t lwe_small_dim, uint32_t glwe_dim,
    uint32_t polynomial_size, uint32_t bsk_level, uint32_t bsk_base_log, uint32_t ksk_level,
    uint32_t ksk_base_log, uint32_t pksk_level, uint32_t pksk_base_log, uint32_t cbs_level,
    uint32_t cbs_base_log) {
  CB_CHECK(cudaSetDevice(gpu_index));
  if (blocks == 0) return;
  cudaStream_t s = as_stream(stream);
  const WopShape p = make_shape(lwe_small_dim, glwe_dim, polynomial_size, bsk_level, bsk_base_log,
                                ksk_level, ksk_base_log, pksk_level, pksk_base_log, cbs_level,
                                cbs_base_log);
  std::vector<u32> nbits(blocks);
  u32 total = 0;
  for (u32 i = 0; i < blocks; ++i) { // wrappers.cpp:896-905
    nbits[i] = (u32)std::ceil(std::log2((double)crt_decomposition[i]));
    if (nbits[i] == 0 || nbits[i] > 58)
      CB_PANIC("Cuda error (WoP-PBS): CRT modulus %llu unsupported",
               (unsigned long long)crt_decomposition[i]);
    total += nbits[i];
  }
  if (total > 30) CB_PANIC("Cuda error (WoP-PBS): %u bits in total (lut of 2^%u entries)", total, total);
  u64 *bits = nullptr;
  CB_CHECK(cudaMallocAsync((void **)&bits, (size_t)total * (lwe_small_dim + 1) * sizeof(u64), s));
  extract_bits_blocks(s, gpu_index, bits, (const u64 *)lwe_array_in, nbits,
                      (const double2 *)fourier_bsk, (const u64 *)ksk, p, true);
  cbs_vertical_packing(s, gpu_index, (u64 *)lwe_array_out, bits, total, (const u64 *)lut_ct, blocks,
                       (size_t)1 << total, (const double2 *)fourier_bsk, (const u64 *)pfpksk, p);
  CB_CHECK(cudaFreeAsync(bits, s));
}

} // extern "C"
