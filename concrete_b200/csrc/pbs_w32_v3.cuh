// pbs_w32_v3.cuh — third generation of the warp-per-polynomial PBS kernel (N=2048, k=1, l=1).
//
// ncu on v2 with ONE ciphertext per CTA (profiles/r1_w32_v2_cpb1_*): a lone warp pair needs
// 21 000 cycles per CMUX, four ciphertexts per SM need 26 500: v2 is bound by per-warp latency
// (two warps per SM sub-partition cannot cover the LDS/DFMA/ALU dependency chains), not by a
// pipe (FP64 45 %, shared-memory data pipe 49 %, issue slots 44 %).  v3 changes the occupancy
// plan:
//   * 6 ciphertexts = 12 warps per CTA, THREE warps per sub-partition, <= 168 registers: the
//     thread's own accumulator pairs are no longer carried in registers between CMUXes but
//     re-read from TMEM in 4-pair chunks, and the CMUX body is an out-of-line function so the
//     caller's loop state is not live across the transforms;
//   * shared memory = 12 x 16 KiB accumulator/transpose buffers + a 4 x 8 KiB key ring; the
//     lane-dependent second-pass twiddles live in TMEM (tw[2^s + b + 1] = i * tw[2^s + b] for
//     even b halves them; 80 columns per lane quadrant);
//   * all 12 warps walk the CMUX index in lock-step and share every key fill: GGSW i streams
//     through the ring in eight 8 KiB groups of 4 Fourier rows.  Slot and mbarrier parity of a
//     group are compile-time constants (8 groups per CMUX, 4 slots); consumers wait on full[slot]
//     and arrive on empty[slot]; one elected lane of warp 0 refills a slot three groups ahead
//     (TMA bulk copies, complete_tx on full[slot]);
//   * the last two forward stages, the product and the cross-term store are fused per group of
//     4 rows (and the cross-term add with the first two inverse stages), so the
//     shared-memory-bound product phase carries FP64 work instead of following it.
// Measured alternative (kept out): de-phasing groups of warps so that FP64-bound and
// shared-memory-bound phases of different ciphertexts overlap.  The unrolled CMUX body is
// ~100 KB of SASS; three different instruction streams per sub-partition turn the kernel
// instruction-fetch bound (stall_no_instruction 60-80 %, 1.6 IPC per SM, 20 % SLOWER than
// lock-step where the 3 warps of a sub-partition share the fetch).
// Arithmetic is unchanged from v1/v2 (same transforms, same rounding, same digit extraction).
#pragma once

#ifndef W32_V3_L2_KEEP
#define W32_V3_L2_KEEP 0 // 1 = key fills carry an L2 evict_last hint (sm100_prims.cuh tma_load_bulk_keep).
                         // Measured (profiles/r2_l2_keep_ab.txt): DRAM reads 826.7 MB with, 824.0 MB
                         // without, same time - the excess reads are not key evictions; off
#endif
#ifndef W32_V3_LF
#define W32_V3_LF 1 // inverse pass 1 with 6-op butterflies (pbs_w32.cuh inv_pass1_lf); 0 = 8-op
#endif

namespace w32 {

constexpr int CPB3 = 6;                 // ciphertexts per CTA
constexpr int NW3 = 2 * CPB3;           // warps per CTA
constexpr int NTHREADS3 = 32 * NW3;     // 384
constexpr u32 V3_BUF_BYTES = (u32)M * 16;                       // 16 KiB per warp
constexpr u32 V3_OFF_RING = (u32)NW3 * V3_BUF_BYTES;            // 196 608
#ifndef W32_V3_GPF
#define W32_V3_GPF 2 // compute groups (4 Fourier rows) per key fill: 2 = four 16 KiB fills per CMUX into two slots
                     // (half the barrier waits and counted hand-backs), 1 = eight 8 KiB fills into four slots
#endif
constexpr int V3_GPF = W32_V3_GPF;
constexpr int V3_GROUPS = 8;            // compute groups of 4 rows per GGSW
constexpr int V3_FILLS = V3_GROUPS / V3_GPF;        // key fills per CMUX
constexpr int V3_SLOTS = 4 / V3_GPF;                // ring slots (the ring stays 32 KiB)
constexpr u32 V3_CH = 2048u * V3_GPF;   // bytes of one GGSW polynomial in a fill (4 * GPF rows x 32 lanes x 16 B)
constexpr u32 V3_SLOT_BYTES = 4 * V3_CH; // the 4 GGSW polynomials
static_assert(V3_GPF == 1 || V3_GPF == 2, "fills of 4 or 8 rows");
// group G reads fill F = G / GPF of its CMUX (sub-block G % GPF of it): slot and mbarrier parity are
// compile-time constants because every slot takes an even number of fills per CMUX
template <int G> struct V3Fill {
  static constexpr int F = G / V3_GPF, SUB = G % V3_GPF;
  static constexpr int SLOT = F & (V3_SLOTS - 1), PARITY = (F / V3_SLOTS) & 1;
  static constexpr bool FIRST = SUB == 0, LAST = SUB == V3_GPF - 1;
};
static_assert((V3_FILLS / V3_SLOTS) % 2 == 0, "static parity");
constexpr u32 V3_OFF_CTRL = V3_OFF_RING + V3_SLOTS * V3_SLOT_BYTES;
constexpr size_t SMEM_BYTES3 = V3_OFF_CTRL + 256;
constexpr u32 V3_TMEM_COLS = 512;
constexpr u32 V3_TW_COL = 384;          // twiddle columns of a lane quadrant (after 3 x 128 own)

struct V3Ctrl {
  unsigned long long full_bar[V3_SLOTS];  // fill landed (1 arrival + tx bytes)
  unsigned int done_cnt[V3_SLOTS];        // warps done with the slot's current fill
  unsigned int tmem_base;
  unsigned int n, base_log, total_fills;
  const double2 *bsk;
  unsigned int wt0[NW3], wi[NW3];         // per warp: fill index base and GGSW index of its CMUX
};

__device__ __forceinline__ u32 v3_ctrl_addr() {
  extern __shared__ __align__(1024) unsigned char smem[];
  return smem_u32(smem) + V3_OFF_CTRL;
}
__device__ __forceinline__ void mbar_wait_s(u32 bar_addr, u32 parity) {
  u32 done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar_addr), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void mbar_arrive_s(u32 bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ double2 lds128(u32 addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(u32 addr, const double2 v) {
  asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(v.x), "d"(v.y) : "memory");
}

// Fill t = cmux * FILLS + f: rows 4*GPF*f .. of the four polynomials of GGSW i -> slot f mod SLOTS.
// One thread.  The caller has waited for empty[slot] of the previous fill.
__device__ __forceinline__ void v3_issue_fill(u32 i, u32 g) {
  extern __shared__ __align__(1024) unsigned char smem[];
  V3Ctrl *ctrl = reinterpret_cast<V3Ctrl *>(smem + V3_OFF_CTRL);
  const u32 slot = g & (V3_SLOTS - 1);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic reads before async write
  mbar_expect_tx(&ctrl->full_bar[slot], V3_SLOT_BYTES);
  const double2 *src = ctrl->bsk + (size_t)i * 4 * M + (size_t)g * (4 * V3_GPF) * 32;
  unsigned char *dst = smem + V3_OFF_RING + slot * V3_SLOT_BYTES;
  // slot layout [r][own, cross]: polynomial (row r, column c) of the GGSW goes to position
  // 2r + (c != r), so a warp addresses both of its polynomials from one base register
#if W32_V3_L2_KEEP
  const unsigned long long pol = l2_policy_evict_last();
#endif
#pragma unroll
  for (int poly = 0; poly < 4; ++poly) {
    const int r = poly >> 1, c = poly & 1, pos = 2 * r + (c != r);
#if W32_V3_L2_KEEP
    tma_load_bulk_keep(dst + pos * V3_CH, src + (size_t)poly * M, V3_CH, &ctrl->full_bar[slot], pol);
#else
    tma_load_bulk(dst + pos * V3_CH, src + (size_t)poly * M, V3_CH, &ctrl->full_bar[slot]);
#endif
  }
}

// consumer side of group G: the fill of rows 4G..4G+3 (slot and parity are static).
// try_once issues one non-looping try_wait whose ~90-cycle latency can be covered by other work.
template <int G> __device__ __forceinline__ u32 v3_try_full() {
  u32 done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(v3_ctrl_addr() + 8 * V3Fill<G>::SLOT), "n"(V3Fill<G>::PARITY)
      : "memory");
  return done;
}
template <int G> __device__ __forceinline__ void v3_wait_full(u32 tried) {
  if (V3Fill<G>::FIRST && !tried) mbar_wait_s(v3_ctrl_addr() + 8 * V3Fill<G>::SLOT, V3Fill<G>::PARITY);
}
// This warp has issued its last read of fill t0 + G: count it (shared-memory operations of a
// warp execute in order, so the count follows the reads).  Returns the previous count to lane 0.
template <int G> __device__ __forceinline__ u32 v3_count_done(int lane) {
  u32 old = 0;
  if (!V3Fill<G>::LAST) return old; // more reads of this fill to come (the next group's)
  __syncwarp();
  if (lane == 0)
    asm volatile("atom.shared.inc.u32 %0, [%1], %2;" // wraps to 0 after the 12th warp
                 : "=r"(old)
                 : "r"(v3_ctrl_addr() + 8 * V3_SLOTS + 4 * V3Fill<G>::SLOT), "n"(NW3 - 1)
                 : "memory");
  return old;
}
// The warp that counted last (all 12 done) refills the slot with the fill SLOTS ahead.
template <int G> __device__ __forceinline__ void v3_refill(u32 old, int lane) {
  extern __shared__ __align__(1024) unsigned char smem[];
  constexpr int F = V3Fill<G>::F;
  if (V3Fill<G>::LAST && lane == 0 && old == NW3 - 1) {
    V3Ctrl *ctrl = reinterpret_cast<V3Ctrl *>(smem + V3_OFF_CTRL);
    const u32 t0 = ctrl->wt0[threadIdx.x >> 5], i = ctrl->wi[threadIdx.x >> 5];
    if (t0 + F + V3_SLOTS < ctrl->total_fills) {
      const u32 n = ctrl->n;
      const u32 inext = (i + 1 == n) ? 0 : i + 1;
      v3_issue_fill((F + V3_SLOTS >= V3_FILLS) ? inext : i, (F + V3_SLOTS) & (V3_FILLS - 1));
    }
  }
}

// a CMUX that adds nothing (a_hat == 0) or an idle ciphertext slot: only walk the ring
template <int G> __device__ __forceinline__ void v3_idle_groups(int lane) {
  v3_wait_full<G>(0);
  v3_refill<G>(v3_count_done<G>(lane), lane);
  if constexpr (G + 1 < V3_GROUPS) v3_idle_groups<G + 1>(lane);
}

__device__ __forceinline__ void tmem_ld8(u32 taddr, u32 (&v)[8]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ void tmem_ld4(u32 taddr, u32 (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr)
               : "memory");
}
// one twiddle (4 columns): an x4 load needs only a 4-aligned register quad, like LDS.128; the
// x8/x16 forms need 8/16 consecutive registers, which ptxas can only find by spilling points
__device__ __forceinline__ double2 tw_of(const u32 (&v)[4]) {
  return make_double2(__hiloint2double((int)v[1], (int)v[0]), __hiloint2double((int)v[3], (int)v[2]));
}

__device__ __forceinline__ double2 mul_i(const double2 w) { return make_double2(-w.y, w.x); }
template <int NV> __device__ __forceinline__ double2 tw_from(const u32 (&v)[NV], int e) {
  return make_double2(__hiloint2double((int)v[4 * e + 1], (int)v[4 * e + 0]),
                      __hiloint2double((int)v[4 * e + 3], (int)v[4 * e + 2]));
}
__device__ __forceinline__ double2 cmul(const double2 a, const double2 b) {
  return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}

// Second-pass twiddles of this lane in TMEM (tw(s2,u) = tw[2^(s2+5) + (lane << s2) + u];
// odd u = i * (u - 1)):
//   columns  0..15 : (0,0) (1,0) (2,0) (2,2)
//   columns 16+8g..: (4,2g) (3, g & ~1)            g = 0..7 (the group of rows 4g..4g+3)
constexpr u32 V3_TW_NCOLS = 16 + 8 * V3_GROUPS;

// stages s2 = 0,1,2 (forward) on all 32 points
__device__ __forceinline__ void v3_fwd_stages012(double2 (&x)[32], u32 tmem_tw) {
  u32 va[4], vb[4];
  tmem_ld4(tmem_tw, va);
  tmem_ld4(tmem_tw + 4, vb);
  tmem_wait_ld();
  const double2 t0 = tw_of(va), t1 = tw_of(vb);
#pragma unroll
  for (int lo = 0; lo < 16; ++lo) cbfft::bfly_ct(x[lo], x[lo | 16], t0);
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int lo = 0; lo < 8; ++lo)
      cbfft::bfly_ct(x[(u << 4) | lo], x[(u << 4) | lo | 8], u ? mul_i(t1) : t1);
  tmem_ld4(tmem_tw + 8, va);
  tmem_ld4(tmem_tw + 12, vb);
  tmem_wait_ld();
  const double2 t2a = tw_of(va), t2b = tw_of(vb);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const double2 b = (u & 2) ? t2b : t2a;
#pragma unroll
    for (int lo = 0; lo < 4; ++lo)
      cbfft::bfly_ct(x[(u << 3) | lo], x[(u << 3) | lo | 4], (u & 1) ? mul_i(b) : b);
  }
}
__device__ __forceinline__ void v3_inv_stages210(double2 (&x)[32], u32 tmem_tw) {
  u32 va[4], vb[4];
  tmem_ld4(tmem_tw + 8, va);
  tmem_ld4(tmem_tw + 12, vb);
  tmem_wait_ld();
  const double2 t2a = tw_of(va), t2b = tw_of(vb);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const double2 b = (u & 2) ? t2b : t2a;
#pragma unroll
    for (int lo = 0; lo < 4; ++lo)
      cbfft::bfly_gs(x[(u << 3) | lo], x[(u << 3) | lo | 4], (u & 1) ? mul_i(b) : b);
  }
  tmem_ld4(tmem_tw, va);
  tmem_ld4(tmem_tw + 4, vb);
  tmem_wait_ld();
  const double2 t0 = tw_of(va), t1 = tw_of(vb);
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int lo = 0; lo < 8; ++lo)
      cbfft::bfly_gs(x[(u << 4) | lo], x[(u << 4) | lo | 8], u ? mul_i(t1) : t1);
#pragma unroll
  for (int lo = 0; lo < 16; ++lo) cbfft::bfly_gs(x[lo], x[lo | 16], t0);
}

__device__ __forceinline__ void pack_pairs(u32 (&v)[16], const u64 (&re)[4], const u64 (&im)[4]) {
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    v[4 * e + 0] = (u32)re[e];
    v[4 * e + 1] = (u32)(re[e] >> 32);
    v[4 * e + 2] = (u32)im[e];
    v[4 * e + 3] = (u32)(im[e] >> 32);
  }
}

// closest_representable + balanced digit of a one-level decomposition from the high word:
// floor((hi + 2^(k-1)) / 2^k) taken as a signed b-bit value, except that the reference maps
// B/2 to +B/2 (decomposer.rs:23-47, iter.rs:121-133 with state >> base_log == 0).  k = 32 - b.
// Two instructions: round-half-up of x equals minus round-half-down of -x, and an arithmetic
// shift of (2^(k-1) - 1 - hi) is exactly that round-half-down; the result lies in [-B/2, B/2),
// so its negation lies in (-B/2, B/2] with +B/2 at the tie, as the reference has it.  The caller
// takes the NEGATED digit: the sign is folded into the first butterflies' FMA operands.
__device__ __forceinline__ int neg_digit_l1_k(u32 hi, u32 kshift) {
  return (int)(((1u << (kshift - 1)) - 1u) - hi) >> kshift;
}
__device__ __forceinline__ int digit_l1_k(u32 hi, u32 kshift) { return -neg_digit_l1_k(hi, kshift); }
__device__ __forceinline__ u64 lds_u64(u32 addr) {
  u64 v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ ulonglong2 lds_u64x2(u32 addr) {
  ulonglong2 v;
  asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr));
  return v;
}
// digits of rows 4c..4c+3 (own pairs in v) into x
template <int DUMMY>
__device__ __forceinline__ void v3_digits(double2 (&x)[32], const u32 (&v)[16], int c, u32 s16,
                                          u32 buf_s, u32 kshift) {
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int p = 4 * c + e;
    const u64 own_re = ((u64)v[4 * e + 1] << 32) | v[4 * e + 0];
    const u64 own_im = ((u64)v[4 * e + 3] << 32) | v[4 * e + 2];
    const u32 sp = s16 + 512u * p;
    // (re, im) swapped when bit 14 of sp is set: fetch the two halves of the pair from
    // address ^ 8 instead of selecting after a 128-bit load (4 SEL per point)
    const u32 a0 = buf_s + (sp & 0x3FF0u) + ((sp >> 11) & 8u);
    const u64 ra = lds_u64(a0), rb = lds_u64(a0 ^ 8u);
    // im is negated when bits 15 and 14 differ, i.e. when bit 15 of sp + 2^14 is set
    const bool nre = (sp & 0x8000u) != 0, nim = ((sp + 0x4000u) & 0x8000u) != 0;
    const u64 c_re = (nre ? (u64)0 - ra : ra) - own_re;
    const u64 c_im = (nim ? (u64)0 - rb : rb) - own_im;
    x[p].x = -__int2double_rn(neg_digit_l1_k((u32)(c_re >> 32), kshift));
    x[p].y = -__int2double_rn(neg_digit_l1_k((u32)(c_im >> 32), kshift));
  }
}

// Shared-memory addresses of this thread inside its warp's 16 KiB buffer (the buffer base is
// 1 KiB aligned).  Each address is ONE just-in-time XOR with an immediate plus an immediate
// offset in the load/store; the XOR is an opaque asm so that ptxas does not keep the eight
// swizzle variants of both layouts live across the transforms (16 registers it does not have):
//   row layout   (index 32*lane + p, key lane): (rowx ^ 16(p & 7)) + 128(p >> 3)
//   column layout(index lane + 32p,  key p)   : (colb ^ 16(p & 7)) + 512p
// with rowx = buf + 512*lane + 16*(lane & 7), colb = buf + 16*lane.
struct V3Addr {
  u32 rowx, colb;
};
template <u32 OFF> __device__ __forceinline__ double2 lds128o(u32 addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(v.x), "=d"(v.y) : "r"(addr), "n"(OFF));
  return v;
}
// load / store at ((base ^ XORI) + OFF): the XOR lives inside the asm so that no address
// register outlives the access
template <u32 XORI, u32 OFF> __device__ __forceinline__ double2 lds128x(u32 base) {
  double2 v;
  asm volatile(
      "{\n"
      ".reg .b32 t;\n"
      "xor.b32 t, %2, %3;\n"
      "ld.shared.v2.f64 {%0, %1}, [t+%4];\n"
      "}\n"
      : "=d"(v.x), "=d"(v.y)
      : "r"(base), "n"(XORI), "n"(OFF));
  return v;
}
template <u32 XORI, u32 OFF> __device__ __forceinline__ void sts128x(u32 base, const double2 v) {
  asm volatile(
      "{\n"
      ".reg .b32 t;\n"
      "xor.b32 t, %0, %1;\n"
      "st.shared.v2.f64 [t+%2], {%3, %4};\n"
      "}\n" ::"r"(base),
      "n"(XORI), "n"(OFF), "d"(v.x), "d"(v.y)
      : "memory");
}
template <int P> __device__ __forceinline__ double2 ld_row(const V3Addr &A, u32 delta = 0) {
  return lds128x<16u * (P & 7), 128u * (P >> 3)>(A.rowx + delta);
}
template <int P> __device__ __forceinline__ void st_row(const V3Addr &A, const double2 v) {
  sts128x<16u * (P & 7), 128u * (P >> 3)>(A.rowx, v);
}
template <int P> __device__ __forceinline__ double2 ld_col(const V3Addr &A) {
  return lds128x<16u * (P & 7), 512u * P>(A.colb);
}
template <int P> __device__ __forceinline__ void st_col(const V3Addr &A, const double2 v) {
  sts128x<16u * (P & 7), 512u * P>(A.colb, v);
}
template <int P> __device__ __forceinline__ void st_col_all(const V3Addr &A, const double2 (&x)[32]) {
  st_col<P>(A, x[P]);
  if constexpr (P + 1 < 32) st_col_all<P + 1>(A, x);
}
template <int P> __device__ __forceinline__ void ld_col_all(const V3Addr &A, double2 (&x)[32]) {
  x[P] = ld_col<P>(A);
  if constexpr (P + 1 < 32) ld_col_all<P + 1>(A, x);
}
template <int P> __device__ __forceinline__ void st_row_all(const V3Addr &A, const double2 (&x)[32]) {
  st_row<P>(A, x[P]);
  if constexpr (P + 1 < 32) st_row_all<P + 1>(A, x);
}
template <int P> __device__ __forceinline__ void ld_row_all(const V3Addr &A, double2 (&x)[32]) {
  x[P] = ld_row<P>(A);
  if constexpr (P + 1 < 32) ld_row_all<P + 1>(A, x);
}

// product of row P with the staged key: cross term to the partner through buf, own term in place
template <int P, u32 SO>
__device__ __forceinline__ void v3_product(double2 (&x)[32], const V3Addr &A, u32 keyr) {
  const double2 w = lds128o<SO + 512 * (P & 3)>(keyr), wo = lds128o<SO + V3_CH + 512 * (P & 3)>(keyr);
  st_row<P>(A, cmul(x[P], wo));
  x[P] = cmul(x[P], w);
}

// forward group G: stages s2 = 3,4 on rows 4G..4G+3, product with the Fourier key (own term in
// place, cross term to the partner warp through this warp's buffer).  Software pipelined across
// groups: on entry tv holds the twiddles of G (tcgen05.ld in flight) and `tried` the result of
// an early try_wait on the fill of G; the body issues both for G + 1 under its own arithmetic.
// wait for the TMEM loads in flight; ta/tb are operands so that every later use reads them after
// the wait (the loads that fill them were issued one group earlier)
__device__ __forceinline__ void tmem_wait_ld8(u32 (&ta)[4], u32 (&tb)[4]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(ta[0]), "+r"(ta[1]), "+r"(ta[2]), "+r"(ta[3]), "+r"(tb[0]), "+r"(tb[1]),
                 "+r"(tb[2]), "+r"(tb[3])
               :
               : "memory");
}
template <int G>
__device__ __forceinline__ void v3_fwd_group(double2 (&x)[32], const V3Addr &A, u32 tmem_tw,
                                             u32 keyr, int lane, u32 (&ta)[4], u32 (&tb)[4]) {
  {
    tmem_wait_ld8(ta, tb); // twiddles of group G, loaded under the products of group G - 1
    const double2 t4 = tw_of(ta);
    const double2 t3 = (G & 1) ? mul_i(tw_of(tb)) : tw_of(tb);
    cbfft::bfly_ct(x[4 * G + 0], x[4 * G + 2], t3);
    cbfft::bfly_ct(x[4 * G + 1], x[4 * G + 3], t3);
    cbfft::bfly_ct(x[4 * G + 0], x[4 * G + 1], t4);
    cbfft::bfly_ct(x[4 * G + 2], x[4 * G + 3], mul_i(t4));
    if constexpr (G + 1 < V3_GROUPS) { // t3/t4 are dead: refill the same registers for G + 1
      tmem_ld4(tmem_tw + 16 + 8 * (G + 1), ta);
      tmem_ld4(tmem_tw + 20 + 8 * (G + 1), tb);
    }
  }
  v3_wait_full<G>(0);
  constexpr u32 so = V3Fill<G>::SLOT * V3_SLOT_BYTES + V3Fill<G>::SUB * 2048u;
  u32 old = 0;
  v3_product<4 * G + 0, so>(x, A, keyr);
  v3_product<4 * G + 1, so>(x, A, keyr);
  v3_product<4 * G + 2, so>(x, A, keyr);
  {
    constexpr int p = 4 * G + 3;
    const double2 w = lds128o<so + 512 * 3>(keyr), wo = lds128o<so + V3_CH + 512 * 3>(keyr);
    old = v3_count_done<G>(lane); // every read of the slot has been issued
    st_row<p>(A, cmul(x[p], wo));
    x[p] = cmul(x[p], w);
  }
  v3_refill<G>(old, lane);
  if constexpr (G + 1 < V3_GROUPS) v3_fwd_group<G + 1>(x, A, tmem_tw, keyr, lane, ta, tb);
}

// inverse group G: add the partner's cross term, inverse stages s2 = 4,3
template <int G>
__device__ __forceinline__ void v3_inv_group(double2 (&x)[32], const V3Addr &A, u32 tmem_tw,
                                             u32 oth_delta) {
  u32 ta[4], tb[4];
  tmem_ld4(tmem_tw + 16 + 8 * G, ta);
  tmem_ld4(tmem_tw + 20 + 8 * G, tb);
  {
    const double2 o0 = ld_row<4 * G + 0>(A, oth_delta), o1 = ld_row<4 * G + 1>(A, oth_delta);
    x[4 * G + 0].x += o0.x;
    x[4 * G + 0].y += o0.y;
    x[4 * G + 1].x += o1.x;
    x[4 * G + 1].y += o1.y;
    const double2 o2 = ld_row<4 * G + 2>(A, oth_delta), o3 = ld_row<4 * G + 3>(A, oth_delta);
    x[4 * G + 2].x += o2.x;
    x[4 * G + 2].y += o2.y;
    x[4 * G + 3].x += o3.x;
    x[4 * G + 3].y += o3.y;
  }
  tmem_wait_ld();
  const double2 t4 = tw_of(ta);
  const double2 t3 = (G & 1) ? mul_i(tw_of(tb)) : tw_of(tb);
  cbfft::bfly_gs(x[4 * G + 0], x[4 * G + 1], t4);
  cbfft::bfly_gs(x[4 * G + 2], x[4 * G + 3], mul_i(t4));
  cbfft::bfly_gs(x[4 * G + 0], x[4 * G + 2], t3);
  cbfft::bfly_gs(x[4 * G + 1], x[4 * G + 3], t3);
  if constexpr (G + 1 < V3_GROUPS) v3_inv_group<G + 1>(x, A, tmem_tw, oth_delta);
}

// Per-phase thread state, rebuilt from a fresh (opaque) read of %tid.x: the values are cheap,
// and separate live ranges keep ptxas from spilling one long-lived address register and
// reloading it from local memory (L2 latency) in front of every shared-memory access.
struct V3Phase {
  int lane, warp;
  u32 buf_s;
  V3Addr A;
  u32 tmem_own, tmem_tw;
};
__device__ __forceinline__ V3Phase v3_phase() {
  extern __shared__ __align__(1024) unsigned char smem[];
  u32 tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  V3Phase P;
  P.lane = tid & 31;
  P.warp = tid >> 5;
  P.buf_s = smem_u32(smem) + (u32)P.warp * V3_BUF_BYTES;
  P.A.rowx = P.buf_s + 512u * P.lane + 16u * (P.lane & 7);
  P.A.colb = P.buf_s + 16u * P.lane;
  u32 tb;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tb) : "r"(v3_ctrl_addr() + 8 * V3_SLOTS + 4 * V3_SLOTS));
  const u32 tmem_quad = tb + ((u32)((P.warp & 3) * 32) << 16);
  P.tmem_own = tmem_quad + (u32)(P.warp >> 2) * 128; // a third of the quadrant's columns
  P.tmem_tw = tmem_quad + V3_TW_COL;
  return P;
}

// One CMUX of one polynomial (one warp).  Kept out of line on purpose: with 128 of the 168
// registers holding the 32 complex points, the loop state of the caller (round, sample, LWE
// pointer, counters) must not stay live across the transforms; everything the body needs beyond
// its three arguments is re-derived from the thread index or read from the control block.
__device__ __noinline__ void v3_cmux(const u32 a_hat) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 x[32];
  { // ct1 = ACC*X^a - ACC, closest_representable, signed digit (l = 1) -> doubles
    const V3Phase P = v3_phase();
    V3Ctrl *ctrl = reinterpret_cast<V3Ctrl *>(smem + V3_OFF_CTRL);
    const u32 kshift = 32 - ctrl->base_log;
    // 16 * ((lane - a_hat) mod 4096): bits 4..13 address the pair, bit 14 swaps (re, im), bit 15
    // negates re, bit 14 ^ bit 15 negates im  (multiplication by (-i)^q, q = bits 15:14)
    const u32 s16 = ((u32)(P.lane - (int)a_hat) << 4) & 0xFFF0u;
    u32 va[16], vb[16];
    tmem_ld16(P.tmem_own, va);
    tmem_wait_ld();
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      tmem_ld16(P.tmem_own + 16 * (2 * cc + 1), vb);
      v3_digits<0>(x, va, 2 * cc, s16, P.buf_s, kshift);
      tmem_wait_ld();
      if (cc < 3) tmem_ld16(P.tmem_own + 16 * (2 * cc + 2), va);
      v3_digits<0>(x, vb, 2 * cc + 1, s16, P.buf_s, kshift);
      if (cc < 3) tmem_wait_ld();
    }
  }

  fwd_pass1(x);
  __syncwarp(); // rotated reads of the accumulator are done in every lane
  {
    const V3Phase P = v3_phase();
    st_col_all<0>(P.A, x);
    __syncwarp();
    ld_row_all<0>(P.A, x);
    v3_fwd_stages012(x, P.tmem_tw);
  }

  __syncwarp(); // transpose reads of the buffer are done in every lane
  {
    const V3Phase P = v3_phase();
    const u32 key = smem_u32(smem) + V3_OFF_RING + 16u * P.lane + (P.warp & 1) * (2 * V3_CH);
    u32 ta[4], tb[4];
    tmem_ld4(P.tmem_tw + 16, ta);
    tmem_ld4(P.tmem_tw + 20, tb);
    v3_fwd_group<0>(x, P.A, P.tmem_tw, key, P.lane, ta, tb);
  }
  {
    const V3Phase P = v3_phase();
    const int r = P.warp & 1;
    asm volatile("bar.sync %0, 64;" ::"r"(1 + (P.warp >> 1)) : "memory");
    v3_inv_group<0>(x, P.A, P.tmem_tw, r ? 0u - V3_BUF_BYTES : V3_BUF_BYTES);
    asm volatile("bar.sync %0, 64;" ::"r"(1 + (P.warp >> 1)) : "memory"); // partner done with buf
  }
  {
    const V3Phase P = v3_phase();
    v3_inv_stages210(x, P.tmem_tw);
    st_row_all<0>(P.A, x);
    __syncwarp();
    ld_col_all<0>(P.A, x);
  }
#if W32_V3_LF
  inv_pass1_lf(x);
#else
  inv_pass1(x);
#endif

  const V3Phase P = v3_phase();
  const int lane = P.lane, warp = P.warp;
  const u32 tmem_own = P.tmem_own;
  // ACC += round(x): own copy from TMEM; the new value goes back to TMEM and to the shared
  // home for the rotated read of the next CMUX
  __syncwarp(); // transpose reads of the buffer are done before it becomes the accumulator again
  ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(smem + (size_t)warp * V3_BUF_BYTES);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    u32 v[16];
    tmem_ld16(tmem_own + 16 * c, v);
    tmem_wait_ld();
    u64 re[4], im[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int p = 4 * c + e;
#if W32_V3_LF
      re[e] = (((u64)v[4 * e + 1] << 32) | v[4 * e + 0]) + from_torus_scaled_lf(x[p].x, c_lf_sigma[p]);
      im[e] = (((u64)v[4 * e + 3] << 32) | v[4 * e + 2]) + from_torus_scaled_lf(x[p].y, c_lf_sigma[p]);
#else
      re[e] = (((u64)v[4 * e + 1] << 32) | v[4 * e + 0]) + from_torus_scaled(x[p].x);
      im[e] = (((u64)v[4 * e + 3] << 32) | v[4 * e + 2]) + from_torus_scaled(x[p].y);
#endif
      acc[lane + 32 * p] = make_ulonglong2(re[e], im[e]);
    }
    pack_pairs(v, re, im);
    tmem_st16(tmem_own + 16 * c, v);
  }
  tmem_wait_st();
  __syncwarp();
}

__global__ void __launch_bounds__(NTHREADS3, 1) pbs_kernel_w32v3(const PbsArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ct = warp >> 1, r = warp & 1;
  ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(smem + (size_t)warp * V3_BUF_BYTES);
  V3Ctrl *ctrl = reinterpret_cast<V3Ctrl *>(smem + V3_OFF_CTRL);
  const u32 n = a.n;
  const size_t out_size = (size_t)N + 1;

  // Sample -> (CTA, round, slot): `full` rounds in which every CTA holds CPB3 ciphertexts, then
  // the remainder spread evenly over ALL CTAs in one last round (a round with fewer active
  // ciphertexts is shorter: idle warps only walk the key ring), so small batches use every SM
  // and the tail of a large batch costs a fraction of a round.
  const u32 wave = gridDim.x * CPB3;
  const u32 full = a.num_samples / wave;
  const u32 rem = a.num_samples - full * wave;
  const u32 rem_base = rem / gridDim.x, rem_extra = rem % gridDim.x;
  const u32 my_rem = rem_base + (blockIdx.x < rem_extra ? 1u : 0u);
  const u32 rem_first = full * wave + blockIdx.x * rem_base + min(blockIdx.x, rem_extra);
  const u32 rounds = full + (my_rem ? 1u : 0u);
  const u32 total_fills = rounds * n * V3_FILLS;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < V3_SLOTS; ++s) {
      mbar_init(&ctrl->full_bar[s], 1);
      ctrl->done_cnt[s] = 0;
    }
    ctrl->n = n;
    ctrl->base_log = a.base_log;
    ctrl->total_fills = total_fills;
    ctrl->bsk = a.bsk;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&ctrl->tmem_base)),
                 "r"(V3_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const u32 tmem_quad = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16);
  const u32 tmem_own = tmem_quad + (u32)(warp >> 2) * 128;
  const u32 tmem_tw = tmem_quad + V3_TW_COL;

  if (warp < 4) { // one warp per lane quadrant stores the quadrant's twiddle columns
#pragma unroll
    for (int c = 0; c < (int)V3_TW_NCOLS / 16; ++c) {
      u32 v[16];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int idx = 4 * c + e; // twiddle slot (4 columns each)
        int s2, u;
        if (idx < 2) { s2 = idx; u = 0; }
        else if (idx < 4) { s2 = 2; u = 2 * (idx - 2); }
        else {
          const int g = (idx - 4) >> 1;
          if ((idx & 1) == 0) { s2 = 4; u = 2 * g; }
          else { s2 = 3; u = g & ~1; }
        }
        const double2 w = a.tw[(1 << (s2 + 5)) + (lane << s2) + u];
        v[4 * e + 0] = (u32)__double2loint(w.x);
        v[4 * e + 1] = (u32)__double2hiint(w.x);
        v[4 * e + 2] = (u32)__double2loint(w.y);
        v[4 * e + 3] = (u32)__double2hiint(w.y);
      }
      tmem_st16(tmem_tw + 16 * c, v);
    }
    tmem_wait_st();
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  if (tid == 0) { // prologue: groups 0..3 of the first GGSW fill the four slots
    for (u32 g = 0; g < (u32)V3_SLOTS && g < total_fills; ++g) v3_issue_fill(0, g);
  }

  u32 cmux = 0; // CMUX rounds of this CTA so far (every warp sees every one of them)
  for (u32 rd = 0; rd < rounds; ++rd) {
    const bool active = rd < full || (u32)ct < my_rem;
    const u32 sample = rd < full ? rd * wave + blockIdx.x * CPB3 + ct : rem_first + ct;
    const u64 *lwe = a.lwe_in + (active ? a.in_idx[sample] : 0) * (u64)(n + 1);

    if (active) {
      const u32 b_hat = modulus_switch(__ldg(&lwe[n]), LOGM + 2);
      const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(2 * N) + (size_t)r * N;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        u64 re[4], im[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = lane + 32 * (4 * c + e);
          {
            const u32 u = (j + b_hat) & (2 * N - 1);
            const u64 w = __ldg(&lut[u & (N - 1)]);
            re[e] = (u < (u32)N) ? w : (u64)0 - w;
          }
          {
            const u32 u = (j + M + b_hat) & (2 * N - 1);
            const u64 w = __ldg(&lut[u & (N - 1)]);
            im[e] = (u < (u32)N) ? w : (u64)0 - w;
          }
          acc[j] = make_ulonglong2(re[e], im[e]);
        }
        u32 v[16];
        pack_pairs(v, re, im);
        tmem_st16(tmem_own + 16 * c, v);
      }
      tmem_wait_st();
    }
    __syncwarp();

    // mod-switched mask elements, 32 at a time: lane k holds a_hat[i0 + k] (one coalesced load
    // per 32 CMUXes instead of a global-load latency in front of every out-of-line call)
    u32 a_blk = 0;
#pragma unroll 1
    for (u32 i = 0; i < n; ++i, ++cmux) {
      if ((i & 31) == 0) {
        const u32 k = i + lane;
        a_blk = (active && k < n) ? modulus_switch(__ldg(&lwe[k]), LOGM + 2) : 0;
      }
      const u32 a_hat = __shfl_sync(0xffffffffu, a_blk, i & 31);
      if (lane == 0) { // what the refill path needs to know about this warp's CMUX
        ctrl->wt0[warp] = cmux * V3_FILLS;
        ctrl->wi[warp] = i;
      }
      if (a_hat == 0) // X^0 - 1 = 0: no CMUX (bootstrap.rs:287); still walk the ring
        v3_idle_groups<0>(lane);
      else
        v3_cmux(a_hat);
    }

    // sample extraction: warp 0 of the pair holds the mask polynomial, warp 1 the body
    if (active) {
      for (u32 li = 0; li < a.lut_count; ++li) {
        const u32 nth = li * a.lut_stride;
        u64 *out = a.lwe_out + ((size_t)li * a.num_samples + a.out_idx[sample]) * out_size;
        if (r == 0) {
          for (int xo = lane; xo < N; xo += 32) {
            const int src = (xo <= (int)nth) ? ((int)nth - xo) : (N + (int)nth - xo);
            const ulonglong2 pr = acc[src & (M - 1)];
            const u64 v = (src < M) ? pr.x : pr.y;
            out[xo] = (xo <= (int)nth) ? v : (u64)0 - v;
          }
        } else if (lane == 0) {
          const ulonglong2 pr = acc[nth & (M - 1)];
          out[N] = (nth < (u32)M) ? pr.x : pr.y;
        }
      }
    }
    __syncwarp();
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(ctrl->tmem_base),
                 "r"(V3_TMEM_COLS)
                 : "memory");
}

} // namespace w32
