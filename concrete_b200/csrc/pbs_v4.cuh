// pbs_v4.cuh — fused PBS for every (k, N, l) with N >= 1024: the v3 data flow (pbs_w32_v3.cuh)
// generalised.  Included by pbs.cu inside its anonymous namespace.
//
// What ncu said about the kernels this replaces (profiles/r2_base_{P4,P6,P7,P8}_*.txt): all four
// are latency-bound (no eligible warp 60-80 % of the cycles) — Fourier-key LDGs straight from
// L2/HBM in the product (long scoreboard), one CTA of 16 warps per SM because accumulator +
// exchange buffers fill the shared memory, spills at N >= 8192 and, at N = 16384, the accumulator
// in global memory.  Here:
//   * a GROUP of T = M/P threads owns one GLWE polynomial of one ciphertext, P = 16 Fourier points
//     per thread (3 register passes for M = 2048..8192 instead of 4 with P = 8);
//   * the polynomial's shared-memory buffer (M x 16 B) is the accumulator's home between CMUXes
//     (rotated gather of ACC*X^a), THEN the exchange buffer of the transforms, THEN the parking
//     place of the forward transform for the partner groups: half the shared memory of the
//     group kernel, so CPB ciphertexts fit where one did;
//   * the thread's own accumulator pairs live in TENSOR MEMORY (tcgen05.ld/st 32x32b), which is
//     what makes that aliasing possible; for l >= 2 the Fourier accumulators are parked there too
//     during the next level's forward transform (no spills, no global-memory accumulator);
//   * the Fourier GGSW streams through a shared-memory ring by bulk TMA copies (cp.async.bulk +
//     mbarrier complete_tx) in blocks of RB rows (row p = the points {t*P + p} of the (k+1)^2
//     polynomials of one level), shared by every ciphertext of the CTA;
//   * exchanges are XOR-swizzled (conflict-free for every pass pattern), the key and the parked
//     transform are stored [p][t] (conflict-free LDS.128).
// Arithmetic: identical to pbs_kernel / pbs_kernel_grp (same rounding, same transforms, same
// key scaling); restates bootstrap.rs:250-368, ggsw.rs:481-602, glwe_sample_extraction.rs:90-165.
#pragma once

#ifndef V4_TW_TMEM
#define V4_TW_TMEM 1 // twiddles of the non-first passes in tensor memory (TwT below); 0 = global
#endif

#ifndef V4_TW_PF
#define V4_TW_PF 0 // 1 = global-memory twiddles fetched one stage ahead (passes without TwT).  Measured at
                   // P8 x 1024 (profiles/r2_p8_twiddle_ab.txt): 631.9 ms against 626.4 ms just in time - off
#endif
#ifndef V4_PROD_PIPE
#define V4_PROD_PIPE 1 // product phase (non-cluster): key and partner loads of row p+1 issued before the
                       // FMAs of row p, own term first and no branch on the polynomial index; 0 = first form
#endif
#ifndef V4_SHFL_X
#define V4_SHFL_X 1 // a last pass of ONE stage (M = 2^(LOGP*j + 1)) is reached by a lane-pair shuffle, not
                    // through shared memory (shfl_exchange below); 0 = every exchange through shared memory
#endif
#ifndef V4_ST_TMEM
#define V4_ST_TMEM 1 // two levels, 32-bit states: the states wait out the first level's transform and product in
                     // the (still empty) Fourier-accumulator columns of tensor memory instead of being spilled
#endif
#ifndef V4_TW_CONST
#define V4_TW_CONST 0 // 1 = warp-uniform twiddles of a non-first pass from the constant bank.  Measured at
                      // P8 x 1024: 625.6 ms against 626.4 ms from global memory - no effect, off
#endif

namespace v4 {
using namespace sm100;

template <int K_, int LOGN_, int LOGP_, int CPB_, bool STASH_, int RB_ = 4, int NB_ = 2,
          bool CL_ = false, bool PACK_ = false>
struct Cfg {
  // PACK: see PACK2 in cmux().  Off by default: measured on B200 at P8 x 1024 it is slower (648 ms
  // against 625 ms) although ptxas reports fewer spill bytes - the packing/unpacking integer work
  // sits on the critical path of a phase in which all 16 warps of the SM wait for each other.
  static constexpr bool PACK = PACK_;
  static constexpr int K = K_, LOGN = LOGN_, LOGM = LOGN_ - 1, M = 1 << LOGM, N = 2 * M;
  static constexpr int LOGP = LOGP_, P = 1 << LOGP_, T = M / P;
  // CL: the k+1 groups of ONE ciphertext are the k+1 CTAs of a thread-block cluster (one
  // polynomial per SM: N = 16384, where one polynomial buffer is 128 KiB).  A CTA then holds one
  // group, consumes only column r of every GGSW level (r = its rank in the cluster) and reads the
  // partners' parked transforms through distributed shared memory.
  static constexpr bool CL = CL_;
  static constexpr int GS = K_ + 1, CPB = CPB_, GROUPS = CL_ ? 1 : CPB_ * GS, THREADS = GROUPS * T;
  static constexpr int KEY_POLYS = CL_ ? GS : GS * GS; // key polynomials of a level a CTA consumes
  static constexpr int NWARPS = THREADS / 32;
  static constexpr bool STASH = STASH_; // l >= 2: Fourier accumulators parked in TMEM
  static constexpr u32 BUF_BYTES = (u32)M * 16;
  static constexpr u32 ROW_BYTES = (u32)T * 16;               // row p of one key polynomial
  // Key ring: NB blocks of RB rows (row p = the T points {t*P + p} of every polynomial of one GGSW
  // level).  One mbarrier wait and one completion count per BLOCK, not per row: a try_wait costs
  // ~90 cycles even when the data is there and the count is an atomic round trip, and with one of
  // each per row they were the critical path of the product (profiles/r2_v4a_*).  RB * NB rows is
  // what the shared memory left by the buffers holds: the whole level when it fits.
  // RB need not divide P: the last block of a level then has LAST_ROWS < RB rows (same slot stride)
  static constexpr int RB = RB_, NB = NB_, BLOCKS_PER_LEVEL = (P + RB_ - 1) / RB_;
  static constexpr int LAST_ROWS = P - (BLOCKS_PER_LEVEL - 1) * RB_;
  static constexpr u32 POLY_BLOCK_BYTES = (u32)RB_ * ROW_BYTES;         // RB rows of one polynomial
  static constexpr u32 BLOCK_BYTES = (u32)KEY_POLYS * POLY_BLOCK_BYTES; // RB rows of one level
  static constexpr u32 OFF_RING = (u32)GROUPS * BUF_BYTES;
  static constexpr u32 OFF_CTRL = OFF_RING + NB_ * BLOCK_BYTES;
  static constexpr size_t SMEM = (size_t)OFF_CTRL + 256;
  static constexpr u32 TM_OWN = 4 * P;                           // columns of a thread's own pairs
  static constexpr u32 TM_PER_WARP = TM_OWN * (STASH_ ? 2 : 1);
  static constexpr int NPASS = (LOGM + LOGP - 1) / LOGP;
  static_assert(T >= 32 && T % 32 == 0, "a group is a whole number of warps");
  static_assert(THREADS <= 1024 && THREADS % 32 == 0, "CTA size");
  static_assert(((NWARPS + 3) / 4) * TM_PER_WARP <= 512, "TMEM columns");
  static_assert(RB_ >= 1 && RB_ <= P && NB_ >= 1 && NB_ <= 16, "ring shape");
  static_assert(SMEM <= 232448, "shared memory");
  static_assert(T == 32 || CPB_ + GROUPS <= 15, "named barriers");
  static_assert(!CL_ || CPB_ == 1, "cluster variant: one ciphertext per cluster");
};

struct Ctrl {
  unsigned long long full_bar[16];
  unsigned int done_cnt[16];
  unsigned int tmem_base;
  unsigned int n, level, base_log, total_blocks;
  const double2 *bsk, *tw;
};

// Fresh, opaque read of the thread index: every phase of the CMUX re-derives its few address
// bases from it instead of keeping them (or, worse, every swizzled address of every exchange)
// live across the transforms — at 128 registers with 64 of them holding Fourier points there is
// no room for hoisted loop invariants (same device as pbs_w32_v3.cuh's V3Phase).
__device__ __forceinline__ int opaque_tid() {
  u32 tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  return (int)tid;
}
// twiddle load that stays where it is written (no hoisting above the previous stage)
__device__ __forceinline__ double2 ldg_tw(const double2 *p) {
  double2 w;
  asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(w.x), "=d"(w.y) : "l"(p));
  return w;
}

template <class C> __device__ __forceinline__ void grp_bar(int grp) {
  if constexpr (C::T == 32) {
    __syncwarp();
  } else {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + C::CPB + grp), "n"(C::T) : "memory");
  }
}
__device__ __forceinline__ u32 cluster_rank() {
  u32 r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// Measured alternatives at P8 x 1024 (release/acquire: 626.8 ms): relaxed arrive / plain wait, which
// order nothing, 591.5 ms (bit-identical results, but not a guarantee of the memory model, so not
// used); ONE fencing thread per CTA between two extra CTA barriers (formally equivalent by
// cumulativity) 642.1 ms: the cluster-scope fence is paid either way, and serialising it behind a
// CTA barrier costs more than letting the 16 warps issue it in parallel.
__device__ __forceinline__ void cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// every group of ciphertext `ct`: a named barrier inside the CTA, or the cluster barrier
template <class C> __device__ __forceinline__ void ct_bar(int ct) {
  if constexpr (C::CL) {
    cluster_arrive();
    cluster_wait();
  } else {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + ct), "n"(C::GS * C::T) : "memory");
  }
}
// (group index inside the CTA, polynomial index r of the group, ciphertext slot of the CTA)
template <class C> __device__ __forceinline__ int poly_of(int grp) {
  if constexpr (C::CL) return (int)cluster_rank();
  else return grp % C::GS;
}

// ---- twiddles of the non-first passes in tensor memory -----------------------------------------
// The first pass of a transform has compile-time twiddles (constant bank).  Every later pass needs
// P/2 twiddles that depend on the thread (tw[2^s + b], b = the index bits above the paired one);
// read from global memory they are an L1 miss per stage at N >= 4096 (the table is 32-64 KiB, the
// L1 left beside 227 KiB of shared memory is not), i.e. an L2 round trip in front of every stage
// of two passes of both transforms (ncu, P7: long scoreboard 41 % of the stalls of the inverse's
// first pass).  With one level (no Fourier-accumulator stash) half of the tensor memory is free:
// each thread keeps its (NPASS - 1) * P/2 twiddles there (tcgen05.ld, 12 cycles) - the v3 device.
// Slot of the twiddle of stage rho (in-thread bit paired) and block pair h >> 1 inside a pass:
// rho = LOGP-1 -> 0; rho = LOGP-2 -> 1; rho <= LOGP-3 -> 2^(LOGP-rho-2) + (h >> 1).
// (32 points per thread with 64 twiddle columns per pass was built and measured for N = 4096 / 8192 -
// 8 warps, 255 registers, 5 + 5 + 1 / 5 + 5 + 2 stages - and lost to 16 points per thread: P6 x 4096
// 167.0 against 150.3 ms, P7 x 4096 417.3 against 407.0 ms; profiles/r2_p32_one_level_ab.txt)
template <class C> struct TwT {
  static constexpr int SLOTS = C::P / 2;
  static constexpr u32 PASS_COLS = (u32)SLOTS * 4;
  static constexpr u32 COLS = (u32)(C::NPASS - 1) * PASS_COLS;
  static constexpr u32 QUAD_WARPS = (C::NWARPS + 3) / 4;
  static constexpr bool ON = V4_TW_TMEM && !C::STASH && C::NPASS > 1 && C::LOGP >= 3 && C::LOGP <= 4 &&
                             QUAD_WARPS * (C::TM_PER_WARP + COLS) <= 512;
  static constexpr u32 BASE = QUAD_WARPS * C::TM_PER_WARP; // first twiddle column of a quadrant
};
__host__ __device__ constexpr int tw_slot_base(int logp, int rho) {
  return rho == logp - 1 ? 0 : 1 << (logp - rho - 2);
}
__host__ __device__ constexpr int tw_slot_count(int logp, int rho) {
  return rho >= logp - 2 ? 1 : 1 << (logp - rho - 2);
}
// twiddle columns of this thread for pass `pass` (>= 1)
template <class C> __device__ __forceinline__ u32 tw_tmem(int pass) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const u32 warp = (u32)opaque_tid() >> 5;
  return reinterpret_cast<const Ctrl *>(smem + C::OFF_CTRL)->tmem_base + (((warp & 3u) * 32u) << 16) +
         TwT<C>::BASE + (warp >> 2) * TwT<C>::COLS + (u32)(pass - 1) * TwT<C>::PASS_COLS;
}
// the CNT (1, 2 or 4) twiddles of one stage: one tcgen05.ld
template <int CNT> __device__ __forceinline__ void tw_load(u32 taddr, double2 (&w)[4]) {
  u32 v[16];
  if constexpr (CNT == 1) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr)
                 : "memory");
  } else if constexpr (CNT == 2) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]),
                   "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr)
                 : "memory");
  } else {
    tmem_ld16(taddr, v);
  }
  tmem_wait_ld();
#pragma unroll
  for (int e = 0; e < CNT; ++e)
    w[e] = make_double2(__hiloint2double((int)v[4 * e + 1], (int)v[4 * e + 0]),
                        __hiloint2double((int)v[4 * e + 3], (int)v[4 * e + 2]));
}
template <class C, int RHO> __device__ __forceinline__ void tw_stage(u32 tm_pass, double2 (&w)[4]) {
  tw_load<tw_slot_count(C::LOGP, RHO)>(tm_pass + 4u * (u32)tw_slot_base(C::LOGP, RHO), w);
}
// stage `rho` is only known after unrolling: dispatch on it (folds to one call)
template <class C> __device__ __forceinline__ void tw_stage_dyn(int rho, u32 tm_pass, double2 (&w)[4]) {
  if (rho == 0) tw_stage<C, 0>(tm_pass, w);
  else if (rho == 1) tw_stage<C, 1>(tm_pass, w);
  else if (rho == 2) tw_stage<C, 2>(tm_pass, w);
  else if constexpr (C::LOGP > 3) tw_stage<C, 3>(tm_pass, w);
}

// ---- twiddles of the non-first passes without TwT (two levels: the tensor memory is full) --------
// Two alternatives to the just-in-time global load, both measured at P8 and both OFF (a negative
// result: the twiddle latency is not what bounds the two-level kernel; the exchanges and the
// product phase are - profiles/r2_v4cl_P8_pf1c1_ncu.txt):
// (a) V4_TW_CONST: a pass whose in-thread bits start at lo >= 5 has block indices that depend on the
//     WARP only: stages s < 8 index the 256-entry constant table with a warp-uniform offset
//     (N = 16384: the second pass);
// (b) V4_TW_PF: every other stage reads its tw_slot_count twiddles from global memory ONE STAGE
//     AHEAD of their use (the first stage of a pass right after the exchange writes, when the
//     points' registers are dead).
template <class C> __host__ __device__ constexpr bool tw_is_const(int lo, int s) {
  return V4_TW_CONST && lo + C::LOGP != C::LOGM && lo >= 5 && (2 << s) <= cbfft::CONST_TW;
}
template <class C>
__device__ __forceinline__ void tw_fetch(const double2 *__restrict__ tw, int t, int lo, int s,
                                         double2 (&w)[C::P / 4]) {
  const int rho = C::LOGM - 1 - s - lo;
#pragma unroll
  for (int kk = 0; kk < tw_slot_count(C::LOGP, rho); ++kk) {
    const int b = ((t >> lo) << (C::LOGP - rho - 1)) | (2 * kk);
    w[kk] = tw_is_const<C>(lo, s) ? cbfft::c_twg[(1 << s) + b] : ldg_tw(&tw[(1 << s) + b]);
  }
}

// ---- transforms: the passes of fft.cuh with an XOR-swizzled exchange ---------------------------
template <class C> __device__ __forceinline__ int swz(int idx) {
  return idx ^ ((idx >> C::LOGP) & 7);
}
template <class C> __device__ __forceinline__ constexpr int pass_lo(int pass) {
  return (pass == C::NPASS - 1) ? 0 : (C::LOGM - C::LOGP * (pass + 1));
}

// stage s of the pass with in-thread bits [lo, lo+LOGP) takes its twiddles from registers filled
// one stage earlier (tw_fetch)
template <class C, bool TWM> __host__ __device__ constexpr bool tw_piped(int lo, int s) {
  return V4_TW_PF && lo + C::LOGP != C::LOGM && !TWM && !tw_is_const<C>(lo, s);
}

// stages [s_begin, s_end) on in-thread data (fft.cuh fwd_stages / inv_stages with pinned loads).
// wq: the twiddles of the first stage when that stage is tw_piped (fetched by the caller).
template <class C, bool TWM = false>
__device__ __forceinline__ void fwd_stages(double2 (&x)[C::P], int t, int lo, int s_begin,
                                           int s_end, const double2 *__restrict__ tw, u32 tm_pass,
                                           double2 (&wq)[C::P / 4]) {
  constexpr int P = C::P, LOGP = C::LOGP, LOGM = C::LOGM;
#pragma unroll
  for (int s = s_begin; s < s_end; ++s) {
    const int beta = LOGM - 1 - s, rho = beta - lo;
    const bool from_tmem = TWM && (lo + LOGP != LOGM);
    const bool piped = tw_piped<C, TWM>(lo, s);
    const bool fetch_next = s + 1 < s_end && tw_piped<C, TWM>(lo, s + 1);
    double2 wt[4], wn[P / 4];
    if (from_tmem) tw_stage_dyn<C>(rho, tm_pass, wt);
    if (fetch_next) tw_fetch<C>(tw, t, lo, s + 1, wn);
#pragma unroll
    for (int r = 0; r < P; ++r) {
      if (r & (1 << rho)) continue;
      if (rho + 1 < LOGP && (r & (1 << (rho + 1)))) continue;
      const int idx = cbfft::elem_index<LOGP>(t, r, lo);
      const bool first_pass = (lo + LOGP == LOGM) && ((1 << s) + (r >> (rho + 1)) < cbfft::CONST_TW);
      const int b = first_pass ? (r >> (rho + 1)) : (idx >> (beta + 1));
      const double2 w = from_tmem                               ? wt[(r >> (rho + 1)) >> 1]
                        : piped                                 ? wq[(r >> (rho + 1)) >> 1]
                        : (first_pass || tw_is_const<C>(lo, s)) ? cbfft::c_twg[(1 << s) + b]
                                                                : ldg_tw(&tw[(1 << s) + b]);
      cbfft::bfly_ct(x[r], x[r | (1 << rho)], w);
      if (rho + 1 < LOGP) {
        const int r2 = r | (1 << (rho + 1));
        cbfft::bfly_ct(x[r2], x[r2 | (1 << rho)], make_double2(-w.y, w.x));
      }
    }
    if (fetch_next) {
#pragma unroll
      for (int e = 0; e < P / 4; ++e) wq[e] = wn[e];
    }
  }
}
template <class C, bool TWM = false>
__device__ __forceinline__ void inv_stages(double2 (&x)[C::P], int t, int lo, int s_hi, int s_lo,
                                           const double2 *__restrict__ tw, u32 tm_pass,
                                           double2 (&wq)[C::P / 4]) {
  constexpr int P = C::P, LOGP = C::LOGP, LOGM = C::LOGM;
#pragma unroll
  for (int s = s_hi - 1; s >= s_lo; --s) {
    const int beta = LOGM - 1 - s, rho = beta - lo;
    const bool from_tmem = TWM && (lo + LOGP != LOGM);
    const bool piped = tw_piped<C, TWM>(lo, s);
    const bool fetch_next = s - 1 >= s_lo && tw_piped<C, TWM>(lo, s - 1);
    double2 wt[4], wn[P / 4];
    if (from_tmem) tw_stage_dyn<C>(rho, tm_pass, wt);
    if (fetch_next) tw_fetch<C>(tw, t, lo, s - 1, wn);
#pragma unroll
    for (int r = 0; r < P; ++r) {
      if (r & (1 << rho)) continue;
      if (rho + 1 < LOGP && (r & (1 << (rho + 1)))) continue;
      const int idx = cbfft::elem_index<LOGP>(t, r, lo);
      const bool first_pass = (lo + LOGP == LOGM) && ((1 << s) + (r >> (rho + 1)) < cbfft::CONST_TW);
      const int b = first_pass ? (r >> (rho + 1)) : (idx >> (beta + 1));
      const double2 w = from_tmem                               ? wt[(r >> (rho + 1)) >> 1]
                        : piped                                 ? wq[(r >> (rho + 1)) >> 1]
                        : (first_pass || tw_is_const<C>(lo, s)) ? cbfft::c_twg[(1 << s) + b]
                                                                : ldg_tw(&tw[(1 << s) + b]);
      cbfft::bfly_gs(x[r], x[r | (1 << rho)], w);
      if (rho + 1 < LOGP) {
        const int r2 = r | (1 << (rho + 1));
        cbfft::bfly_gs(x[r2], x[r2 | (1 << rho)], make_double2(-w.y, w.x));
      }
    }
    if (fetch_next) {
#pragma unroll
      for (int e = 0; e < P / 4; ++e) wq[e] = wn[e];
    }
  }
}

// ---- exchange by shuffle when the last pass has ONE stage --------------------------------------------
// Layout lo = 1 (bits [1, LOGP] of the index in-thread, bit 0 = bit 0 of t) <-> layout lo = 0 (bits
// [0, LOGP) in-thread): only index bit 0 and bit LOGP trade places, and both layouts keep bits
// above LOGP in t >> 1 - the exchange is a swap of half of the points between the lanes t and
// t ^ 1.  Forward (lo 1 -> 0): old slot r of lane parity h goes to lane parity r >> (LOGP-1), new
// slot 2 (r mod P/2) + h; the inverse direction is the mirror.  No shared memory, no barrier:
// N = 16384 (13 = 4 + 4 + 4 + 1 stages) saves three of its nine 2 x 128 KiB exchanges per CMUX,
// k = 2 / N = 1024 (9 = 4 + 4 + 1) three of six.  Same values in the same registers as the
// shared-memory exchange: results are bit-identical.
template <class C> struct ShX {
  static constexpr int LAST_G = C::LOGM - C::LOGP * (C::NPASS - 1);
  static constexpr bool ON = V4_SHFL_X && C::NPASS >= 2 && LAST_G == 1 && C::LOGP >= 2;
};
__device__ __forceinline__ double2 shfl_xor1(const double2 v) {
  return make_double2(__shfl_xor_sync(0xffffffffu, v.x, 1), __shfl_xor_sync(0xffffffffu, v.y, 1));
}
template <class C> __device__ __forceinline__ void shfl_exchange_fwd(double2 (&x)[C::P], int t) {
  constexpr int H = C::P / 2;
  const bool h = t & 1;
  double2 y[C::P];
#pragma unroll
  for (int m = 0; m < H; ++m) {
    const double2 a = x[m], b = x[H + m];
    const double2 keep = h ? b : a, recv = shfl_xor1(h ? a : b);
    y[2 * m] = h ? recv : keep;
    y[2 * m + 1] = h ? keep : recv;
  }
#pragma unroll
  for (int r = 0; r < C::P; ++r) x[r] = y[r];
}
template <class C> __device__ __forceinline__ void shfl_exchange_inv(double2 (&x)[C::P], int t) {
  constexpr int H = C::P / 2;
  const bool h = t & 1;
  double2 y[C::P];
#pragma unroll
  for (int m = 0; m < H; ++m) {
    const double2 a = x[2 * m], b = x[2 * m + 1];
    const double2 keep = h ? b : a, recv = shfl_xor1(h ? a : b);
    y[m] = h ? recv : keep;
    y[H + m] = h ? keep : recv;
  }
#pragma unroll
  for (int r = 0; r < C::P; ++r) x[r] = y[r];
}

// In: x[r] = z_{t + r*T}.  Out: x[r] = F_{q = t*P + r}.  Every exchange is barrier / write /
// barrier / read, so the buffer may have other readers right up to the call.
// `pending` (cluster variant): the partners may still be reading this CTA's buffer (their product
// phase); the cluster barrier they arrive at afterwards is waited for right before the first
// exchange write, i.e. behind the register-only stages of the first pass.
template <class C, bool TWM = false>
__device__ __forceinline__ void forward(double2 (&x)[C::P], const double2 *__restrict__ tw,
                                        bool pending = false) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 wq[C::P / 4];
#pragma unroll
  for (int pass = 0; pass < C::NPASS; ++pass) {
    const int tid = opaque_tid(), t = tid % C::T, grp = tid / C::T;
    double2 *buf = reinterpret_cast<double2 *>(smem + (size_t)grp * C::BUF_BYTES);
    const int lo = pass_lo<C>(pass);
    const int s0 = pass * C::LOGP;
    const int s1 = (pass == C::NPASS - 1) ? C::LOGM : s0 + C::LOGP;
    fwd_stages<C, TWM>(x, t, lo, s0, s1, tw, (TWM && pass > 0) ? tw_tmem<C>(pass) : 0u, wq);
    if (ShX<C>::ON && pass + 2 == C::NPASS) {
      if constexpr (C::CL) { // (a two-pass transform has no shared-memory exchange to wait in front of)
        if (pass == 0 && pending) cluster_wait();
      }
      shfl_exchange_fwd<C>(x, t);
      if (tw_piped<C, TWM>(0, s1)) tw_fetch<C>(tw, t, 0, s1, wq);
    } else if (pass + 1 < C::NPASS) {
      const int lo2 = pass_lo<C>(pass + 1);
      if constexpr (C::CL) {
        if (pass == 0 && pending) cluster_wait();
      }
      grp_bar<C>(grp);
#pragma unroll
      for (int r = 0; r < C::P; ++r) buf[swz<C>(cbfft::elem_index<C::LOGP>(t, r, lo))] = x[r];
      // the points' registers are dead until the exchange reads: the next pass's first twiddles
      if (tw_piped<C, TWM>(lo2, s1)) tw_fetch<C>(tw, t, lo2, s1, wq);
      grp_bar<C>(grp);
#pragma unroll
      for (int r = 0; r < C::P; ++r) x[r] = buf[swz<C>(cbfft::elem_index<C::LOGP>(t, r, lo2))];
    }
  }
}

// In: x[r] = F_{t*P + r}.  Out: x[r] = M * z_{t + r*T} (unnormalised).
template <class C, bool TWM = false>
__device__ __forceinline__ void inverse(double2 (&x)[C::P], const double2 *__restrict__ tw,
                                        bool pending = false) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 wq[C::P / 4];
  if (C::NPASS > 1 && tw_piped<C, TWM>(0, C::LOGM - 1))
    tw_fetch<C>(tw, opaque_tid() % C::T, 0, C::LOGM - 1, wq);
#pragma unroll
  for (int pass = C::NPASS - 1; pass >= 0; --pass) {
    const int tid = opaque_tid(), t = tid % C::T, grp = tid / C::T;
    double2 *buf = reinterpret_cast<double2 *>(smem + (size_t)grp * C::BUF_BYTES);
    const int lo = pass_lo<C>(pass);
    const int s0 = pass * C::LOGP;
    const int s1 = (pass == C::NPASS - 1) ? C::LOGM : s0 + C::LOGP;
    inv_stages<C, TWM>(x, t, lo, s1, s0, tw, (TWM && pass > 0) ? tw_tmem<C>(pass) : 0u, wq);
    // `pending`: the cluster barrier is waited for in front of the FIRST write to shared memory
    constexpr int FIRST_SMEM_PASS = ShX<C>::ON ? C::NPASS - 2 : C::NPASS - 1;
    if (ShX<C>::ON && pass == C::NPASS - 1) {
      shfl_exchange_inv<C>(x, t);
      const int lo2 = pass_lo<C>(pass - 1);
      if (tw_piped<C, TWM>(lo2, s0 - 1)) tw_fetch<C>(tw, t, lo2, s0 - 1, wq);
      if constexpr (C::CL) {
        if (FIRST_SMEM_PASS == 0 && pending) cluster_wait(); // two-pass transform: nothing else waits
      }
    } else if (pass > 0) {
      const int lo2 = pass_lo<C>(pass - 1);
      if constexpr (C::CL) {
        if (pass == FIRST_SMEM_PASS && pending) cluster_wait();
      }
      grp_bar<C>(grp);
#pragma unroll
      for (int r = 0; r < C::P; ++r) buf[swz<C>(cbfft::elem_index<C::LOGP>(t, r, lo))] = x[r];
      if (tw_piped<C, TWM>(lo2, s0 - 1)) tw_fetch<C>(tw, t, lo2, s0 - 1, wq);
      grp_bar<C>(grp);
#pragma unroll
      for (int r = 0; r < C::P; ++r) x[r] = buf[swz<C>(cbfft::elem_index<C::LOGP>(t, r, lo2))];
    }
  }
}

// ---- key ring -----------------------------------------------------------------------------------
// Block number `blk` (counted from the start of the kernel) = rows [b*RB, (b+1)*RB) of the
// (k+1)^2 Fourier polynomials of level m of GGSW i -> slot blk % NB.  The RB rows of one
// polynomial are contiguous in the key (layout [p][t]): one bulk copy per polynomial.
template <class C> __device__ __noinline__ void issue_block(u32 i, u32 m, u32 b, u32 slot) {
  extern __shared__ __align__(1024) unsigned char smem[];
  Ctrl *ctrl = reinterpret_cast<Ctrl *>(smem + C::OFF_CTRL);
  const double2 *bsk = ctrl->bsk;
  const u32 level = ctrl->level;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic reads before async write
  const u32 poly_bytes = (b + 1 == (u32)C::BLOCKS_PER_LEVEL ? (u32)C::LAST_ROWS : (u32)C::RB) * C::ROW_BYTES;
  mbar_expect_tx(&ctrl->full_bar[slot], (u32)C::KEY_POLYS * poly_bytes);
  // cluster variant: this CTA consumes column r only: key polynomials (rr, r), rr = 0..k
  const size_t col = C::CL ? (size_t)cluster_rank() : 0, pstride = C::CL ? C::GS : 1;
  const double2 *src = bsk + (((size_t)i * level + m) * C::GS * C::GS + col) * C::M +
                       (size_t)b * C::RB * C::T;
  unsigned char *dst = smem + C::OFF_RING + slot * C::BLOCK_BYTES;
#pragma unroll
  for (int poly = 0; poly < C::KEY_POLYS; ++poly)
    tma_load_bulk(dst + poly * C::POLY_BLOCK_BYTES, src + (size_t)poly * pstride * C::M, poly_bytes,
                  &ctrl->full_bar[slot]);
  // Keys that do not fit the 126 MB L2 (N >= 4096: 115 MB .. 240 MB): all CTAs walk the GGSW index
  // together, so ONE early request per block turns every later fill into an L2 hit.  An eighth of
  // the CTAs asks for the same block two GGSWs ahead.
  if (C::LOGN >= 12 && (blockIdx.x & 7u) == (b & 7u) && i + 2 < ctrl->n) {
    const double2 *ahead = src + (size_t)2 * level * C::GS * C::GS * C::M;
#pragma unroll
    for (int poly = 0; poly < C::KEY_POLYS; ++poly)
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ahead +
                                                                       (size_t)poly * pstride * C::M),
                   "r"(poly_bytes)
                   : "memory");
  }
}

// (slot, parity) of a block: slot = blk % NB, parity = (blk / NB) & 1
struct RingPos {
  u32 slot, parity;
};
template <class C> __device__ __forceinline__ RingPos ring_pos(u32 blk) {
  RingPos r;
  r.slot = blk % (u32)C::NB;
  r.parity = (blk / (u32)C::NB) & 1u;
  return r;
}
template <class C> __device__ __forceinline__ void ring_next(RingPos &r) {
  if (++r.slot == (u32)C::NB) {
    r.slot = 0;
    r.parity ^= 1u;
  }
}
template <class C> __device__ __forceinline__ void wait_block(const RingPos &r) {
  extern __shared__ __align__(1024) unsigned char smem[];
  Ctrl *ctrl = reinterpret_cast<Ctrl *>(smem + C::OFF_CTRL);
  mbar_wait(&ctrl->full_bar[r.slot], r.parity);
}

// This warp has issued its last read of block `blk` (level m, index b in the level): count it;
// the warp that counts last refills the slot with the block NB ahead (if there is one).
template <class C>
__device__ __forceinline__ void block_done(int lane, u32 i, u32 m, u32 b, u32 blk, u32 slot) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __syncwarp();
  if (lane == 0) {
    const Ctrl *ctrl = reinterpret_cast<const Ctrl *>(smem + C::OFF_CTRL);
    u32 old;
    asm volatile("atom.shared.inc.u32 %0, [%1], %2;" // wraps to 0 after the last warp
                 : "=r"(old)
                 : "r"(smem_u32(smem) + C::OFF_CTRL + 128 + 4 * slot), "n"(C::NWARPS - 1)
                 : "memory");
    if (old == (u32)C::NWARPS - 1 && blk + C::NB < ctrl->total_blocks) {
      u32 bn = b + C::NB, mn = m, in = i;
      while (bn >= (u32)C::BLOCKS_PER_LEVEL) {
        bn -= C::BLOCKS_PER_LEVEL;
        if (++mn == ctrl->level) {
          mn = 0;
          in = (in + 1 == ctrl->n) ? 0 : in + 1;
        }
      }
      issue_block<C>(in, mn, bn, slot);
    }
  }
}

// a ciphertext slot with nothing to do in this CMUX (idle slot of the last round, or a_hat == 0)
// still walks the ring: every warp of the CTA consumes every block
template <class C> __device__ __noinline__ void idle_cmux(u32 i, u32 blk0) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int lane = opaque_tid() & 31;
  const u32 level = reinterpret_cast<const Ctrl *>(smem + C::OFF_CTRL)->level;
  RingPos rp = ring_pos<C>(blk0);
  u32 blk = blk0;
  for (u32 m = 0; m < level; ++m) {
#pragma unroll 1
    for (u32 b = 0; b < (u32)C::BLOCKS_PER_LEVEL; ++b, ++blk) {
      wait_block<C>(rp);
      block_done<C>(lane, i, m, b, blk, rp.slot);
      ring_next<C>(rp);
    }
  }
}

__device__ __forceinline__ void unpack4(const u32 (&v)[16], u64 (&re)[4], u64 (&im)[4]) {
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    re[e] = ((u64)v[4 * e + 1] << 32) | v[4 * e + 0];
    im[e] = ((u64)v[4 * e + 3] << 32) | v[4 * e + 2];
  }
}
__device__ __forceinline__ void pack4(u32 (&v)[16], const u64 (&re)[4], const u64 (&im)[4]) {
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    v[4 * e + 0] = (u32)re[e];
    v[4 * e + 1] = (u32)(re[e] >> 32);
    v[4 * e + 2] = (u32)im[e];
    v[4 * e + 3] = (u32)(im[e] >> 32);
  }
}

// One CMUX of polynomial r of one ciphertext (one group).  Out of line on purpose: the caller's
// loop state (round, sample, LWE pointer) must not stay live across the transforms; everything
// the body needs beyond its three arguments comes from the thread index or the control block.
template <class C, bool NARROW, int LV>
__device__ __noinline__ void cmux(const u32 a_hat, const u32 i, const u32 blk0) {
  using st_t = typename std::conditional<NARROW, u32, u64>::type;
  constexpr int P = C::P, T = C::T, M = C::M, N = C::N, LOGM = C::LOGM, GS = C::GS;
  extern __shared__ __align__(1024) unsigned char smem[];
  const Ctrl *ctrl = reinterpret_cast<const Ctrl *>(smem + C::OFF_CTRL);
  const double2 *__restrict__ tw = ctrl->tw;
  const u32 base_log = ctrl->base_log, level = LV ? (u32)LV : ctrl->level;

  // ---- phase A: ct1_r = ACC_r * X^a - ACC_r, rounding, decomposition states
  // PACK2 (two levels, 32-bit states): the first level's digits are taken right here and what is
  // left of the two states of a point (the top digit plus a carry: <= 2^base_log each, and
  // base_log <= 15 since 2 * base_log <= 31) is packed into ONE register, so 16 registers less
  // are live across the first forward transform and product (ptxas: 532 -> 128 B of spills)
  constexpr bool PACK2 = (LV == 2) && NARROW && C::PACK;
  // ST_TM (two levels, 32-bit states, no packing): after the first level's digits the 2 P state words
  // of a thread are parked in the first 2 P of the 4 P tensor-memory columns that will hold the
  // Fourier accumulators of level 0 - those are written only after the level-0 product, right
  // after the states have been read back.  ptxas spilled them instead (256 B of stores and 128 B
  // of loads per thread and CMUX through a 28 KiB L1 next to 227 KiB of shared memory, i.e. L2).
  constexpr bool ST_TM = V4_ST_TMEM && C::STASH && LV == 2 && NARROW && !PACK2 && (2 * P) % 16 == 0;
  // FAST1 (one level): closest_representable + the balanced digit straight from the HIGH word of
  // ct1 in two integer instructions (the rounding bit 63 - base_log lies in the high word for every
  // base_log <= 31; same derivation as pbs_w32_v3.cuh neg_digit_l1_k: round-half-up of x is minus
  // round-half-down of -x, taken by an arithmetic shift; the tie lands on +B/2 as in
  // decomposer.rs:23-47 + iter.rs:121-133).  ncu on P7 (profiles/r2_v4_P7_ncu.txt): phase A was
  // 15.7 % of the CMUX, bound by the integer pipe.
  constexpr bool FAST1 = (LV == 1);
  st_t st_re[(PACK2 || FAST1) ? 1 : P], st_im[(PACK2 || FAST1) ? 1 : P];
  u32 pk[PACK2 ? P : 1];
  double2 x[P];
  {
    const int tid = opaque_tid(), t = tid % T, grp = tid / T, warp = tid >> 5;
    const ulonglong2 *acc = reinterpret_cast<const ulonglong2 *>(smem + (size_t)grp * C::BUF_BYTES);
    const u32 tm_own = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) +
                       (u32)(warp >> 2) * C::TM_PER_WARP;
    const u32 round_shift = 64 - base_log * level - 1;
    const u32 state_shift = 64 - base_log * level;
#pragma unroll
    for (int c4 = 0; c4 < P / 4; ++c4) {
      u32 v[16];
      tmem_ld16(tm_own + 16 * c4, v);
      tmem_wait_ld();
      u64 own_re[4], own_im[4];
      unpack4(v, own_re, own_im);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int p = 4 * c4 + e;
        const int j = t + p * T;
        const u32 s = (u32)(j - (int)a_hat) & (2 * N - 1);
        const u32 q = s >> LOGM;
        const ulonglong2 src = acc[s & (M - 1)];
        const u64 ra = (q & 1) ? src.y : src.x;
        const u64 rb = (q & 1) ? src.x : src.y;
        const u64 rot_re = (q == 2 || q == 3) ? (u64)0 - ra : ra;
        const u64 rot_im = (q == 1 || q == 2) ? (u64)0 - rb : rb;
        if constexpr (FAST1) {
          const u32 kshift = 32 - base_log;
          const u32 half_m1 = (1u << (kshift - 1)) - 1u;
          const u32 h_re = (u32)((rot_re - own_re[e]) >> 32), h_im = (u32)((rot_im - own_im[e]) >> 32);
          x[p].x = -__int2double_rn((int)(half_m1 - h_re) >> kshift);
          x[p].y = -__int2double_rn((int)(half_m1 - h_im) >> kshift);
          continue;
        }
        const st_t s_re = (st_t)(closest_representable(rot_re - own_re[e], round_shift) >> state_shift);
        const st_t s_im = (st_t)(closest_representable(rot_im - own_im[e], round_shift) >> state_shift);
        if constexpr (PACK2) {
          const u32 mask = (1u << base_log) - 1u;
          u32 a = (u32)s_re, b = (u32)s_im;
          x[p].x = __int2double_rn((int)grp_decompose<u32>(a, mask, base_log));
          x[p].y = __int2double_rn((int)grp_decompose<u32>(b, mask, base_log));
          pk[p] = a | (b << 16);
        } else {
          st_re[p] = s_re;
          st_im[p] = s_im;
        }
      }
    }
  }

#pragma unroll 1
  for (u32 m = 0; m < level; ++m) { // memory level index m <-> level l-m (ggsw.rs:231-245)
    if constexpr (FAST1) {
      // digits taken in phase A
    } else if constexpr (PACK2) {
      if (m > 0) {
        const u32 mask = (1u << base_log) - 1u;
#pragma unroll
        for (int p = 0; p < P; ++p) {
          u32 a = pk[p] & 0xFFFFu, b = pk[p] >> 16;
          x[p].x = __int2double_rn((int)grp_decompose<u32>(a, mask, base_log));
          x[p].y = __int2double_rn((int)grp_decompose<u32>(b, mask, base_log));
        }
      }
    } else {
      const st_t mask = (st_t)(((u64)1 << base_log) - 1);
#pragma unroll
      for (int p = 0; p < P; ++p) {
        x[p].x = __int2double_rn((int)(u32)grp_decompose<st_t>(st_re[p], mask, base_log));
        x[p].y = __int2double_rn((int)(u32)grp_decompose<st_t>(st_im[p], mask, base_log));
      }
      if constexpr (ST_TM) {
        if (m == 0) {
          const int warp = opaque_tid() >> 5;
          const u32 tm_st = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) +
                            (u32)(warp >> 2) * C::TM_PER_WARP + C::TM_OWN;
#pragma unroll
          for (int c = 0; c < 2 * P / 16; ++c) {
            u32 v[16];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              v[2 * e + 0] = (u32)st_re[8 * c + e];
              v[2 * e + 1] = (u32)st_im[8 * c + e];
            }
            tmem_st16(tm_st + 16 * c, v);
          }
          tmem_wait_st();
        }
      }
    }
    // its first barrier also ends the rotated reads of ACC; for m > 0 of the cluster variant the
    // partners' reads of the previous level's parked transform end at the pending cluster barrier
    forward<C, TwT<C>::ON>(x, tw, C::CL && m > 0);
    {
      const int tid = opaque_tid(), t = tid % T, grp = tid / T;
      double2 *buf = reinterpret_cast<double2 *>(smem + (size_t)grp * C::BUF_BYTES);
      grp_bar<C>(grp); // last exchange fully read
#pragma unroll
      for (int p = 0; p < P; ++p) buf[p * T + t] = x[p]; // F_{r,m} parked in key order
      ct_bar<C>(grp / GS);                               // every F_{.,m} of this ciphertext visible
    }

    // ---- phase B: column r of the external product, in place in x
    {
      const int tid = opaque_tid(), t = tid % T, grp = tid / T, warp = tid >> 5, lane = tid & 31;
      const int ct = C::CL ? 0 : grp / GS, r = poly_of<C>(grp);
      const u32 tm_stash = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) +
                           (u32)(warp >> 2) * C::TM_PER_WARP + C::TM_OWN;
      // key polynomial (rr, r) of the block: at rr * GS + r (whole level staged) or at rr (cluster
      // variant: only column r is staged)
      const u32 ring = smem_u32(smem) + C::OFF_RING + (u32)t * 16 +
                       (C::CL ? 0u : (u32)r * C::POLY_BLOCK_BYTES);
      constexpr u32 KEY_RR_STRIDE = (u32)GS * C::POLY_BLOCK_BYTES;
      const u32 sib = smem_u32(smem) + (u32)(ct * GS) * C::BUF_BYTES + (u32)t * 16;
      constexpr bool PIPE = V4_PROD_PIPE && !C::CL;
      u32 pk_key[GS], pk_f[GS];
#pragma unroll
      for (int j = 0; j < GS; ++j) {
        const u32 rr = (u32)(r + j) % (u32)GS;
        pk_key[j] = rr * KEY_RR_STRIDE;
        pk_f[j] = sib + rr * C::BUF_BYTES;
      }
      // cluster variant: the partners' buffers through distributed shared memory, fetched one
      // ring block ahead of their use (a remote load takes ~200 cycles)
      // (partner j of this CTA is the CTA of rank (r + 1 + j) mod (k+1))
      constexpr int NSIB = C::CL ? GS - 1 : 1;
      u32 sib_cl[NSIB], key_cl[NSIB];
      double2 fs[NSIB][C::RB];
      if constexpr (C::CL) {
#pragma unroll
        for (int j = 0; j < NSIB; ++j) {
          const u32 rr = (u32)(r + 1 + j) % (u32)GS;
          key_cl[j] = rr * C::POLY_BLOCK_BYTES;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;"
                       : "=r"(sib_cl[j])
                       : "r"(smem_u32(smem) + (u32)t * 16), "r"(rr));
        }
#pragma unroll
        for (int j = 0; j < NSIB; ++j)
#pragma unroll
          for (int pr = 0; pr < C::RB; ++pr)
            asm volatile("ld.shared::cluster.v2.f64 {%0, %1}, [%2];"
                         : "=d"(fs[j][pr].x), "=d"(fs[j][pr].y)
                         : "r"(sib_cl[j] + (u32)(pr * T * 16)));
      }
      u32 blk = blk0 + m * C::BLOCKS_PER_LEVEL;
      RingPos rp = ring_pos<C>(blk);
#pragma unroll
      for (int b = 0; b < C::BLOCKS_PER_LEVEL; ++b, ++blk) {
        wait_block<C>(rp);
        const u32 slot = ring + rp.slot * C::BLOCK_BYTES;
        double2 fn[NSIB][C::RB];
        if constexpr (C::CL) {
          if (b + 1 < C::BLOCKS_PER_LEVEL) {
#pragma unroll
            for (int j = 0; j < NSIB; ++j)
#pragma unroll
              for (int pr = 0; pr < C::RB; ++pr) {
                if ((b + 1) * C::RB + pr >= P) continue; // short last block
                asm volatile("ld.shared::cluster.v2.f64 {%0, %1}, [%2];"
                             : "=d"(fn[j][pr].x), "=d"(fn[j][pr].y)
                             : "r"(sib_cl[j] + (u32)(((b + 1) * C::RB + pr) * T * 16)));
              }
          }
        }
        // V4_PROD_PIPE: (key, partner transform) operands of the current and of the next row; slot j is
        // polynomial (r + j) mod (k+1), j = 0 the own term (operand x[p], key only)
        double2 wc[GS], fc[GS], wn[GS], fn2[GS];
        auto load_row = [&](int pr, double2(&w)[GS], double2(&f)[GS]) {
#pragma unroll
          for (int j = 0; j < GS; ++j) {
            asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                         : "=d"(w[j].x), "=d"(w[j].y)
                         : "r"(slot + pk_key[j] + (u32)pr * C::ROW_BYTES));
            if (j > 0)
              asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                           : "=d"(f[j].x), "=d"(f[j].y)
                           : "r"(pk_f[j] + (u32)((b * C::RB + pr) * T * 16)));
          }
        };
        if constexpr (PIPE) load_row(0, wc, fc);
#pragma unroll
        for (int pr = 0; pr < C::RB; ++pr) {
          const int p = b * C::RB + pr;
          if (p >= P) continue; // short last block (RB does not divide P)
          const bool more_rows = pr + 1 < C::RB && p + 1 < P;
          double2 o = make_double2(0.0, 0.0);
          if (C::STASH && m > 0) {
            u32 v[4];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                         : "r"(tm_stash + 4 * p)
                         : "memory");
            tmem_wait_ld();
            o = make_double2(__hiloint2double((int)v[1], (int)v[0]),
                             __hiloint2double((int)v[3], (int)v[2]));
          }
          if constexpr (C::CL) { // own term, then the partners' (order of the sum is free)
            double2 w;
            asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                         : "=d"(w.x), "=d"(w.y)
                         : "r"(slot + (u32)r * C::POLY_BLOCK_BYTES + (u32)pr * C::ROW_BYTES));
            o.x = fma(x[p].x, w.x, fma(-x[p].y, w.y, o.x));
            o.y = fma(x[p].x, w.y, fma(x[p].y, w.x, o.y));
#pragma unroll
            for (int j = 0; j < NSIB; ++j) {
              asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                           : "=d"(w.x), "=d"(w.y)
                           : "r"(slot + key_cl[j] + (u32)pr * C::ROW_BYTES));
              const double2 f = fs[j][pr];
              o.x = fma(f.x, w.x, fma(-f.y, w.y, o.x));
              o.y = fma(f.x, w.y, fma(f.y, w.x, o.y));
            }
          } else if constexpr (PIPE) {
            if (more_rows) load_row(pr + 1, wn, fn2);
            o.x = fma(x[p].x, wc[0].x, fma(-x[p].y, wc[0].y, o.x));
            o.y = fma(x[p].x, wc[0].y, fma(x[p].y, wc[0].x, o.y));
#pragma unroll
            for (int j = 1; j < GS; ++j) {
              o.x = fma(fc[j].x, wc[j].x, fma(-fc[j].y, wc[j].y, o.x));
              o.y = fma(fc[j].x, wc[j].y, fma(fc[j].y, wc[j].x, o.y));
            }
            if (more_rows) {
#pragma unroll
              for (int j = 0; j < GS; ++j) {
                wc[j] = wn[j];
                fc[j] = fn2[j];
              }
            }
          } else {
#pragma unroll
            for (int rr = 0; rr < GS; ++rr) {
              double2 w, f;
              asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                           : "=d"(w.x), "=d"(w.y)
                           : "r"(slot + (u32)rr * KEY_RR_STRIDE + (u32)pr * C::ROW_BYTES));
              if (rr == r) {
                f = x[p];
              } else {
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                             : "=d"(f.x), "=d"(f.y)
                             : "r"(sib + (u32)rr * C::BUF_BYTES + (u32)(p * T * 16)));
              }
              o.x = fma(f.x, w.x, fma(-f.y, w.y, o.x));
              o.y = fma(f.x, w.y, fma(f.y, w.x, o.y));
            }
          }
          x[p] = o;
        }
        if constexpr (C::CL) {
#pragma unroll
          for (int j = 0; j < NSIB; ++j)
#pragma unroll
            for (int pr = 0; pr < C::RB; ++pr) fs[j][pr] = fn[j][pr];
        }
        block_done<C>(lane, i, m, (u32)b, blk, rp.slot);
        ring_next<C>(rp);
      }
      // every read of the parked transforms done before the buffers are reused; the cluster
      // variant only ARRIVES here and waits behind the register stages of the next transform
      if constexpr (C::CL) cluster_arrive();
      else ct_bar<C>(ct);
      if constexpr (ST_TM) { // states back (unconditionally: every path to the next level redefines them)
#pragma unroll
        for (int c = 0; c < 2 * P / 16; ++c) {
          u32 v[16];
          tmem_ld16(tm_stash + 16 * c, v);
          tmem_wait_ld();
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            st_re[8 * c + e] = (st_t)v[2 * e + 0];
            st_im[8 * c + e] = (st_t)v[2 * e + 1];
          }
        }
      }
      if (C::STASH && m + 1 < level) {
#pragma unroll
        for (int c4 = 0; c4 < P / 4; ++c4) {
          u32 v[16];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            v[4 * e + 0] = (u32)__double2loint(x[4 * c4 + e].x);
            v[4 * e + 1] = (u32)__double2hiint(x[4 * c4 + e].x);
            v[4 * e + 2] = (u32)__double2loint(x[4 * c4 + e].y);
            v[4 * e + 3] = (u32)__double2hiint(x[4 * c4 + e].y);
          }
          tmem_st16(tm_stash + 16 * c4, v);
        }
        tmem_wait_st();
      }
    }
  }

  // ---- phase C: inverse transform, torus rounding, ACC += (own copy in TMEM, home in smem)
  inverse<C, TwT<C>::ON>(x, tw, C::CL);
  {
    const int tid = opaque_tid(), t = tid % T, grp = tid / T, warp = tid >> 5;
    ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(smem + (size_t)grp * C::BUF_BYTES);
    const u32 tm_own = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) +
                       (u32)(warp >> 2) * C::TM_PER_WARP;
    grp_bar<C>(grp); // last exchange fully read before the buffer becomes the accumulator again
#pragma unroll
    for (int c4 = 0; c4 < P / 4; ++c4) {
      u32 v[16];
      tmem_ld16(tm_own + 16 * c4, v);
      tmem_wait_ld();
      u64 re[4], im[4];
      unpack4(v, re, im);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int p = 4 * c4 + e;
        re[e] += from_torus(x[p].x);
        im[e] += from_torus(x[p].y);
        acc[t + p * T] = make_ulonglong2(re[e], im[e]);
      }
      pack4(v, re, im);
      tmem_st16(tm_own + 16 * c4, v);
    }
    tmem_wait_st();
    grp_bar<C>(grp); // ACC_r complete before the rotated reads of the next CMUX
  }
}

template <class C, bool NARROW, int LV>
__global__ void __launch_bounds__(C::THREADS, 1) pbs_kernel_v4(const PbsArgs a) {
  constexpr int P = C::P, T = C::T, M = C::M, N = C::N, GS = C::GS, K = C::K, CPB = C::CPB;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int grp = tid / T, t = tid % T, ct = C::CL ? 0 : grp / GS, r = poly_of<C>(grp);
  ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(smem + (size_t)grp * C::BUF_BYTES);
  Ctrl *ctrl = reinterpret_cast<Ctrl *>(smem + C::OFF_CTRL);
  const u32 n = a.n, level = LV ? (u32)LV : a.level;
  const size_t out_size = (size_t)K * N + 1;
  // scheduling unit: a CTA, or a cluster of k+1 CTAs working on one ciphertext
  const u32 num_units = C::CL ? gridDim.x / GS : gridDim.x;
  const u32 unit = C::CL ? blockIdx.x / GS : blockIdx.x;

  // Sample -> (CTA, round, slot), as in pbs_kernel_w32v3: `full` rounds with CPB ciphertexts in
  // every CTA, then the remainder spread evenly over all CTAs in one last (shorter) round.
  const u32 wave = num_units * CPB;
  const u32 full = a.num_samples / wave;
  const u32 rem = a.num_samples - full * wave;
  const u32 rem_base = rem / num_units, rem_extra = rem % num_units;
  const u32 my_rem = rem_base + (unit < rem_extra ? 1u : 0u);
  const u32 rem_first = full * wave + unit * rem_base + min(unit, rem_extra);
  const u32 rounds = full + (my_rem ? 1u : 0u);
  const u32 blocks_per_cmux = level * C::BLOCKS_PER_LEVEL;
  const u32 total_blocks = rounds * n * blocks_per_cmux;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < C::NB; ++s) {
      mbar_init(&ctrl->full_bar[s], 1);
      ctrl->done_cnt[s] = 0;
    }
    ctrl->n = n;
    ctrl->level = level;
    ctrl->base_log = a.base_log;
    ctrl->total_blocks = total_blocks;
    ctrl->bsk = a.bsk;
    ctrl->tw = a.tw;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&ctrl->tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const u32 tm_own = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) +
                     (u32)(warp >> 2) * C::TM_PER_WARP;

  if constexpr (TwT<C>::ON) { // this thread's twiddles of the non-first passes -> tensor memory
    constexpr int LOGP = C::LOGP, LOGM = C::LOGM;
    const u32 tm_tw = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) + TwT<C>::BASE +
                      (u32)(warp >> 2) * TwT<C>::COLS;
#pragma unroll
    for (int pass = 1; pass < C::NPASS; ++pass) {
      const int lo = pass_lo<C>(pass);
      const int s0 = pass * LOGP, s1 = (pass == C::NPASS - 1) ? LOGM : s0 + LOGP;
      u32 v[TwT<C>::SLOTS * 4];
#pragma unroll
      for (int e = 0; e < TwT<C>::SLOTS * 4; ++e) v[e] = 0;
#pragma unroll
      for (int s = s0; s < s1; ++s) {
        const int rho = LOGM - 1 - s - lo;
#pragma unroll
        for (int kk = 0; kk < tw_slot_count(LOGP, rho); ++kk) {
          const int slot = tw_slot_base(LOGP, rho) + kk;
          const int b = ((t >> lo) << (LOGP - rho - 1)) | (2 * kk);
          const double2 w = a.tw[(1 << s) + b];
          v[4 * slot + 0] = (u32)__double2loint(w.x);
          v[4 * slot + 1] = (u32)__double2hiint(w.x);
          v[4 * slot + 2] = (u32)__double2loint(w.y);
          v[4 * slot + 3] = (u32)__double2hiint(w.y);
        }
      }
#pragma unroll
      for (int c = 0; c < TwT<C>::SLOTS * 4 / 16; ++c) {
        u32 v16[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) v16[e] = v[16 * c + e];
        tmem_st16(tm_tw + (u32)(pass - 1) * TwT<C>::PASS_COLS + 16 * c, v16);
      }
      if constexpr (TwT<C>::SLOTS * 4 % 16 != 0) { // P = 8: 16 columns per pass, handled above
        static_assert(TwT<C>::SLOTS * 4 % 16 == 0, "twiddle columns per pass");
      }
    }
    tmem_wait_st();
  }

  if (tid == 0) { // prologue: the first NB blocks
    for (u32 f = 0; f < (u32)C::NB && f < total_blocks; ++f)
      issue_block<C>((f / C::BLOCKS_PER_LEVEL) / level, (f / C::BLOCKS_PER_LEVEL) % level,
                     f % C::BLOCKS_PER_LEVEL, f);
  }

  u32 blk0 = 0; // block number of (current CMUX, level 0, block 0)
  for (u32 rd = 0; rd < rounds; ++rd) {
    const bool active = rd < full || (u32)ct < my_rem;
    const u32 sample = rd < full ? rd * wave + unit * CPB + ct : rem_first + ct;
    const u64 *lwe = a.lwe_in + (active ? a.in_idx[sample] : 0) * (u64)(n + 1);

    if (active) { // ACC_r <- LUT_r * X^(-b_hat)   (polynomial_algorithms.rs:500-539)
      const u32 b_hat = modulus_switch(__ldg(&lwe[n]), C::LOGN + 1);
      const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(GS * N) + (size_t)r * N;
#pragma unroll
      for (int c4 = 0; c4 < P / 4; ++c4) {
        u64 re[4], im[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = t + (4 * c4 + e) * T;
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
        pack4(v, re, im);
        tmem_st16(tm_own + 16 * c4, v);
      }
      tmem_wait_st();
    }
    grp_bar<C>(grp);

    u32 a_next = active ? modulus_switch(__ldg(&lwe[0]), C::LOGN + 1) : 0;
#pragma unroll 1
    for (u32 i = 0; i < n; ++i, blk0 += blocks_per_cmux) {
      const u32 a_hat = a_next;
      if (active && i + 1 < n) a_next = modulus_switch(__ldg(&lwe[i + 1]), C::LOGN + 1);
      if (a_hat == 0) // X^0 - 1 = 0: no CMUX (bootstrap.rs:287), or an idle slot
        idle_cmux<C>(i, blk0);
      else
        cmux<C, NARROW, LV>(a_hat, i, blk0);
    }

    // sample extraction (glwe_sample_extraction.rs:90-165) at nth = li*lut_stride: group c < k
    // writes mask polynomial c from its own accumulator, group k the body
    if (active) {
      for (u32 li = 0; li < a.lut_count; ++li) {
        const u32 nth = li * a.lut_stride;
        u64 *out = a.lwe_out + ((size_t)li * a.num_samples + a.out_idx[sample]) * out_size;
        if (r < K) {
          for (int j = t; j < N; j += T) {
            const int src = (j <= (int)nth) ? ((int)nth - j) : (N + (int)nth - j);
            const ulonglong2 pr = acc[src & (M - 1)];
            const u64 v = (src < M) ? pr.x : pr.y;
            out[(size_t)r * N + j] = (j <= (int)nth) ? v : (u64)0 - v;
          }
        } else if (t == 0) {
          const ulonglong2 pr = acc[nth & (M - 1)];
          out[(size_t)K * N] = (nth < (u32)M) ? pr.x : pr.y;
        }
      }
    }
    grp_bar<C>(grp);
  }

  if constexpr (C::CL) { // no CTA leaves while a partner may still read its shared memory
    cluster_arrive();
    cluster_wait();
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(ctrl->tmem_base),
                 "r"(512u)
                 : "memory");
}

// ---- K4 for this layout: dst[p*T + t] = F_{t*P + p} / M, in place, one group per polynomial ----
template <class C>
__global__ void __launch_bounds__(C::T) bsk_convert_kernel_v4(double2 *bsk,
                                                              const double2 *__restrict__ tw,
                                                              size_t num_polys) {
  constexpr int P = C::P, T = C::T, M = C::M;
  const int t = threadIdx.x;
  for (size_t poly = blockIdx.x; poly < num_polys; poly += gridDim.x) {
    const u64 *src = reinterpret_cast<const u64 *>(bsk + poly * M);
    double2 x[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const int j = t + p * T;
      // convert_forward_torus (fft/mod.rs:201-222): signed value * 2^-64
      x[p].x = __ll2double_rn((i64)src[j]) * 5.421010862427522170037e-20;
      x[p].y = __ll2double_rn((i64)src[j + M]) * 5.421010862427522170037e-20;
    }
    __syncthreads(); // every read of this polynomial precedes the in-place writes
    // (the CTA is one group: tid / T == 0, exchange buffer at offset 0, barrier of T threads)
    forward<C>(x, tw);
    double2 *dst = bsk + poly * M;
    const double inv_m = 1.0 / (double)M;
#pragma unroll
    for (int p = 0; p < P; ++p) dst[p * T + t] = make_double2(x[p].x * inv_m, x[p].y * inv_m);
    __syncthreads();
  }
}

} // namespace v4
