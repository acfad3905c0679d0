// fft.cuh — K2: negacyclic f64 transform for polynomials mod X^N+1, N = 2M.
//
// Computes what tfhe/src/core_crypto/fft_impl/fft64/math/fft/mod.rs:201-308,491-561 computes
// (fold z_j = p_j + i p_{j+M}, twist by exp(i*pi*j/N), M-point complex FFT, "unordered" output)
// and what the reference GPU does in cuda/src/fft/bnsmfft.cuh:26-202, with a different schedule:
//
//  * The twist is merged into the butterflies.  The M outputs are Z(psi^(4q'+1)), the roots of
//    X^M - i (psi = exp(i*pi/N)); the transform is the recursive reduction
//       A mod (X^2m - c^2)  ->  (A_lo + c*A_hi) mod (X^m - c),  (A_lo - c*A_hi) mod (X^m + c)
//    i.e. only Cooley-Tukey butterflies (6 FMA-pipe ops each: 4 FMA + 2 FMA), no separate twist
//    pass and no trivial stages.  Stage s (0-based), block b uses
//       tw[2^s + b] = exp(i*pi*(1 + 4*bitrev_s(b)) / 2^(s+2)).
//  * P = 2^LOGP points per thread live in registers; log2(P) stages are done per pass without
//    any synchronisation, and passes exchange data through a padded shared-memory buffer
//    (one 16-byte complex per bank group, conflict-free for every pass pattern).
//  * The inverse is the exact mirror (Gentleman-Sande butterflies with conjugate twiddles), left
//    unnormalised: the 1/M factor is folded into the Fourier bootstrapping key at conversion.
//  * The thread that holds natural-order points j = t + r*T before the forward transform holds
//    Fourier positions q = t*P + r after it; pointwise products happen there, in registers, and
//    the inverse returns to j = t + r*T.  The Fourier key is stored in exactly that order, so no
//    bit-reversal or transposition is ever materialised.
#pragma once
#include <cuda_runtime.h>

namespace cbfft {

// tw[2^s + b] does not depend on M, so one small table serves every transform size: the first
// pass of every transform pairs the top index bits, its block index b is a compile-time constant
// (t >> lo == 0) and the twiddle becomes a constant-bank operand of the FMA (no load at all).
constexpr int CONST_TW = 256; // stages s < 8
__constant__ double2 c_twg[CONST_TW];

template <int LOGM, int LOGP> struct Cfg {
  static constexpr int M = 1 << LOGM;
  static constexpr int P = 1 << LOGP;
  static constexpr int T = M / P; // threads per transform
  static constexpr int NPASS = (LOGM + LOGP - 1) / LOGP;
  static constexpr int LAST_G = LOGM - LOGP * (NPASS - 1); // stages in the last forward pass
  static constexpr int BUF_ELEMS = M + (M >> 3);           // padded exchange buffer (double2)
};

__device__ __forceinline__ int pad(int idx) { return idx + (idx >> 3); }

// lowest index bit held in-thread during forward pass `pass`
template <int LOGM, int LOGP> __device__ __forceinline__ constexpr int pass_lo(int pass) {
  using C = Cfg<LOGM, LOGP>;
  return (pass == C::NPASS - 1) ? 0 : (LOGM - LOGP * (pass + 1));
}

// element index of in-thread slot r for thread t when bits [lo, lo+LOGP) are in-thread
template <int LOGP> __device__ __forceinline__ int elem_index(int t, int r, int lo) {
  return ((t >> lo) << (lo + LOGP)) | (r << lo) | (t & ((1 << lo) - 1));
}

// CT butterfly: (a, b) <- (a + w*b, a - w*b)      6 FMA-pipe ops
__device__ __forceinline__ void bfly_ct(double2 &a, double2 &b, const double2 w) {
  const double tr = fma(-w.y, b.y, fma(w.x, b.x, a.x));
  const double ti = fma(w.y, b.x, fma(w.x, b.y, a.y));
  b.x = fma(2.0, a.x, -tr);
  b.y = fma(2.0, a.y, -ti);
  a.x = tr;
  a.y = ti;
}

// GS butterfly with conjugate twiddle: (u, v) <- (u + v, (u - v)*conj(w))   8 ops
__device__ __forceinline__ void bfly_gs(double2 &u, double2 &v, const double2 w) {
  const double dr = u.x - v.x, di = u.y - v.y;
  u.x += v.x;
  u.y += v.y;
  v.x = fma(w.y, di, w.x * dr);
  v.y = fma(-w.y, dr, w.x * di);
}

// Stages [s_begin, s_end) of the forward transform on in-thread data (bits [lo, lo+LOGP) held).
template <int LOGM, int LOGP>
__device__ __forceinline__ void fwd_stages(double2 (&x)[1 << LOGP], int t, int lo, int s_begin,
                                           int s_end, const double2 *__restrict__ tw) {
  constexpr int P = 1 << LOGP;
#pragma unroll
  for (int s = s_begin; s < s_end; ++s) {
    const int beta = LOGM - 1 - s; // index bit paired by this stage
    const int rho = beta - lo;     // in-thread bit
    // in-thread slots r and r | 2^(rho+1) sit in blocks b (even) and b + 1, and
    // tw[2^s + b + 1] = i * tw[2^s + b]: one load serves both (the twiddle loads are a third of
    // the L1 traffic of the fused PBS kernels, profiles/r1_grp_p4_ncu.txt)
#pragma unroll
    for (int r = 0; r < P; ++r) {
      if (r & (1 << rho)) continue;
      if (rho + 1 < LOGP && (r & (1 << (rho + 1)))) continue;
      const int idx = elem_index<LOGP>(t, r, lo);
      const bool first_pass = (lo + LOGP == LOGM) && ((1 << s) + (r >> (rho + 1)) < CONST_TW);
      const int b = first_pass ? (r >> (rho + 1)) : (idx >> (beta + 1));
      const double2 w = first_pass ? c_twg[(1 << s) + b] : __ldg(&tw[(1 << s) + b]);
      bfly_ct(x[r], x[r | (1 << rho)], w);
      if (rho + 1 < LOGP) {
        const int r2 = r | (1 << (rho + 1));
        bfly_ct(x[r2], x[r2 | (1 << rho)], make_double2(-w.y, w.x));
      }
    }
  }
}

template <int LOGM, int LOGP>
__device__ __forceinline__ void inv_stages(double2 (&x)[1 << LOGP], int t, int lo, int s_hi,
                                           int s_lo, const double2 *__restrict__ tw) {
  // stages s_hi-1 down to s_lo
  constexpr int P = 1 << LOGP;
#pragma unroll
  for (int s = s_hi - 1; s >= s_lo; --s) {
    const int beta = LOGM - 1 - s;
    const int rho = beta - lo;
#pragma unroll
    for (int r = 0; r < P; ++r) {
      if (r & (1 << rho)) continue;
      if (rho + 1 < LOGP && (r & (1 << (rho + 1)))) continue;
      const int idx = elem_index<LOGP>(t, r, lo);
      const bool first_pass = (lo + LOGP == LOGM) && ((1 << s) + (r >> (rho + 1)) < CONST_TW);
      const int b = first_pass ? (r >> (rho + 1)) : (idx >> (beta + 1));
      const double2 w = first_pass ? c_twg[(1 << s) + b] : __ldg(&tw[(1 << s) + b]);
      bfly_gs(x[r], x[r | (1 << rho)], w);
      if (rho + 1 < LOGP) {
        const int r2 = r | (1 << (rho + 1));
        bfly_gs(x[r2], x[r2 | (1 << rho)], make_double2(-w.y, w.x));
      }
    }
  }
}

// Barrier used between passes.  GROUP_T == blockDim.x -> __syncthreads, otherwise a named
// barrier per transform group (several transforms per CTA).
template <int GROUP_T> __device__ __forceinline__ void group_sync(int bar_id) {
  if (bar_id < 0) {
    __syncthreads();
  } else {
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(GROUP_T) : "memory");
  }
}

// Forward transform.  In: x[r] = z_{t + r*T}.  Out: x[r] = F_{q = t*P + r}.
// `buf` must hold Cfg::BUF_ELEMS double2.  Each exchange is barrier / write / barrier / read, so
// back-to-back transforms may share one buffer.
template <int LOGM, int LOGP>
__device__ __forceinline__ void forward(double2 (&x)[1 << LOGP], double2 *buf, int t,
                                        const double2 *__restrict__ tw, int bar_id = -1) {
  using C = Cfg<LOGM, LOGP>;
  constexpr int P = C::P;
#pragma unroll
  for (int pass = 0; pass < C::NPASS; ++pass) {
    const int lo = pass_lo<LOGM, LOGP>(pass);
    const int s0 = pass * LOGP;
    const int s1 = (pass == C::NPASS - 1) ? LOGM : s0 + LOGP;
    fwd_stages<LOGM, LOGP>(x, t, lo, s0, s1, tw);
    if (pass + 1 < C::NPASS) {
      const int lo2 = pass_lo<LOGM, LOGP>(pass + 1);
      group_sync<C::T>(bar_id); // every earlier reader of `buf` is done
#pragma unroll
      for (int r = 0; r < P; ++r) buf[pad(elem_index<LOGP>(t, r, lo))] = x[r];
      group_sync<C::T>(bar_id);
#pragma unroll
      for (int r = 0; r < P; ++r) x[r] = buf[pad(elem_index<LOGP>(t, r, lo2))];
    }
  }
}

// Inverse transform (unnormalised: returns M * z).  In: x[r] = F_{t*P + r}.  Out: x[r] = M*z_{t+r*T}.
template <int LOGM, int LOGP>
__device__ __forceinline__ void inverse(double2 (&x)[1 << LOGP], double2 *buf, int t,
                                        const double2 *__restrict__ tw, int bar_id = -1) {
  using C = Cfg<LOGM, LOGP>;
  constexpr int P = C::P;
#pragma unroll
  for (int pass = C::NPASS - 1; pass >= 0; --pass) {
    const int lo = pass_lo<LOGM, LOGP>(pass);
    const int s0 = pass * LOGP;
    const int s1 = (pass == C::NPASS - 1) ? LOGM : s0 + LOGP;
    inv_stages<LOGM, LOGP>(x, t, lo, s1, s0, tw);
    if (pass > 0) {
      const int lo2 = pass_lo<LOGM, LOGP>(pass - 1);
      group_sync<C::T>(bar_id); // every earlier reader of `buf` is done
#pragma unroll
      for (int r = 0; r < P; ++r) buf[pad(elem_index<LOGP>(t, r, lo))] = x[r];
      group_sync<C::T>(bar_id);
#pragma unroll
      for (int r = 0; r < P; ++r) x[r] = buf[pad(elem_index<LOGP>(t, r, lo2))];
    }
  }
}

} // namespace cbfft
