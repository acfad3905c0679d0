// pbs_w32.cuh — "warp per polynomial" building blocks for N = 2048 (M = 1024 complex points):
// the register passes, the 16 KiB transpose, torus rounding, the BSK -> Fourier converter and the
// FFT test hook.  The PBS kernel that uses them is pbs_w32_v3.cuh; its two predecessors (v1: this
// file's original kernel, v2: TMEM + one-stage key buffer) are kept as measured baselines under
// experiments/ and are not compiled into the library.
// Original header of the first kernel follows (the data-flow description still applies to v3):
// "warp per polynomial" fast path of K3/K4 for N = 2048 (M = 1024 complex points),
// k = 1, one decomposition level: the shape of BASELINE.json configs[1] (n=742, k=1, N=2048, l=1)
// and of the reference's own benchmark (benchmark_pbs.cpp:388-403).
//
// Why this shape gets its own kernel: ncu on the generic kernel (profiles/r1_pbs_generic_*.txt)
// shows it is bound by the L1/shared-memory data pipe (128 B/clk/SM), not by FP64: with 8 points
// per thread an M = 1024 transform needs three shared-memory exchanges, the accumulator is read
// twice and written once per CMUX, and register spills go through the same pipe.  Here
//   * one WARP owns one polynomial of one ciphertext: 32 points per thread, the transform is
//     5 register stages | ONE 16 KiB transpose through shared memory | 5 register stages, and
//     the transpose needs only __syncwarp — no CTA barrier inside a transform;
//   * the first 5 forward stages (last 5 inverse stages) use compile-time twiddles taken straight
//     from the constant bank as DFMA operands; the other 31 per-lane twiddles come from a 16 KiB
//     shared-memory table laid out [stage][sub-block][lane] (conflict-free LDS.128);
//   * accumulator polynomial r is touched only by warp r, as (coef j, coef j+M) 16-byte pairs;
//     the new value computed by the torus rounding of iteration i stays in registers and is the
//     "own" operand of ACC*X^a - ACC of iteration i+1, so per CMUX the accumulator costs one
//     rotated read, one read and one write of shared memory;
//   * the two warps of a ciphertext meet once per CMUX (named barrier, 64 threads) to swap the
//     cross terms X_r * G[r][1-r] of the external product;
//   * the Fourier key is scaled by 1/M (not 2^-64/M), so the inverse transform delivers the
//     result already in units of 2^-64 and torus rounding is 3 DADD + 1 F2I.
// Restates the same reference code as pbs.cu (bootstrap.rs:250-368, ggsw.rs:481-602).
#pragma once

namespace w32 {
using namespace sm100;

constexpr int LOGM = 10, M = 1024, N = 2048;
constexpr int CPB = 3;                    // ciphertexts per CTA (2 warps each)
constexpr int NTHREADS = 64 * CPB;
constexpr int ACC_PAIRS = 2 * M;          // two polynomials, M pairs each
constexpr size_t SMEM_TW = 31 * 32 * sizeof(double2);                    // 15 872 B
constexpr size_t SMEM_PER_CT = (size_t)ACC_PAIRS * 16 + 2 * (size_t)M * 16; // 64 KiB
constexpr size_t SMEM_BYTES = SMEM_TW + CPB * SMEM_PER_CT;

// tw[2^s + b], s = 0..4 (31 entries at indices 1..31): compile-time indexed -> c[bank] operands
__constant__ double2 c_tw[32];

__device__ __forceinline__ int swz(int idx, int key) { return idx ^ (key & 7); }

// ---- register passes -------------------------------------------------------------------------
// Pass 1 (forward stages 0..4): thread holds j = lane + 32*p, in-thread bits are j[9:5] = p.
__device__ __forceinline__ void fwd_pass1(double2 (&x)[32]) {
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const int rho = 4 - s;
#pragma unroll
    for (int u = 0; u < (1 << s); ++u) {
      const double2 w = c_tw[(1 << s) + u];
#pragma unroll
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo;
        cbfft::bfly_ct(x[p], x[p | (1 << rho)], w);
      }
    }
  }
}
// Pass 2 (forward stages 5..9): thread holds j = 32*lane + p, in-thread bits are j[4:0] = p.
// tws[((2^(s-5) - 1) + u) * 32 + lane] = tw[2^s + (lane << (s-5)) + u].
__device__ __forceinline__ void fwd_pass2(double2 (&x)[32], const double2 *tws, int lane) {
#pragma unroll
  for (int s2 = 0; s2 < 5; ++s2) {
    const int rho = 4 - s2;
#pragma unroll
    for (int u = 0; u < (1 << s2); ++u) {
      const double2 w = tws[(((1 << s2) - 1) + u) * 32 + lane];
#pragma unroll
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo;
        cbfft::bfly_ct(x[p], x[p | (1 << rho)], w);
      }
    }
  }
}
__device__ __forceinline__ void inv_pass2(double2 (&x)[32], const double2 *tws, int lane) {
#pragma unroll
  for (int s2 = 4; s2 >= 0; --s2) {
    const int rho = 4 - s2;
#pragma unroll
    for (int u = 0; u < (1 << s2); ++u) {
      const double2 w = tws[(((1 << s2) - 1) + u) * 32 + lane];
#pragma unroll
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo;
        cbfft::bfly_gs(x[p], x[p | (1 << rho)], w);
      }
    }
  }
}
__device__ __forceinline__ void inv_pass1(double2 (&x)[32]) {
#pragma unroll
  for (int s = 4; s >= 0; --s) {
    const int rho = 4 - s;
#pragma unroll
    for (int u = 0; u < (1 << s); ++u) {
      const double2 w = c_tw[(1 << s) + u];
#pragma unroll
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo;
        cbfft::bfly_gs(x[p], x[p | (1 << rho)], w);
      }
    }
  }
}

// ---- inverse pass 1 with 6-op butterflies (Linzer-Feig scaling) ----------------------------------
// A Gentleman-Sande butterfly (u, v) <- (u + v, (u - v) * conj(w)) is 8 FP64 instructions.  With
// the twiddle written as c * (1 + i t) (t = tan, or s * (cot + i) when |sin| > |cos|) the real
// factor is NOT multiplied in: every slot carries a pending real scale sigma[p] that is a
// compile-time function of the slot, the next butterfly reads its v operand through
// rho = sigma[v] / sigma[u] (an FMA operand, free), and the last pending scale is the multiplier
// of the FMA that starts the torus rounding.  (u + rho v, u - rho v) are 4 FMA, the rotation by
// (1, t) is 2 FMA: 6 instructions, 160 fewer per thread and CMUX.  Only the constant-twiddle pass
// can do this: its rho / t / sigma are the same in every lane (constant-bank operands); in the
// lane-dependent pass they would be 32 more values per lane and a multiply per point to resolve.
// Tables are built on the host next to c_tw (lf_tables below), butterfly index
// bi = 16 * (4 - s) + (u << rho) + lo in the order inv_pass1 walks them.
__host__ __device__ constexpr int lf_bitrev(int b, int bits) {
  int r = 0;
  for (int i = 0; i < bits; ++i) r |= ((b >> i) & 1) << (bits - 1 - i);
  return r;
}
// angle of tw[2^s + u] is pi * num / den with num = 1 + 4 bitrev_s(u), den = 2^(s+2):
// |cos| >= |sin| iff the angle is within pi/4 of 0 or pi
__host__ __device__ constexpr bool lf_tan_form(int s, int u) {
  const int num = 1 + 4 * lf_bitrev(u, s), den = 1 << (s + 2);
  return 4 * num <= den || 4 * num >= 3 * den;
}
__constant__ double c_lf_rho[80], c_lf_t[80], c_lf_sigma[32];

__device__ __forceinline__ void inv_pass1_lf(double2 (&x)[32]) {
#pragma unroll
  for (int s = 4; s >= 0; --s) {
    const int rho = 4 - s;
#pragma unroll
    for (int u = 0; u < (1 << s); ++u) {
#pragma unroll
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo, q = p | (1 << rho);
        const int bi = 16 * (4 - s) + (u << rho) + lo;
        const double r = c_lf_rho[bi], t = c_lf_t[bi];
        const double dr = fma(-r, x[q].x, x[p].x), di = fma(-r, x[q].y, x[p].y);
        x[p].x = fma(r, x[q].x, x[p].x);
        x[p].y = fma(r, x[q].y, x[p].y);
        if (lf_tan_form(s, u)) { // c * (dr + t di, di - t dr)
          x[q].x = fma(t, di, dr);
          x[q].y = fma(-t, dr, di);
        } else {                 // s * (di + t dr, t di - dr), t = cot
          x[q].x = fma(t, dr, di);
          x[q].y = fma(t, di, -dr);
        }
      }
    }
  }
}

// host: the tables of inv_pass1_lf, from the same angles as the twiddle table (get_twiddles)
inline void lf_tables(double (&rho_t)[80], double (&t_t)[80], double (&sigma_t)[32]) {
  const long double pi = 3.14159265358979323846264338327950288L;
  long double sigma[32];
  for (int p = 0; p < 32; ++p) sigma[p] = 1.0L;
  for (int s = 4; s >= 0; --s) {
    const int rho = 4 - s;
    for (int u = 0; u < (1 << s); ++u) {
      const long double ang = pi * (long double)(1 + 4 * lf_bitrev(u, s)) / (long double)(1 << (s + 2));
      const long double c = cosl(ang), sn = sinl(ang);
      const bool tan_form = lf_tan_form(s, u);
      for (int lo = 0; lo < (1 << rho); ++lo) {
        const int p = (u << (rho + 1)) | lo, q = p | (1 << rho);
        const int bi = 16 * (4 - s) + (u << rho) + lo;
        rho_t[bi] = (double)(sigma[q] / sigma[p]);
        t_t[bi] = (double)(tan_form ? sn / c : c / sn);
        sigma[q] = sigma[p] * (tan_form ? c : sn);
      }
    }
  }
  for (int p = 0; p < 32; ++p) sigma_t[p] = (double)sigma[p];
}

// transpose between the two register layouts through a 16 KiB buffer owned by this warp
__device__ __forceinline__ void xchg_1to2(double2 (&x)[32], double2 *buf, int lane) {
  __syncwarp();
#pragma unroll
  for (int p = 0; p < 32; ++p) buf[swz(lane + 32 * p, p)] = x[p];
  __syncwarp();
#pragma unroll
  for (int p = 0; p < 32; ++p) x[p] = buf[swz(32 * lane + p, lane)];
}
__device__ __forceinline__ void xchg_2to1(double2 (&x)[32], double2 *buf, int lane) {
  __syncwarp();
#pragma unroll
  for (int p = 0; p < 32; ++p) buf[swz(32 * lane + p, lane)] = x[p];
  __syncwarp();
#pragma unroll
  for (int p = 0; p < 32; ++p) x[p] = buf[swz(lane + 32 * p, p)];
}

// In: x[p] = z_{lane + 32p}.  Out: x[p] = F_{q = 32*lane + p}.
__device__ __forceinline__ void forward(double2 (&x)[32], double2 *buf, const double2 *tws,
                                        int lane) {
  fwd_pass1(x);
  xchg_1to2(x, buf, lane);
  fwd_pass2(x, tws, lane);
}
// In: x[p] = F_{32*lane + p}.  Out: x[p] = M * z_{lane + 32p} (unnormalised).
__device__ __forceinline__ void inverse(double2 (&x)[32], double2 *buf, const double2 *tws,
                                        int lane) {
  inv_pass2(x, tws, lane);
  xchg_2to1(x, buf, lane);
  inv_pass1(x);
}

__device__ __forceinline__ void load_tws(double2 *tws, const double2 *__restrict__ tw, int tid,
                                         int nthreads) {
  for (int e = tid; e < 31 * 32; e += nthreads) {
    const int row = e >> 5, lane = e & 31;
    int s2 = 0;
    while (((1 << (s2 + 1)) - 1) <= row) ++s2;
    const int u = row - ((1 << s2) - 1);
    tws[e] = tw[(1 << (s2 + 5)) + (lane << s2) + u];
  }
}

// round(v) mod 2^64 for v already in units of 2^-64 (torus/mod.rs:72-78 with the 2^64 folded in)
__device__ __forceinline__ u64 from_torus_scaled(double v) {
  const double C = 1.5 * 4503599627370496.0 * 18446744073709551616.0; // 1.5 * 2^52 * 2^64
  const double r = (v + C) - C;                                        // nearest multiple of 2^64
  return (u64)__double2ll_rn(v - r);
}

// the same for v = sigma * x with the pending scale of inv_pass1_lf as an FMA operand
__device__ __forceinline__ u64 from_torus_scaled_lf(double x, double sigma) {
  const double C = 1.5 * 4503599627370496.0 * 18446744073709551616.0;
  const double r = fma(sigma, x, C) - C;
  return (u64)__double2ll_rn(fma(sigma, x, -r));
}

// closest_representable + the single balanced digit of a one-level decomposition, from the high
// word only (base_log <= 31, so every bit that matters lives in bits 63..32):
// decomposer.rs:23-47 then iter.rs:121-133 with state >> base_log == 0.
__device__ __forceinline__ int digit_l1(u32 hi, u32 base_log) {
  u32 r = ((hi >> (31 - base_log)) + 1) >> 1;
  r &= (1u << base_log) - 1;
  const u32 carry = ((r - 1) & r) >> (base_log - 1);
  return (int)(r - (carry << base_log));
}

// ---- K4 fast path: one warp per key polynomial, output in the [p][lane] order pbs_kernel_w32
// reads, scaled by 1/M only (see header).
__global__ void __launch_bounds__(128) bsk_convert_kernel_w32(double2 *bsk,
                                                              const double2 *__restrict__ tw,
                                                              size_t num_polys) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *tws = reinterpret_cast<double2 *>(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double2 *buf = reinterpret_cast<double2 *>(smem + SMEM_TW) + warp * M;
  load_tws(tws, tw, tid, blockDim.x);
  __syncthreads();
  const size_t wpb = blockDim.x >> 5;
  for (size_t poly = blockIdx.x * wpb + warp; poly < num_polys; poly += gridDim.x * wpb) {
    const u64 *src = reinterpret_cast<const u64 *>(bsk + poly * M);
    double2 x[32];
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const int j = lane + 32 * p;
      x[p].x = __ll2double_rn((i64)src[j]) * (1.0 / (double)M);
      x[p].y = __ll2double_rn((i64)src[j + M]) * (1.0 / (double)M);
    }
    __syncwarp(); // every read of this polynomial precedes the in-place writes
    forward(x, buf, tws, lane);
    double2 *dst = bsk + poly * M;
#pragma unroll
    for (int p = 0; p < 32; ++p) dst[p * 32 + lane] = x[p];
  }
}

// FFT test hook on the warp-level transform (N = 2048 only)
__global__ void __launch_bounds__(128) poly_mul_kernel_w32(const double *in1, const double *in2,
                                                           double *out,
                                                           const double2 *__restrict__ tw,
                                                           u32 total) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *tws = reinterpret_cast<double2 *>(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double2 *buf = reinterpret_cast<double2 *>(smem + SMEM_TW) + warp * M;
  load_tws(tws, tw, tid, blockDim.x);
  __syncthreads();
  const u32 wpb = blockDim.x >> 5;
  for (u32 poly = blockIdx.x * wpb + warp; poly < total; poly += gridDim.x * wpb) {
    double2 x[32], y[32];
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const int j = lane + 32 * p;
      x[p] = make_double2(in1[(size_t)poly * N + j], in1[(size_t)poly * N + j + M]);
    }
    forward(x, buf, tws, lane);
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const int j = lane + 32 * p;
      y[p] = make_double2(in2[(size_t)poly * N + j], in2[(size_t)poly * N + j + M]);
    }
    forward(y, buf, tws, lane);
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const double2 u = x[p], v = y[p];
      x[p] = make_double2(u.x * v.x - u.y * v.y, u.x * v.y + u.y * v.x);
    }
    inverse(x, buf, tws, lane);
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const int j = lane + 32 * p;
      out[(size_t)poly * N + j] = x[p].x * (1.0 / (double)M);
      out[(size_t)poly * N + j + M] = x[p].y * (1.0 / (double)M);
    }
  }
}

} // namespace w32
