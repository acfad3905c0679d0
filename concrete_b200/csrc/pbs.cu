// pbs.cu — K3 (fused programmable bootstrap) + K4 (BSK -> Fourier) + scratch/cleanup.
//
// Replaces, behind the same C ABI, the reference's classic PBS:
//   third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda/src/pbs/programmable_bootstrap_classic.cuh
//   (DEFAULT: 2*n launches, accumulators round-trip through global memory :434-448),
//   programmable_bootstrap_cg_classic.cuh, programmable_bootstrap_tbc_classic.cuh,
//   programmable_bootstrap_amortized.cuh:48-213 and pbs/bootstrapping_key.cuh:75-209,
// and computes what concrete-cpu computes:
//   tfhe/src/core_crypto/fft_impl/fft64/crypto/bootstrap.rs:250-368  (blind rotate + extract)
//   tfhe/src/core_crypto/fft_impl/fft64/crypto/ggsw.rs:481-602       (external product)
//   tfhe/src/core_crypto/algorithms/glwe_sample_extraction.rs:90-165
//
// K3 design: ONE launch for the whole batch.  One CTA owns one ciphertext for all n CMUX
// iterations: the GLWE accumulator stays in shared memory as (coef j, coef j+N/2) pairs, every
// thread keeps P Fourier points of all k+1 output polynomials in registers, the Fourier key is
// read straight from L2 with 16-byte coalesced loads in the very order the registers need it,
// and mod-switch / rotation / gadget decomposition / FFT / product / inverse FFT / torus rounding
// / sample extraction never leave the SM.
#include "pbs_common.cuh"
#include <atomic>
#include <cooperative_groups.h>

// pbs_v4.cu
bool cb_pbs_v4_shape(u32 k, u32 N);
int cb_pbs_v4_logp(u32 k, u32 N);
void cb_pbs_v4_convert(cudaStream_t stream, int gpu, void *dest, size_t polys, u32 k, u32 N);
void cb_pbs_v4_launch(cudaStream_t stream, int gpu, u64 *lwe_out, const u64 *out_idx,
                      const u64 *lut_vector, const u64 *lut_idx, const u64 *lwe_in,
                      const u64 *in_idx, const double2 *bsk, u32 n, u32 k, u32 N, u32 base_log,
                      u32 level, u32 num_samples, u32 lut_count, u32 lut_stride);

namespace {

// ------------------------------------------------------------------------------------------
// Kernel configuration per (K, LOGN)
// ------------------------------------------------------------------------------------------
template <int K, int LOGN> struct PbsCfg {
  static constexpr int LOGM = LOGN - 1;
  // points per thread: keep (K+1)*P*2 doubles of Fourier accumulators in registers
  // N = 16384: 16 points per thread (T = 512, 128 registers) for the cluster kernel of pbs_grp.cuh
  static constexpr int LOGP_WANT = (K == 1 && LOGN == 14) ? 4 : ((K <= 2) ? 3 : (K <= 5 ? 2 : 1));
  static constexpr int LOGP_MIN = (LOGM - LOGP_WANT > 10) ? (LOGM - 10) : LOGP_WANT; // T <= 1024
  static constexpr int LOGP = (LOGM - LOGP_MIN < 5) ? (LOGM - 5) : LOGP_MIN;         // T >= 32
  using F = cbfft::Cfg<LOGM, LOGP>;
  static constexpr int T = F::T;
  static constexpr int M = F::M;
  static constexpr int N = 2 * M;
  static constexpr int P = F::P;
  static constexpr int REGS = (K == 1) ? (LOGN == 12 ? 255 : 128) : ((K == 2) ? 255 : 128);
  static constexpr int MINB_R = (65536 / REGS) / T;
  static constexpr int MINB = MINB_R < 1 ? 1 : (MINB_R > 16 ? 16 : MINB_R);
  static constexpr size_t FBUF_BYTES = (size_t)F::BUF_ELEMS * sizeof(double2);
  static constexpr size_t ACC_BYTES = (size_t)(K + 1) * N * sizeof(u64);
  // accumulator in shared memory if it fits in the 227 KB opt-in limit (with ~4 KB for a_hat)
  static constexpr bool ACC_GLOBAL = (FBUF_BYTES + ACC_BYTES + 4096) > 232448;
};


} // namespace
#include "sm100_prims.cuh"
#include "pbs_w32.cuh"
#include "pbs_w32_v3.cuh"
namespace {

void cb_w32_set_const_twiddles(const double2 *host_tw) {
  CB_CHECK(cudaMemcpyToSymbol(w32::c_tw, host_tw, 32 * sizeof(double2)));
  double rho[80], t[80], sigma[32];
  w32::lf_tables(rho, t, sigma);
  CB_CHECK(cudaMemcpyToSymbol(w32::c_lf_rho, rho, sizeof(rho)));
  CB_CHECK(cudaMemcpyToSymbol(w32::c_lf_t, t, sizeof(t)));
  CB_CHECK(cudaMemcpyToSymbol(w32::c_lf_sigma, sigma, sizeof(sigma)));
}
const bool g_tw_hook_installed = (g_tw_hook_logm10 = cb_w32_set_const_twiddles, true);

// ------------------------------------------------------------------------------------------
// Kernel plan.  Which fused kernel serves a shape is a pure function of (k, N, level) — it has to
// be: the Fourier key layout is private to the kernel family and is chosen when the key is
// converted, long before the first PBS call.  Families:
//   W32      pbs_w32_v3.cuh   k = 1, N = 2048, l = 1 (BASELINE.json configs[1])
//   V4       pbs_v4.cuh       k = 1, N = 1024..8192 and k = 2, N = 1024, any l
//   GENERIC  pbs_grp.cuh / pbs_kernel: every other (k, N) of the reference, and N = 16384
// concrete_b200_set_pbs_plan() is a TEST hook (tests keep the older kernels covered): it must be
// called before the key is converted.  0 = automatic.
// ------------------------------------------------------------------------------------------
enum { PLAN_AUTO = 0, PLAN_GENERIC = 1, PLAN_GRP = 2, PLAN_CLUSTER = 3, PLAN_V4 = 4 };
std::atomic<int> g_pbs_plan{PLAN_AUTO};
enum Family { FAM_W32, FAM_V4, FAM_GENERIC };

// the warp-per-polynomial kernel covers exactly this shape
inline bool use_w32(u32 k, u32 N, u32 level) { return k == 1 && N == 2048 && level == 1; }

template <int K, int LOGN>
__global__ void __launch_bounds__(PbsCfg<K, LOGN>::T, PbsCfg<K, LOGN>::MINB)
pbs_kernel(const PbsArgs a) {
  using C = PbsCfg<K, LOGN>;
  constexpr int LOGM = C::LOGM, LOGP = C::LOGP, M = C::M, N = C::N, P = C::P, T = C::T;
  constexpr int GS = K + 1;
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *fbuf = reinterpret_cast<double2 *>(smem);
  ulonglong2 *acc;
  unsigned short *s_ahat;
  if (C::ACC_GLOBAL) {
    acc = a.global_acc + (size_t)blockIdx.x * GS * M;
    s_ahat = reinterpret_cast<unsigned short *>(smem + C::FBUF_BYTES);
  } else {
    acc = reinterpret_cast<ulonglong2 *>(smem + C::FBUF_BYTES);
    s_ahat = reinterpret_cast<unsigned short *>(smem + C::FBUF_BYTES + C::ACC_BYTES);
  }
  const int t = threadIdx.x;
  const u32 n = a.n, base_log = a.base_log, level = a.level;
  const u64 mask = ((u64)1 << base_log) - 1;
  const u32 round_shift = 64 - base_log * level - 1; // closest_representable shift
  const u32 state_shift = 64 - base_log * level;
  const double2 *__restrict__ tw = a.tw;
  const size_t out_size = (size_t)K * N + 1;

  for (u32 sample = blockIdx.x; sample < a.num_samples; sample += gridDim.x) {
    const u64 *lwe = a.lwe_in + a.in_idx[sample] * (u64)(n + 1);
    for (u32 i = t; i < n; i += T) s_ahat[i] = (unsigned short)modulus_switch(lwe[i], LOGN + 1);
    const u32 b_hat = modulus_switch(lwe[n], LOGN + 1);

    // ACC <- LUT * X^(-b_hat)   (polynomial_algorithms.rs:500-539)
    const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(GS * N);
#pragma unroll
    for (int c = 0; c < GS; ++c) {
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        ulonglong2 v;
        {
          const u32 u = (j + b_hat) & (2 * N - 1);
          const u64 w = lut[c * N + (u & (N - 1))];
          v.x = (u < (u32)N) ? w : (u64)0 - w;
        }
        {
          const u32 u = (j + M + b_hat) & (2 * N - 1);
          const u64 w = lut[c * N + (u & (N - 1))];
          v.y = (u < (u32)N) ? w : (u64)0 - w;
        }
        acc[c * M + j] = v;
      }
    }
    __syncthreads();

    for (u32 i = 0; i < n; ++i) {
      const u32 a_hat = s_ahat[i];
      if (a_hat == 0) continue; // X^0 - 1 = 0: the CMUX adds nothing (bootstrap.rs:287)

      double2 oacc[GS][P];
#pragma unroll
      for (int c = 0; c < GS; ++c)
#pragma unroll
        for (int p = 0; p < P; ++p) oacc[c][p] = make_double2(0.0, 0.0);

#pragma unroll 1
      for (int r = 0; r < GS; ++r) {
        // ct1_r = ACC_r * X^a_hat - ACC_r  (polynomial_algorithms.rs:610-690), then
        // closest_representable and the decomposition state (fft64/math/decomposition.rs:27-46)
        u64 st_re[P], st_im[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
          const int j = t + p * T;
          const u32 s = (u32)(j - (int)a_hat) & (2 * N - 1);
          const u32 q = s >> LOGM;
          const ulonglong2 src = acc[r * M + (s & (M - 1))];
          const ulonglong2 own = acc[r * M + j];
          const u64 ra = (q & 1) ? src.y : src.x;
          const u64 rb = (q & 1) ? src.x : src.y;
          // (-i)^q * (A + iB): q=0 (A,B) q=1 (B,-A) q=2 (-A,-B) q=3 (-B,A)
          const u64 rot_re = (q == 2 || q == 3) ? (u64)0 - ra : ra;
          const u64 rot_im = (q == 1 || q == 2) ? (u64)0 - rb : rb;
          st_re[p] = closest_representable(rot_re - own.x, round_shift) >> state_shift;
          st_im[p] = closest_representable(rot_im - own.y, round_shift) >> state_shift;
        }
#pragma unroll 1
        for (u32 m = 0; m < level; ++m) { // memory level index m <-> level l-m (ggsw.rs:231-245)
          double2 x[P];
#pragma unroll
          for (int p = 0; p < P; ++p) {
            x[p].x = __int2double_rn((int)(u32)decompose_one_level(st_re[p], mask, base_log));
            x[p].y = __int2double_rn((int)(u32)decompose_one_level(st_im[p], mask, base_log));
          }
          cbfft::forward<LOGM, LOGP>(x, fbuf, t, tw);
          const double2 *__restrict__ g =
              a.bsk + ((((size_t)i * level + m) * GS + r) * GS) * M + t;
#pragma unroll
          for (int c = 0; c < GS; ++c) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
              const double2 w = __ldg(&g[(size_t)c * M + p * T]);
              oacc[c][p].x = fma(x[p].x, w.x, fma(-x[p].y, w.y, oacc[c][p].x));
              oacc[c][p].y = fma(x[p].x, w.y, fma(x[p].y, w.x, oacc[c][p].y));
            }
          }
        }
      }
      if (C::F::NPASS == 1) __syncthreads(); // all rotated reads done before ACC changes
#pragma unroll
      for (int c = 0; c < GS; ++c) {
        cbfft::inverse<LOGM, LOGP>(oacc[c], fbuf, t, tw);
#pragma unroll
        for (int p = 0; p < P; ++p) {
          const int j = t + p * T;
          ulonglong2 v = acc[c * M + j];
          v.x += from_torus(oacc[c][p].x);
          v.y += from_torus(oacc[c][p].y);
          acc[c * M + j] = v;
        }
      }
      __syncthreads();
    }

    // sample extraction (glwe_sample_extraction.rs:90-165) at nth = li*lut_stride
    for (u32 li = 0; li < a.lut_count; ++li) {
      const u32 nth = li * a.lut_stride;
      u64 *out = a.lwe_out + ((size_t)li * a.num_samples + a.out_idx[sample]) * out_size;
      for (int x = t; x < K * N; x += T) {
        const int c = x >> LOGN, j = x & (N - 1);
        const int src = (j <= (int)nth) ? ((int)nth - j) : (N + (int)nth - j);
        const ulonglong2 pr = acc[c * M + (src & (M - 1))];
        const u64 v = (src < M) ? pr.x : pr.y;
        out[x] = (j <= (int)nth) ? v : (u64)0 - v;
      }
      if (t == 0) {
        const ulonglong2 pr = acc[K * M + (nth & (M - 1))];
        out[(size_t)K * N] = (nth < (u32)M) ? pr.x : pr.y;
      }
    }
    __syncthreads();
  }
}

#include "pbs_grp.cuh"

// ------------------------------------------------------------------------------------------
// K4: standard-domain BSK -> Fourier, in place, one CTA per polynomial.
// Output polynomial `poly` occupies the same N*8 bytes as the input polynomial `poly`:
//   dst[p*T + t] = F_{q = t*P + p} / M      (1/M = the inverse transform's normalisation)
// ------------------------------------------------------------------------------------------
template <int K, int LOGN>
__global__ void __launch_bounds__(PbsCfg<K, LOGN>::T)
bsk_convert_kernel(double2 *bsk, const double2 *__restrict__ tw, size_t num_polys) {
  using C = PbsCfg<K, LOGN>;
  constexpr int LOGM = C::LOGM, LOGP = C::LOGP, M = C::M, P = C::P, T = C::T;
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *fbuf = reinterpret_cast<double2 *>(smem);
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
    cbfft::forward<LOGM, LOGP>(x, fbuf, t, tw);
    double2 *dst = bsk + poly * M;
    const double inv_m = 1.0 / (double)M;
#pragma unroll
    for (int p = 0; p < P; ++p) dst[p * T + t] = make_double2(x[p].x * inv_m, x[p].y * inv_m);
    __syncthreads();
  }
}

// FFT test hook: out = in1 (*) in2 negacyclic, doubles.
template <int K, int LOGN>
__global__ void __launch_bounds__(PbsCfg<K, LOGN>::T)
poly_mul_kernel(const double *in1, const double *in2, double *out,
                const double2 *__restrict__ tw, u32 total) {
  using C = PbsCfg<K, LOGN>;
  constexpr int LOGM = C::LOGM, LOGP = C::LOGP, M = C::M, N = C::N, P = C::P, T = C::T;
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *fbuf = reinterpret_cast<double2 *>(smem);
  const int t = threadIdx.x;
  for (u32 poly = blockIdx.x; poly < total; poly += gridDim.x) {
    double2 x[P], y[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const int j = t + p * T;
      x[p] = make_double2(in1[(size_t)poly * N + j], in1[(size_t)poly * N + j + M]);
      y[p] = make_double2(in2[(size_t)poly * N + j], in2[(size_t)poly * N + j + M]);
    }
    cbfft::forward<LOGM, LOGP>(x, fbuf, t, tw);
    cbfft::forward<LOGM, LOGP>(y, fbuf, t, tw);
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const double2 u = x[p], v = y[p];
      x[p] = make_double2(u.x * v.x - u.y * v.y, u.x * v.y + u.y * v.x);
    }
    cbfft::inverse<LOGM, LOGP>(x, fbuf, t, tw);
    const double inv_m = 1.0 / (double)M;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const int j = t + p * T;
      out[(size_t)poly * N + j] = x[p].x * inv_m;
      out[(size_t)poly * N + j + M] = x[p].y * inv_m;
    }
    __syncthreads();
  }
}

// CMUX of the vertical-packing tree (WoP-PBS, wop_pbs.cu): out[p] = t0 + GGSW (x) (t1 - t0) with
// t0 = in[2p], t1 = in[2p + 1] (GLWEs of (K+1)N words) — `cmux` of ggsw.rs applied to a whole tree
// layer (fft64/crypto/wop_pbs/mod.rs:465-587).  One CTA per pair; the GGSW is one Fourier GGSW
// in the generic kernel's layout (bsk_convert_kernel), shared by every pair of the layer; pairs
// of lut `q` use ggsw + q * 0 (same selector bit for every lut).
template <int K, int LOGN>
__global__ void __launch_bounds__(PbsCfg<K, LOGN>::T, PbsCfg<K, LOGN>::MINB)
cmux_kernel(u64 *__restrict__ out, const u64 *__restrict__ in, const double2 *__restrict__ ggsw,
            const double2 *__restrict__ tw, u32 base_log, u32 level, u32 pairs) {
  using C = PbsCfg<K, LOGN>;
  constexpr int LOGM = C::LOGM, LOGP = C::LOGP, M = C::M, N = C::N, P = C::P, T = C::T;
  constexpr int GS = K + 1;
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *fbuf = reinterpret_cast<double2 *>(smem);
  const int t = threadIdx.x;
  const u64 mask = ((u64)1 << base_log) - 1;
  const u32 round_shift = 64 - base_log * level - 1, state_shift = 64 - base_log * level;
  for (u32 pair = blockIdx.x; pair < pairs; pair += gridDim.x) {
    const u64 *t0 = in + (size_t)(2 * pair) * GS * N, *t1 = t0 + (size_t)GS * N;
    double2 oacc[GS][P];
#pragma unroll
    for (int c = 0; c < GS; ++c)
#pragma unroll
      for (int p = 0; p < P; ++p) oacc[c][p] = make_double2(0.0, 0.0);
#pragma unroll 1
    for (int r = 0; r < GS; ++r) {
      u64 st_re[P], st_im[P];
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        st_re[p] = closest_representable(t1[r * N + j] - t0[r * N + j], round_shift) >> state_shift;
        st_im[p] = closest_representable(t1[r * N + j + M] - t0[r * N + j + M], round_shift) >> state_shift;
      }
#pragma unroll 1
      for (u32 m = 0; m < level; ++m) {
        double2 x[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
          x[p].x = __int2double_rn((int)(u32)decompose_one_level(st_re[p], mask, base_log));
          x[p].y = __int2double_rn((int)(u32)decompose_one_level(st_im[p], mask, base_log));
        }
        cbfft::forward<LOGM, LOGP>(x, fbuf, t, tw);
        const double2 *__restrict__ g = ggsw + (((size_t)m * GS + r) * GS) * M + t;
#pragma unroll
        for (int c = 0; c < GS; ++c) {
#pragma unroll
          for (int p = 0; p < P; ++p) {
            const double2 w = __ldg(&g[(size_t)c * M + p * T]);
            oacc[c][p].x = fma(x[p].x, w.x, fma(-x[p].y, w.y, oacc[c][p].x));
            oacc[c][p].y = fma(x[p].x, w.y, fma(x[p].y, w.x, oacc[c][p].y));
          }
        }
      }
    }
    u64 *o = out + (size_t)pair * GS * N;
#pragma unroll
    for (int c = 0; c < GS; ++c) {
      cbfft::inverse<LOGM, LOGP>(oacc[c], fbuf, t, tw);
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        o[c * N + j] = t0[c * N + j] + from_torus(oacc[c][p].x);
        o[c * N + j + M] = t0[c * N + j + M] + from_torus(oacc[c][p].y);
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Host-side dispatch
// ------------------------------------------------------------------------------------------
struct Variant {
  void (*pbs)(const PbsArgs);
  void (*pbs_grp)(const PbsArgs); // group-per-polynomial kernel (pbs_grp.cuh), if it fits
  void (*pbs_grp_w)(const PbsArgs);    // same with 64-bit decomposition states (b*l > 31)
  void (*pbs_grp_fold)(const PbsArgs); // same two reading the v4 key layout (k = 2, N = 1024 only)
  void (*pbs_grp_fold_w)(const PbsArgs);
  void (*pbs_grp_cl_w)(const PbsArgs);
  int grp_threads;
  size_t grp_smem_fixed;
  bool grp_fits;
  void (*pbs_grp_cl)(const PbsArgs); // cluster variant: one CTA per polynomial
  size_t cl_smem_fixed, cl_stash_bytes;
  bool cl_fits;
  void (*convert)(double2 *, const double2 *, size_t);
  void (*polymul)(const double *, const double *, double *, const double2 *, u32);
  void (*cmux)(u64 *, const u64 *, const double2 *, const double2 *, u32, u32, u32);
  int T;
  size_t fbuf_bytes, acc_bytes;
  bool acc_global;
};

template <int K, int LOGN> Variant make_variant() {
  using C = PbsCfg<K, LOGN>;
  Variant v;
  v.pbs = pbs_kernel<K, LOGN>;
  v.pbs_grp = pbs_kernel_grp<K, LOGN, true>;
  v.pbs_grp_w = pbs_kernel_grp<K, LOGN, false>;
  v.pbs_grp_fold = v.pbs_grp_fold_w = nullptr;
  if constexpr (K == 2 && LOGN == 10) { // same key as the v4 kernel of this shape (16 points/thread)
    static_assert(PbsCfg<K, LOGN>::LOGP == 3, "fold 1 = 8 -> 16 points per thread");
    v.pbs_grp_fold = pbs_kernel_grp<K, LOGN, true, 1>;
    v.pbs_grp_fold_w = pbs_kernel_grp<K, LOGN, false, 1>;
  }
  v.pbs_grp_cl_w = pbs_kernel_grp_cluster<K, LOGN, false>;
  v.grp_threads = GrpCfg<K, LOGN>::THREADS;
  v.grp_smem_fixed = GrpCfg<K, LOGN>::SMEM_FIXED;
  v.grp_fits = GrpCfg<K, LOGN>::FITS;
  v.pbs_grp_cl = pbs_kernel_grp_cluster<K, LOGN, true>;
  v.cl_smem_fixed = GrpCfg<K, LOGN>::CL_SMEM_FIXED;
  v.cl_stash_bytes = GrpCfg<K, LOGN>::CL_STASH_BYTES;
  v.cl_fits = GrpCfg<K, LOGN>::CL_FITS;
  v.convert = bsk_convert_kernel<K, LOGN>;
  v.polymul = poly_mul_kernel<K, LOGN>;
  v.cmux = cmux_kernel<K, LOGN>;
  v.T = C::T;
  v.fbuf_bytes = C::FBUF_BYTES;
  v.acc_bytes = C::ACC_BYTES;
  v.acc_global = C::ACC_GLOBAL;
  return v;
}

bool get_variant(u32 k, u32 N, Variant &v) {
  int logn = 0;
  while ((1u << logn) < N) ++logn;
  if ((1u << logn) != N) return false;
#define CB_CASE(KK, LL)                                                                  \
  if (k == KK && logn == LL) {                                                           \
    v = make_variant<KK, LL>();                                                          \
    return true;                                                                         \
  }
  CB_CASE(1, 8) CB_CASE(1, 9) CB_CASE(1, 10) CB_CASE(1, 11) CB_CASE(1, 12) CB_CASE(1, 13)
  CB_CASE(1, 14)
  CB_CASE(2, 8) CB_CASE(2, 9) CB_CASE(2, 10) CB_CASE(2, 11) CB_CASE(2, 12)
  CB_CASE(3, 8) CB_CASE(3, 9) CB_CASE(3, 10) CB_CASE(3, 11)
  CB_CASE(4, 8) CB_CASE(4, 9)
  CB_CASE(5, 8) CB_CASE(5, 9)
  CB_CASE(6, 8)
#undef CB_CASE
  return false;
}

Family key_family(u32 k, u32 N, u32 level) {
  const int plan = g_pbs_plan.load(std::memory_order_relaxed);
  if (plan == PLAN_GENERIC || plan == PLAN_GRP || plan == PLAN_CLUSTER) return FAM_GENERIC;
  if (plan == PLAN_V4 && cb_pbs_v4_shape(k, N)) return FAM_V4;
  if (use_w32(k, N, level)) return FAM_W32;
  // k = 2 (P4): the KEY is v4's too.  Measured on B200 (profiles/r2_p4_v4_warp_per_poly.txt): v4 with
  // one warp per polynomial, 5 ciphertexts per CTA and its twiddles in tensor memory runs 16 384
  // samples in 191.8 ms against 228.5 ms for the group kernel, but 64 samples in 4.7 ms against
  // 2.7 ms (one warp per polynomial is a long dependent chain: the latency of ONE PBS).  Both read
  // the same converted key (pbs_grp.cuh FOLD), so the batch size picks the kernel at launch.
  if (cb_pbs_v4_shape(k, N)) return FAM_V4;
  return FAM_GENERIC;
}

// k = 2, N = 1024 under FAM_V4: batches below this run the group kernel on the v4 key (one
// ciphertext per CTA, several CTAs per SM).  Measured on B200 (profiles/r2_p4_v4_warp_per_poly.txt):
// 64 samples 2.9 ms (group) against 4.7 ms (v4); 739 samples 13.6 ms against 8.6 ms: the v4 kernel
// is ahead as soon as its CTAs hold more than two ciphertexts each
constexpr u32 V4_K2_MIN_BATCH = 2 * 148 + 1;

constexpr u32 PBS_MAGIC = 0xB200CB01u;

struct PbsBuffer {
  u32 magic;
  int gpu;
  u32 k, N, level, max_samples;
  int grid_cap;           // resident CTAs for the persistent launch
  int grp_grid_cap;       // same for the group-per-polynomial kernel (0: not usable)
  int cl_cluster_cap;     // resident clusters of the cluster variant (0: not usable)
  size_t smem_fixed;      // dynamic smem without the a_hat table
  ulonglong2 *global_acc; // device scratch when the accumulator does not fit in smem
  Variant v;
};

size_t pbs_smem_bytes(const Variant &v, u32 n) {
  size_t ahat = (((size_t)n * 2 + 15) / 16) * 16;
  return v.fbuf_bytes + (v.acc_global ? 0 : v.acc_bytes) + ahat;
}

} // namespace

// internal entry points shared with legacy.cu / chain
void cb_pbs_scratch(cudaStream_t stream, int gpu, int8_t **buffer, u32 k, u32 N, u32 level,
                    u32 count, bool allocate) {
  Variant v;
  if (!get_variant(k, N, v))
    CB_PANIC("Cuda error (classical PBS): unsupported (glwe_dimension=%u, polynomial_size=%u). "
             "Supported N: powers of two in [256..16384].",
             k, N);
  PbsBuffer *b = new PbsBuffer();
  b->magic = PBS_MAGIC;
  b->gpu = gpu;
  b->k = k;
  b->N = N;
  b->level = level;
  b->max_samples = count;
  b->v = v;
  b->global_acc = nullptr;
  get_twiddles(gpu, (int)std::lround(std::log2((double)N)) - 1);
  // size the persistent grid from the real occupancy with a typical a_hat table (n <= 2048)
  const size_t smem = pbs_smem_bytes(v, 2048);
  CB_CHECK(cudaFuncSetAttribute(v.pbs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0, sms = 0;
  CB_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, v.pbs, v.T, smem));
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  if (per_sm < 1)
    CB_PANIC("Cuda error (classical PBS): kernel does not fit on an SM (smem %zu B)", smem);
  b->grid_cap = per_sm * sms;
  b->grp_grid_cap = 0;
  if (v.grp_fits) {
    const size_t gsmem = v.grp_smem_fixed + 4096;
    int occ = 0;
    CB_CHECK(cudaFuncSetAttribute(v.pbs_grp, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)gsmem));
    CB_CHECK(cudaFuncSetAttribute(v.pbs_grp_w, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)gsmem));
    if (v.pbs_grp_fold) {
      CB_CHECK(cudaFuncSetAttribute(v.pbs_grp_fold, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)gsmem));
      CB_CHECK(cudaFuncSetAttribute(v.pbs_grp_fold_w, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)gsmem));
    }
    CB_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, v.pbs_grp, v.grp_threads, gsmem));
    b->grp_grid_cap = occ * sms;
  }
  b->cl_cluster_cap = 0;
  if (v.cl_fits) {
    const size_t csmem = v.cl_smem_fixed + 4096 + v.cl_stash_bytes;
    CB_CHECK(cudaFuncSetAttribute(v.pbs_grp_cl, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)csmem));
    CB_CHECK(cudaFuncSetAttribute(v.pbs_grp_cl_w, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)csmem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((k + 1) * (u32)sms, 1, 1);
    cfg.blockDim = dim3((u32)v.T, 1, 1);
    cfg.dynamicSmemBytes = csmem;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = k + 1;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, v.pbs_grp_cl, &cfg) == cudaSuccess)
      b->cl_cluster_cap = clusters;
    else
      (void)cudaGetLastError();
  }
  if (v.acc_global && allocate) {
    size_t ctas = (size_t)std::min<u32>((u32)b->grid_cap, count ? count : 1);
    CB_CHECK(cudaMallocAsync((void **)&b->global_acc, ctas * v.acc_bytes, stream));
  }
  *buffer = reinterpret_cast<int8_t *>(b);
}

void cb_pbs_cleanup(cudaStream_t stream, int gpu, int8_t **buffer) {
  (void)gpu;
  if (!buffer || !*buffer) return;
  PbsBuffer *b = reinterpret_cast<PbsBuffer *>(*buffer);
  if (b->magic != PBS_MAGIC) CB_PANIC("Cuda error (classical PBS): invalid pbs buffer");
  if (b->global_acc) CB_CHECK(cudaFreeAsync(b->global_acc, stream));
  b->magic = 0;
  delete b; // the reference leaks the host object (programmable_bootstrap.cu cleanup); we do not
  *buffer = nullptr;
}

void cb_pbs_launch(cudaStream_t stream, int gpu, u64 *lwe_out, const u64 *out_idx,
                   const u64 *lut_vector, const u64 *lut_idx, const u64 *lwe_in,
                   const u64 *in_idx, const double2 *bsk, int8_t *buffer, u32 n, u32 k, u32 N,
                   u32 base_log, u32 level, u32 num_samples, u32 lut_count, u32 lut_stride) {
  if (num_samples == 0) return;
  PbsBuffer *b = reinterpret_cast<PbsBuffer *>(buffer);
  if (!b || b->magic != PBS_MAGIC) CB_PANIC("Cuda error (classical PBS): invalid pbs buffer");
  if (b->k != k || b->N != N)
    CB_PANIC("Cuda error (classical PBS): buffer was created for (k=%u,N=%u), called with "
             "(k=%u,N=%u)", b->k, b->N, k, N);
  if (base_log == 0 || level == 0 || base_log * level >= 64 || base_log > 31)
    CB_PANIC("Cuda error (classical PBS): unsupported decomposition base_log=%u level=%u",
             base_log, level);
  if (lut_count == 0) lut_count = 1;
  if ((u64)(lut_count - 1) * lut_stride >= N)
    CB_PANIC("Cuda error (classical PBS): lut_count*lut_stride exceeds polynomial size");
  const Family fam = key_family(k, N, level);
  // small batch on a v4 key that the group kernel can read (k = 2, N = 1024): see key_family
  const bool grp_on_v4_key = fam == FAM_V4 && b->v.pbs_grp_fold && b->grp_grid_cap > 0 && n <= 2048 &&
                             g_pbs_plan.load(std::memory_order_relaxed) == PLAN_AUTO &&
                             cb_pbs_v4_logp(k, N) == 4 && num_samples < V4_K2_MIN_BATCH;
  if (fam == FAM_V4 && !grp_on_v4_key) {
    cb_pbs_v4_launch(stream, gpu, lwe_out, out_idx, lut_vector, lut_idx, lwe_in, in_idx, bsk, n, k,
                     N, base_log, level, num_samples, lut_count, lut_stride);
    return;
  }
  if (fam == FAM_W32) {
    PbsArgs a;
    a.lwe_out = lwe_out;
    a.out_idx = out_idx;
    a.lut_vector = lut_vector;
    a.lut_idx = lut_idx;
    a.lwe_in = lwe_in;
    a.in_idx = in_idx;
    a.bsk = bsk;
    a.tw = get_twiddles(gpu, 10);
    a.global_acc = nullptr;
    a.n = n;
    a.base_log = base_log;
    a.level = level;
    a.num_samples = num_samples;
    a.lut_count = lut_count;
    a.lut_stride = lut_stride;
    int sms = 0;
    CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
    auto kern = w32::pbs_kernel_w32v3;
    CB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)w32::SMEM_BYTES3));
    const u32 grid = std::min<u32>((u32)sms, num_samples); // the kernel balances the tail
    kern<<<grid, w32::NTHREADS3, w32::SMEM_BYTES3, stream>>>(a);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    return;
  }
  const Variant &v = b->v;
  u32 grid = std::min<u32>((u32)b->grid_cap, num_samples);
  if (v.acc_global) {
    if (!b->global_acc)
      CB_PANIC("Cuda error (classical PBS): scratch was created without GPU memory");
    grid = std::min<u32>(grid, std::max<u32>(1, std::min<u32>((u32)b->grid_cap, b->max_samples)));
  }
  PbsArgs a;
  a.lwe_out = lwe_out;
  a.out_idx = out_idx;
  a.lut_vector = lut_vector;
  a.lut_idx = lut_idx;
  a.lwe_in = lwe_in;
  a.in_idx = in_idx;
  a.bsk = bsk;
  a.tw = get_twiddles(gpu, (int)std::lround(std::log2((double)N)) - 1);
  a.global_acc = b->global_acc;
  a.n = n;
  a.base_log = base_log;
  a.level = level;
  a.num_samples = num_samples;
  a.lut_count = lut_count;
  a.lut_stride = lut_stride;
  // The cluster variant (one CTA per polynomial, F through DSMEM) is bit-compatible and fits
  // N = 8192, but measured slower than the first kernel there (P7: 671 vs 633 ms: two cluster
  // barriers per CMUX with one CTA per SM); it stays selectable for experiments and tests.
  // Exception: when the accumulator lives in global memory anyway (N = 16384) the cluster
  // variant is the default: one polynomial per CTA keeps 16 points per thread in 128 registers
  // where the first kernel is capped at 64 registers with 2 KB of spills per thread.
  const int plan = g_pbs_plan.load(std::memory_order_relaxed);
  const bool force_cluster =
      plan == PLAN_CLUSTER || (plan != PLAN_GENERIC && v.acc_global && b->global_acc);
  if (plan != PLAN_GENERIC && b->grp_grid_cap > 0 && n <= 2048 &&
      !(force_cluster && b->cl_cluster_cap > 0)) {
    const size_t gsmem = v.grp_smem_fixed + (((size_t)n * 2 + 15) / 16) * 16;
    const u32 ggrid = std::min<u32>((u32)b->grp_grid_cap, num_samples);
    const bool narrow = base_log * level <= 31;
    (grp_on_v4_key ? (narrow ? v.pbs_grp_fold : v.pbs_grp_fold_w)
                   : (narrow ? v.pbs_grp : v.pbs_grp_w))<<<ggrid, v.grp_threads, gsmem, stream>>>(a);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    return;
  }
  if (force_cluster && b->cl_cluster_cap > 0 && n <= 2048) {
    u32 clusters = std::min<u32>((u32)b->cl_cluster_cap, num_samples);
    if (v.acc_global) // one accumulator slot per cluster in the scratch buffer
      clusters = std::min<u32>(clusters, std::max<u32>(1, std::min<u32>((u32)b->grid_cap, b->max_samples)));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(clusters * (k + 1), 1, 1);
    cfg.blockDim = dim3((u32)v.T, 1, 1);
    cfg.dynamicSmemBytes = v.cl_stash_bytes ? v.cl_smem_fixed + 4096 + v.cl_stash_bytes
                                            : v.cl_smem_fixed + (((size_t)n * 2 + 15) / 16) * 16;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = k + 1;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    CB_CHECK(cudaLaunchKernelEx(&cfg, base_log * level <= 31 ? v.pbs_grp_cl : v.pbs_grp_cl_w, a));
    CB_LAUNCHED();
    return;
  }
  const size_t smem = pbs_smem_bytes(v, n);
  CB_CHECK(cudaFuncSetAttribute(v.pbs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  v.pbs<<<grid, v.T, smem, stream>>>(a);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

void cb_bsk_convert(cudaStream_t stream, int gpu, void *dest, const void *src_host, u32 n, u32 k,
                    u32 level, u32 N, bool src_on_device) {
  Variant v;
  if (!get_variant(k, N, v))
    CB_PANIC("Cuda error (convert BSK): unsupported (glwe_dimension=%u, polynomial_size=%u)", k,
             N);
  const size_t polys = (size_t)n * level * (k + 1) * (k + 1);
  const size_t bytes = polys * N * sizeof(u64);
  if (bytes == 0) return;
  if (dest != src_host) // (the seeded-key path decompresses straight into dest: seeded.cu)
    CB_CHECK(cudaMemcpyAsync(dest, src_host, bytes,
                             src_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                             stream));
  const double2 *tw = get_twiddles(gpu, (int)std::lround(std::log2((double)N)) - 1);
  int sms = 0;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  const Family fam = key_family(k, N, level);
  if (fam == FAM_V4) {
    cb_pbs_v4_convert(stream, gpu, dest, polys, k, N);
    return;
  }
  if (fam == FAM_W32) {
    const size_t smem = w32::SMEM_TW + 4 * (size_t)w32::M * sizeof(double2);
    CB_CHECK(cudaFuncSetAttribute(w32::bsk_convert_kernel_w32,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const u32 grid = (u32)std::min<size_t>((polys + 3) / 4, (size_t)sms * 2);
    w32::bsk_convert_kernel_w32<<<grid, 128, smem, stream>>>(reinterpret_cast<double2 *>(dest),
                                                             tw, polys);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    return;
  }
  CB_CHECK(cudaFuncSetAttribute(v.convert, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)v.fbuf_bytes));
  const u32 grid = (u32)std::min<size_t>(polys, (size_t)sms * 8);
  v.convert<<<grid, v.T, v.fbuf_bytes, stream>>>(reinterpret_cast<double2 *>(dest), tw, polys);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

// One Fourier GGSW list in the GENERIC kernel's layout (what cmux_kernel reads), in place on the
// device, whatever family serves (k, N) for the PBS.
void cb_ggsw_convert_generic(cudaStream_t stream, int gpu, void *data, size_t polys, u32 k, u32 N) {
  Variant v;
  if (!get_variant(k, N, v))
    CB_PANIC("Cuda error (convert GGSW): unsupported (glwe_dimension=%u, polynomial_size=%u)", k, N);
  if (polys == 0) return;
  const double2 *tw = get_twiddles(gpu, (int)std::lround(std::log2((double)N)) - 1);
  int sms = 0;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  CB_CHECK(cudaFuncSetAttribute(v.convert, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)v.fbuf_bytes));
  const u32 grid = (u32)std::min<size_t>(polys, (size_t)sms * 8);
  v.convert<<<grid, v.T, v.fbuf_bytes, stream>>>(reinterpret_cast<double2 *>(data), tw, polys);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

// out[p] = in[2p] + ggsw (x) (in[2p+1] - in[2p]) for p < pairs (GLWEs of (k+1)N words)
void cb_cmux_launch(cudaStream_t stream, int gpu, u64 *out, const u64 *in, const double2 *ggsw,
                    u32 k, u32 N, u32 base_log, u32 level, u32 pairs) {
  Variant v;
  if (!get_variant(k, N, v))
    CB_PANIC("Cuda error (cmux): unsupported (glwe_dimension=%u, polynomial_size=%u)", k, N);
  if (pairs == 0) return;
  const double2 *tw = get_twiddles(gpu, (int)std::lround(std::log2((double)N)) - 1);
  int sms = 0;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  CB_CHECK(cudaFuncSetAttribute(v.cmux, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)v.fbuf_bytes));
  const u32 grid = std::min<u32>(pairs, (u32)sms * 4);
  v.cmux<<<grid, v.T, v.fbuf_bytes, stream>>>(out, in, ggsw, tw, base_log, level, pairs);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

extern "C" {

// test hook, see "Kernel plan" above
void concrete_b200_set_pbs_plan(int plan) { g_pbs_plan.store(plan, std::memory_order_relaxed); }

void cuda_convert_lwe_programmable_bootstrap_key_64(void *stream, uint32_t gpu_index, void *dest,
                                                    void *src, uint32_t input_lwe_dim,
                                                    uint32_t glwe_dim, uint32_t level_count,
                                                    uint32_t polynomial_size) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cudaPointerAttributes attr;
  bool on_device = false;
  if (cudaPointerGetAttributes(&attr, src) == cudaSuccess)
    on_device = (attr.type == cudaMemoryTypeDevice);
  else
    (void)cudaGetLastError();
  cb_bsk_convert(as_stream(stream), gpu_index, dest, src, input_lwe_dim, glwe_dim, level_count,
                 polynomial_size, on_device);
}

void scratch_cuda_programmable_bootstrap_64(void *stream, uint32_t gpu_index, int8_t **buffer,
                                            uint32_t glwe_dimension, uint32_t polynomial_size,
                                            uint32_t level_count,
                                            uint32_t input_lwe_ciphertext_count,
                                            bool allocate_gpu_memory) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_scratch(as_stream(stream), gpu_index, buffer, glwe_dimension, polynomial_size,
                 level_count, input_lwe_ciphertext_count, allocate_gpu_memory);
}

void cuda_programmable_bootstrap_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lut_vector, void *lut_vector_indexes, void *lwe_array_in, void *lwe_input_indexes,
    void *bootstrapping_key, int8_t *buffer, uint32_t lwe_dimension, uint32_t glwe_dimension,
    uint32_t polynomial_size, uint32_t base_log, uint32_t level_count, uint32_t num_samples,
    uint32_t lut_count, uint32_t lut_stride) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_launch(as_stream(stream), gpu_index, (u64 *)lwe_array_out,
                (const u64 *)lwe_output_indexes, (const u64 *)lut_vector,
                (const u64 *)lut_vector_indexes, (const u64 *)lwe_array_in,
                (const u64 *)lwe_input_indexes, (const double2 *)bootstrapping_key, buffer,
                lwe_dimension, glwe_dimension, polynomial_size, base_log, level_count,
                num_samples, lut_count, lut_stride);
}

void cleanup_cuda_programmable_bootstrap(void *stream, uint32_t gpu_index, int8_t **pbs_buffer) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_cleanup(as_stream(stream), gpu_index, pbs_buffer);
}

// "amortized" names: same fused kernel (one CTA per ciphertext is exactly the amortized design,
// programmable_bootstrap_amortized.cuh:48-213, without its 48 KiB shared-memory ceiling).
void scratch_cuda_programmable_bootstrap_amortized_64(void *stream, uint32_t gpu_index,
                                                      int8_t **pbs_buffer,
                                                      uint32_t glwe_dimension,
                                                      uint32_t polynomial_size,
                                                      uint32_t input_lwe_ciphertext_count,
                                                      bool allocate_gpu_memory) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_scratch(as_stream(stream), gpu_index, pbs_buffer, glwe_dimension, polynomial_size, 0,
                 input_lwe_ciphertext_count, allocate_gpu_memory);
}

void cuda_programmable_bootstrap_amortized_lwe_ciphertext_vector_64(
    void *stream, uint32_t gpu_index, void *lwe_array_out, void *lwe_output_indexes,
    void *lut_vector, void *lut_vector_indexes, void *lwe_array_in, void *lwe_input_indexes,
    void *bootstrapping_key, int8_t *pbs_buffer, uint32_t lwe_dimension,
    uint32_t glwe_dimension, uint32_t polynomial_size, uint32_t base_log, uint32_t level_count,
    uint32_t num_samples) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_launch(as_stream(stream), gpu_index, (u64 *)lwe_array_out,
                (const u64 *)lwe_output_indexes, (const u64 *)lut_vector,
                (const u64 *)lut_vector_indexes, (const u64 *)lwe_array_in,
                (const u64 *)lwe_input_indexes, (const double2 *)bootstrapping_key, pbs_buffer,
                lwe_dimension, glwe_dimension, polynomial_size, base_log, level_count,
                num_samples, 1, 1);
}

void cleanup_cuda_programmable_bootstrap_amortized(void *stream, uint32_t gpu_index,
                                                   int8_t **pbs_buffer) {
  CB_CHECK(cudaSetDevice(gpu_index));
  cb_pbs_cleanup(as_stream(stream), gpu_index, pbs_buffer);
}

void cuda_fourier_polynomial_mul(void *stream, uint32_t gpu_index, void *input1, void *input2,
                                 void *output, uint32_t polynomial_size,
                                 uint32_t total_polynomials) {
  CB_CHECK(cudaSetDevice(gpu_index));
  if (total_polynomials == 0) return;
  Variant v;
  if (!get_variant(1, polynomial_size, v))
    CB_PANIC("Cuda error (fourier mul): unsupported polynomial size %u", polynomial_size);
  const double2 *tw =
      get_twiddles(gpu_index, (int)std::lround(std::log2((double)polynomial_size)) - 1);
  if (polynomial_size == 2048) {
    const size_t smem = w32::SMEM_TW + 4 * (size_t)w32::M * sizeof(double2);
    CB_CHECK(cudaFuncSetAttribute(w32::poly_mul_kernel_w32,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const u32 grid = std::min<u32>((total_polynomials + 3) / 4, 148 * 2);
    w32::poly_mul_kernel_w32<<<grid, 128, smem, as_stream(stream)>>>(
        (const double *)input1, (const double *)input2, (double *)output, tw, total_polynomials);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    return;
  }
  CB_CHECK(cudaFuncSetAttribute(v.polymul, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)v.fbuf_bytes));
  const u32 grid = std::min<u32>(total_polynomials, 148 * 8);
  v.polymul<<<grid, v.T, v.fbuf_bytes, as_stream(stream)>>>(
      (const double *)input1, (const double *)input2, (double *)output, tw, total_polynomials);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

} // extern "C"
