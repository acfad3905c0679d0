// pbs_grp.cuh — fused PBS with one THREAD GROUP per GLWE polynomial (every (k, N) whose working
// set fits in shared memory; included by pbs.cu inside its anonymous namespace).
//
// Same arithmetic and the same Fourier-key layout as pbs_kernel (pbs.cu); what changes is who
// holds what.  ncu/ptxas on pbs_kernel: every thread carries the Fourier accumulators of ALL k+1
// output polynomials (8(k+1)P registers) next to its transform points and the decomposition
// states, ptxas spills (P4: 236 B, P6: 472 B per thread) and the (k+1)·l forward transforms of a
// CMUX run one after the other.  Here a CTA has k+1 groups of T = M/P threads:
//   phase A  group r: ct1_r = ACC_r·X^a − ACC_r, rounding, digits of level m, forward transform,
//            result F_{r,m} parked in the group's exchange buffer            (groups in parallel)
//   phase B  group c: out_c += Σ_r F_{r,m} ⊙ GGSW[m][r][c]     (F from shared memory, key from L2)
//   phase C  group c: inverse transform of out_c, ACC_c += round(·)          (groups in parallel)
// so a thread holds P points + P accumulators of ONE polynomial (128 registers, no spills), the
// k+1 transforms of a phase overlap, and occupancy is set by shared memory, not registers.
// CTA-wide barriers: two per decomposition level; transforms use one named barrier per group.
//
// CLUSTER variant (same body): when k+1 exchange buffers + the accumulator exceed one SM's shared
// memory (N = 8192), the k+1 groups become the k+1 CTAs of a thread-block cluster, each holding
// ONE polynomial's accumulator and exchange buffer; phase B reads the partners' F_{r,m} through
// distributed shared memory and the CTA-wide barriers become cluster barriers.
#pragma once
// (pbs.cu includes <cooperative_groups.h> at file scope: this header sits in its namespace)

template <int K, int LOGN> struct GrpCfg {
  using C = PbsCfg<K, LOGN>;
  static constexpr int GS = K + 1;
  static constexpr int T = C::T;            // threads per group (same split as pbs_kernel, so the
  static constexpr int P = C::P;            // Fourier key layout dst[p*T + t] is shared)
  static constexpr int THREADS = GS * T;
  static constexpr size_t SMEM_FIXED = (size_t)GS * C::FBUF_BYTES + C::ACC_BYTES;
  static constexpr bool FITS = THREADS <= 512 && SMEM_FIXED + 4096 <= 232448; // >= 128 regs/thread
  static constexpr int BY_SMEM = FITS ? (int)(232448 / (SMEM_FIXED + 4096)) : 1;
  static constexpr int BY_REGS = (65536 / 128) / THREADS;  // spill-free at ~128 registers
  static constexpr int MINB_ = BY_SMEM < BY_REGS ? BY_SMEM : BY_REGS;
  static constexpr int MINB = MINB_ < 1 ? 1 : (MINB_ > 8 ? 8 : MINB_);
  // cluster variant: one CTA per polynomial
  static constexpr size_t CL_SMEM_FIXED = C::FBUF_BYTES + (C::ACC_GLOBAL ? 0 : C::ACC_BYTES / GS);
  static constexpr bool CL_FITS = T <= 512 && GS <= 8 && CL_SMEM_FIXED + 4096 <= 232448;
  static constexpr int CL_MINB_ = (int)(232448 / (CL_SMEM_FIXED + 4096)) < (65536 / 128) / T
                                      ? (int)(232448 / (CL_SMEM_FIXED + 4096))
                                      : (65536 / 128) / T;
  static constexpr int CL_MINB = CL_MINB_ < 1 ? 1 : CL_MINB_;
  // N = 16384 (accumulator in global memory): the shared memory left next to the exchange buffer
  // parks the real parts of the Fourier accumulators during the forward transforms of levels
  // >= 1 (32 registers at P = 16; ptxas: 940 -> ? bytes of spills)
  static constexpr size_t CL_STASH_OFF = CL_SMEM_FIXED + 4096;
  static constexpr size_t CL_STASH_BYTES = C::ACC_GLOBAL ? (size_t)T * P * sizeof(double) : 0;
};

// FOLD > 0: the Fourier key is in the layout of the v4 kernel with 2^FOLD times as many points per
// thread (pbs_v4.cuh: F_q at (q mod P4) * T4 + q / P4 with P4 = P << FOLD, T4 = T >> FOLD), so that
// ONE converted key serves both kernels and the batch size alone picks between them at launch
// (k = 2, N = 1024: this kernel for small batches - the latency of one PBS -, v4 for large ones).
// Thread t still owns the points q = t*P + p: they sit at ((t mod 2^FOLD) * P + p) * T4 + (t >> FOLD),
// i.e. a warp reads 2^FOLD runs of 32 >> FOLD consecutive elements per row instead of one run of 32.
template <int K, int LOGN, bool CLUSTER, bool NARROW, int FOLD = 0>
__device__ __forceinline__ void pbs_grp_body(const PbsArgs &a) {
  using st_t = typename std::conditional<NARROW, u32, u64>::type;
  namespace cg = cooperative_groups;
  using C = PbsCfg<K, LOGN>;
  using G = GrpCfg<K, LOGN>;
  constexpr int LOGM = C::LOGM, LOGP = C::LOGP, M = C::M, N = C::N, P = C::P, T = C::T;
  constexpr int GS = K + 1, THREADS = CLUSTER ? T : G::THREADS;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x;
  // polynomial index (row in phase A, column in B/C) and index inside the group
  const int grp = CLUSTER ? (int)cg::this_cluster().block_rank() : tid / T;
  const int t = CLUSTER ? tid : tid % T;
  // shared-memory map: CTA mode [GS exchange buffers][GS accumulators][a_hat];
  //                    cluster mode [1 exchange buffer][1 accumulator][a_hat] per CTA
  double2 *fbuf = reinterpret_cast<double2 *>(smem) + (CLUSTER ? 0 : (size_t)grp * C::F::BUF_ELEMS);
  // (N = 16384 in cluster mode: the accumulator of this CTA's polynomial lives in global memory)
  constexpr bool ACC_G = CLUSTER && C::ACC_GLOBAL;
  ulonglong2 *acc =
      ACC_G ? a.global_acc + (size_t)blockIdx.x * M
            : reinterpret_cast<ulonglong2 *>(smem + (CLUSTER ? 1 : GS) * C::FBUF_BYTES) +
                  (CLUSTER ? 0 : (size_t)grp * M);
  unsigned short *s_ahat = reinterpret_cast<unsigned short *>(
      smem + (CLUSTER ? G::CL_SMEM_FIXED : GS * C::FBUF_BYTES + C::ACC_BYTES));
  const double2 *fsrc[GS]; // where phase B finds F_{r,m} (own SM or a partner's, through DSMEM)
#pragma unroll
  for (int r = 0; r < GS; ++r) {
    if (CLUSTER)
      fsrc[r] = cg::this_cluster().map_shared_rank(reinterpret_cast<double2 *>(smem), r) + t;
    else
      fsrc[r] = reinterpret_cast<double2 *>(smem) + (size_t)r * C::F::BUF_ELEMS + t;
  }
  auto sync_all = [&]() {
    if (CLUSTER) cg::this_cluster().sync(); else __syncthreads();
  };
  const u32 first_sample = CLUSTER ? blockIdx.x / GS : blockIdx.x;
  const u32 sample_stride = CLUSTER ? gridDim.x / GS : gridDim.x;
  const int bar_id = CLUSTER ? 1 : 1 + grp;
  constexpr int KT = T >> FOLD; // key row stride and this thread's offset inside a key row
  const int koff = FOLD ? (t & ((1 << FOLD) - 1)) * P * KT + (t >> FOLD) : t;
  const u32 n = a.n, base_log = a.base_log, level = a.level;
  const st_t mask = (st_t)(((u64)1 << base_log) - 1);
  const u32 round_shift = 64 - base_log * level - 1;
  const bool fast1 = (level == 1); // uniform over the grid
  const u32 state_shift = 64 - base_log * level;
  const double2 *__restrict__ tw = a.tw;
  const size_t out_size = (size_t)K * N + 1;

  for (u32 sample = first_sample; sample < a.num_samples; sample += sample_stride) {
    const u64 *lwe = a.lwe_in + a.in_idx[sample] * (u64)(n + 1);
    for (u32 i = tid; i < n; i += THREADS)
      s_ahat[i] = (unsigned short)modulus_switch(lwe[i], LOGN + 1);
    const u32 b_hat = modulus_switch(lwe[n], LOGN + 1);
    { // ACC_grp <- LUT_grp * X^(-b_hat)   (polynomial_algorithms.rs:500-539)
      const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(GS * N) + (size_t)grp * N;
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        ulonglong2 v;
        {
          const u32 u = (j + b_hat) & (2 * N - 1);
          const u64 w = lut[u & (N - 1)];
          v.x = (u < (u32)N) ? w : (u64)0 - w;
        }
        {
          const u32 u = (j + M + b_hat) & (2 * N - 1);
          const u64 w = lut[u & (N - 1)];
          v.y = (u < (u32)N) ? w : (u64)0 - w;
        }
        acc[j] = v;
      }
    }
    sync_all();

    for (u32 i = 0; i < n; ++i) {
      const u32 a_hat = s_ahat[i];
      if (a_hat == 0) continue; // uniform over the CTA (bootstrap.rs:287)

      // ---- phase A (first part): rotation, rounding, decomposition state of polynomial grp
      st_t st_re[P], st_im[P];
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        const u32 s = (u32)(j - (int)a_hat) & (2 * N - 1);
        const u32 q = s >> LOGM;
        const ulonglong2 src = acc[s & (M - 1)];
        const ulonglong2 own = acc[j];
        const u64 ra = (q & 1) ? src.y : src.x;
        const u64 rb = (q & 1) ? src.x : src.y;
        const u64 rot_re = (q == 2 || q == 3) ? (u64)0 - ra : ra;
        const u64 rot_im = (q == 1 || q == 2) ? (u64)0 - rb : rb;
        if (fast1) { // one level: rounding + balanced digit from the high word (pbs_v4.cuh FAST1)
          const u32 kshift = 32 - base_log, half_m1 = (1u << (kshift - 1)) - 1u;
          st_re[p] = (st_t)(u32)(-((int)(half_m1 - (u32)((rot_re - own.x) >> 32)) >> kshift));
          st_im[p] = (st_t)(u32)(-((int)(half_m1 - (u32)((rot_im - own.y) >> 32)) >> kshift));
        } else {
          st_re[p] = (st_t)(closest_representable(rot_re - own.x, round_shift) >> state_shift);
          st_im[p] = (st_t)(closest_representable(rot_im - own.y, round_shift) >> state_shift);
        }
      }
      double2 oacc[P];
#pragma unroll
      for (int p = 0; p < P; ++p) oacc[p] = make_double2(0.0, 0.0);

#pragma unroll 1
      for (u32 m = 0; m < level; ++m) { // memory level index m <-> level l-m (ggsw.rs:231-245)
        if (ACC_G && m > 0) {
          double *stash = reinterpret_cast<double *>(smem + G::CL_STASH_OFF);
#pragma unroll
          for (int p = 0; p < P; ++p) stash[p * T + t] = oacc[p].x;
        }
        {
          double2 x[P];
#pragma unroll
          for (int p = 0; p < P; ++p) {
            x[p].x = __int2double_rn(fast1 ? (int)(u32)st_re[p]
                                           : (int)(u32)grp_decompose<st_t>(st_re[p], mask, base_log));
            x[p].y = __int2double_rn(fast1 ? (int)(u32)st_im[p]
                                           : (int)(u32)grp_decompose<st_t>(st_im[p], mask, base_log));
          }
          cbfft::forward<LOGM, LOGP>(x, fbuf, t, tw, bar_id);
          cbfft::group_sync<T>(bar_id); // last exchange of the transform fully read
#pragma unroll
          for (int p = 0; p < P; ++p) fbuf[p * T + t] = x[p]; // F_{grp,m} in key order
        }
        if (ACC_G) { // accumulators come (back) to life only after the transform
          const double *stash = reinterpret_cast<const double *>(smem + G::CL_STASH_OFF);
#pragma unroll
          for (int p = 0; p < P; ++p) {
            if (m == 0) oacc[p] = make_double2(0.0, 0.0);
            else oacc[p].x = stash[p * T + t];
          }
        }
        const double2 *__restrict__ g =
            a.bsk + (((size_t)i * level + m) * GS * GS + grp) * M + koff;
        // ---- phase B: column grp of the external product.  The own term (r = grp) reads
        // back exactly what this thread stored, so in cluster mode it runs between the arrive
        // and the wait of the cluster barrier and covers the barrier's latency.
        if (CLUSTER) {
          asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
#pragma unroll
          for (int p = 0; p < P; ++p) {
            const double2 w = __ldg(&g[(size_t)grp * GS * M + p * KT]);
            const double2 f = fbuf[p * T + t];
            oacc[p].x = fma(f.x, w.x, fma(-f.y, w.y, oacc[p].x));
            oacc[p].y = fma(f.x, w.y, fma(f.y, w.x, oacc[p].y));
          }
          asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        } else {
          sync_all(); // every F_{r,m} is visible
        }
#pragma unroll
        for (int r = 0; r < GS; ++r) {
          if (CLUSTER && r == grp) continue; // done above
          const double2 *fr = fsrc[r];
#pragma unroll
          for (int p = 0; p < P; ++p) {
            const double2 w = __ldg(&g[(size_t)r * GS * M + p * KT]);
            const double2 f = fr[p * T];
            oacc[p].x = fma(f.x, w.x, fma(-f.y, w.y, oacc[p].x));
            oacc[p].y = fma(f.x, w.y, fma(f.y, w.x, oacc[p].y));
          }
        }
        sync_all(); // all reads of the F buffers done before they are reused
      }

      // ---- phase C: inverse transform of column grp, torus rounding, accumulate
      cbfft::inverse<LOGM, LOGP>(oacc, fbuf, t, tw, bar_id);
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int j = t + p * T;
        ulonglong2 v = acc[j];
        v.x += from_torus(oacc[p].x);
        v.y += from_torus(oacc[p].y);
        acc[j] = v;
      }
      cbfft::group_sync<T>(bar_id); // ACC_grp complete before its rotated reads of the next CMUX
    }

    // sample extraction (glwe_sample_extraction.rs:90-165) at nth = li*lut_stride: group c < k
    // writes mask polynomial c from its own accumulator, group k the body
    for (u32 li = 0; li < a.lut_count; ++li) {
      const u32 nth = li * a.lut_stride;
      u64 *out = a.lwe_out + ((size_t)li * a.num_samples + a.out_idx[sample]) * out_size;
      if (grp < K) {
        for (int j = t; j < N; j += T) {
          const int src = (j <= (int)nth) ? ((int)nth - j) : (N + (int)nth - j);
          const ulonglong2 pr = acc[src & (M - 1)];
          const u64 v = (src < M) ? pr.x : pr.y;
          out[(size_t)grp * N + j] = (j <= (int)nth) ? v : (u64)0 - v;
        }
      } else if (t == 0) {
        const ulonglong2 pr = acc[nth & (M - 1)];
        out[(size_t)K * N] = (nth < (u32)M) ? pr.x : pr.y;
      }
    }
    sync_all(); // a_hat table and buffers are reused by the next sample
  }
}

template <int K, int LOGN, bool NARROW, int FOLD = 0>
__global__ void __launch_bounds__(GrpCfg<K, LOGN>::THREADS, GrpCfg<K, LOGN>::MINB)
pbs_kernel_grp(const PbsArgs a) {
  pbs_grp_body<K, LOGN, false, NARROW, FOLD>(a);
}

// launched with cluster dimension (k+1, 1, 1): CTA rank in the cluster = polynomial index
template <int K, int LOGN, bool NARROW>
__global__ void __launch_bounds__(PbsCfg<K, LOGN>::T, GrpCfg<K, LOGN>::CL_MINB)
pbs_kernel_grp_cluster(const PbsArgs a) {
  pbs_grp_body<K, LOGN, true, NARROW>(a);
}
