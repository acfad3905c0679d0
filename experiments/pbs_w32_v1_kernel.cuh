// experiments/pbs_w32_v1_kernel.cuh — first warp-per-polynomial PBS kernel (round 1, 388 ms at
// P2048 x 16 384 = 22.5 % of the FP64 peak).  Superseded by pbs_w32_v3.cuh; kept as a measured
// baseline.  To build it again include it from pbs.cu right after pbs_w32.cuh.
#pragma once
namespace w32 {
// ---- K3 fast path ------------------------------------------------------------------------------
__global__ void __launch_bounds__(NTHREADS, 1) pbs_kernel_w32(const PbsArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *tws = reinterpret_cast<double2 *>(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ct = warp >> 1, r = warp & 1;
  unsigned char *base = smem + SMEM_TW + (size_t)ct * SMEM_PER_CT;
  ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(base) + r * M; // this warp's polynomial
  double2 *buf_own = reinterpret_cast<double2 *>(base + (size_t)ACC_PAIRS * 16) + r * M;
  double2 *buf_oth = reinterpret_cast<double2 *>(base + (size_t)ACC_PAIRS * 16) + (1 - r) * M;
  const int bar_id = 1 + ct;

  load_tws(tws, a.tw, tid, NTHREADS);
  __syncthreads();

  const u32 n = a.n, base_log = a.base_log;
  const size_t out_size = (size_t)N + 1;
  const double2 *__restrict__ gown = a.bsk + (size_t)(r * 2 + r) * M + lane;
  const double2 *__restrict__ goth = a.bsk + (size_t)(r * 2 + (1 - r)) * M + lane;

  for (u32 sample = blockIdx.x * CPB + ct; sample < a.num_samples; sample += gridDim.x * CPB) {
    const u64 *lwe = a.lwe_in + a.in_idx[sample] * (u64)(n + 1);
    const u32 b_hat = modulus_switch(__ldg(&lwe[n]), LOGM + 2);
    const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(2 * N) + (size_t)r * N;

    // ACC_r <- LUT_r * X^(-b_hat); own[] keeps this thread's pairs in registers
    u64 own_re[32], own_im[32];
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      const int j = lane + 32 * p;
      {
        const u32 u = (j + b_hat) & (2 * N - 1);
        const u64 w = __ldg(&lut[u & (N - 1)]);
        own_re[p] = (u < (u32)N) ? w : (u64)0 - w;
      }
      {
        const u32 u = (j + M + b_hat) & (2 * N - 1);
        const u64 w = __ldg(&lut[u & (N - 1)]);
        own_im[p] = (u < (u32)N) ? w : (u64)0 - w;
      }
      acc[j] = make_ulonglong2(own_re[p], own_im[p]);
    }
    __syncwarp();

    u32 a_next = modulus_switch(__ldg(&lwe[0]), LOGM + 2);
#pragma unroll 1
    for (u32 i = 0; i < n; ++i) {
      const u32 a_hat = a_next;
      if (i + 1 < n) a_next = modulus_switch(__ldg(&lwe[i + 1]), LOGM + 2);
      if (a_hat != 0) { // X^0 - 1 = 0: the CMUX adds nothing (bootstrap.rs:287)

      // ct1 = ACC*X^a - ACC, closest_representable, signed digit (l = 1) -> doubles
      double2 x[32];
#pragma unroll
      for (int p = 0; p < 32; ++p) {
        const int j = lane + 32 * p;
        const u32 s = (u32)(j - (int)a_hat) & (2 * N - 1);
        const u32 q = s >> LOGM;
        const ulonglong2 src = acc[s & (M - 1)];
        const u64 ra = (q & 1) ? src.y : src.x;
        const u64 rb = (q & 1) ? src.x : src.y;
        // (-i)^q * (A + iB): negate re for q in {2,3}, im for q in {1,2}
        const u64 m_re = (u64)0 - (u64)(q >> 1);
        const u64 m_im = (u64)0 - (u64)(((q + 1) >> 1) & 1);
        const u64 c_re = ((ra ^ m_re) - m_re) - own_re[p];
        const u64 c_im = ((rb ^ m_im) - m_im) - own_im[p];
        x[p].x = __int2double_rn(digit_l1((u32)(c_re >> 32), base_log));
        x[p].y = __int2double_rn(digit_l1((u32)(c_im >> 32), base_log));
      }

      forward(x, buf_own, tws, lane);

      // external product: cross term to the partner warp, own term in place
      const double2 *__restrict__ g0 = gown + (size_t)i * 4 * M;
      const double2 *__restrict__ g1 = goth + (size_t)i * 4 * M;
      __syncwarp(); // every lane is done reading buf_own (transpose) before it is overwritten
#pragma unroll
      for (int p = 0; p < 32; ++p) {
        const double2 wo = __ldg(&g1[p * 32]);
        double2 cr;
        cr.x = fma(x[p].x, wo.x, -x[p].y * wo.y);
        cr.y = fma(x[p].x, wo.y, x[p].y * wo.x);
        buf_own[swz(32 * lane + p, lane)] = cr;
        const double2 w = __ldg(&g0[p * 32]);
        const double re = fma(x[p].x, w.x, -x[p].y * w.y);
        const double im = fma(x[p].x, w.y, x[p].y * w.x);
        x[p].x = re;
        x[p].y = im;
      }
      asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
#pragma unroll
      for (int p = 0; p < 32; ++p) {
        const double2 o = buf_oth[swz(32 * lane + p, lane)];
        x[p].x += o.x;
        x[p].y += o.y;
      }
      asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory"); // partner done with buf_own

      inverse(x, buf_own, tws, lane);

      // ACC += round(x) ; the new value stays in registers as next iteration's "own" (own[] is
      // dead during the transforms, so it is re-read here instead of being carried through them)
#pragma unroll
      for (int p = 0; p < 32; ++p) {
        const ulonglong2 o = acc[lane + 32 * p];
        own_re[p] = o.x + from_torus_scaled(x[p].x);
        own_im[p] = o.y + from_torus_scaled(x[p].y);
        acc[lane + 32 * p] = make_ulonglong2(own_re[p], own_im[p]);
      }
      __syncwarp();
      } // a_hat != 0
    }

    // sample extraction (glwe_sample_extraction.rs:90-165): warp 0 owns the mask polynomial,
    // warp 1 the body polynomial
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
    __syncwarp();
  }
}

} // namespace w32
