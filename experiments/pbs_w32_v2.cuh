// pbs_w32_v2.cuh — second generation of the warp-per-polynomial PBS kernel (N=2048, k=1, l=1).
//
// ncu on v1 (profiles/r1_w32_v1_pbs_kernel_ncu.txt): latency-bound — 6 warps/SM at 255 registers,
// 39 % of the stall samples on the Fourier-key LDGs of the external product.  v2 changes the
// memory plan, not the arithmetic:
//   * TMEM as per-thread scratch: the thread's own accumulator pairs (64 u64 = 128 columns) live
//     in tensor memory (tcgen05.st / tcgen05.ld, 32x32b shape: lane i of a warp <-> TMEM lane
//     32*(warp%4)+i), so the shared-memory copy of the accumulator is only needed between the
//     torus rounding of CMUX i and the rotated read of CMUX i+1 and can alias the transpose
//     buffer: 32 KiB of shared memory per ciphertext instead of 64 KiB;
//   * 4 ciphertexts (8 warps) per CTA advance in lock-step over the CMUX index, and each 64 KiB
//     GGSW is brought into shared memory ONCE per CTA by TMA bulk copies (cp.async.bulk +
//     mbarrier complete_tx) issued by whichever warp releases the stage last; the load of GGSW
//     i+1 overlaps the inverse transform of CMUX i and the decomposition + forward transform of
//     CMUX i+1.  L2->SM key traffic drops 4x and the product reads the key with LDS latency.
#pragma once

namespace w32 {
using namespace sm100;
using namespace sm100;

constexpr int CPB2 = 4;
constexpr int NTHREADS2 = 64 * CPB2;
constexpr size_t V2_OFF_STAGE = 16384;                              // tws padded to 16 KiB
constexpr size_t V2_STAGE_BYTES = 4 * (size_t)M * sizeof(double2);  // one GGSW: 64 KiB
constexpr size_t V2_OFF_BUF = V2_OFF_STAGE + V2_STAGE_BYTES;
constexpr size_t V2_BUF_BYTES = (size_t)M * 16;                     // 16 KiB per warp
constexpr size_t V2_OFF_CTRL = V2_OFF_BUF + 2 * CPB2 * V2_BUF_BYTES;
constexpr size_t SMEM_BYTES2 = V2_OFF_CTRL + 32;

struct V2Ctrl {
  unsigned long long full_bar; // mbarrier: GGSW stage filled
  unsigned int released;       // warps done with the current stage
  unsigned int tmem_base;
};

__device__ __forceinline__ u32 smem_u32(const void *p) {
  return (u32)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, u32 parity) {
  u32 done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_bulk(void *dst_smem, const void *src_gmem, u32 bytes,
                                              unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// one thread: fill the stage with GGSW `i` (64 KiB contiguous in the Fourier key)
__device__ __forceinline__ void issue_stage(unsigned char *stage, const double2 *bsk, u32 i,
                                            unsigned long long *bar) {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic reads before async write
  mbar_expect_tx(bar, (u32)V2_STAGE_BYTES);
  const unsigned char *src = reinterpret_cast<const unsigned char *>(bsk + (size_t)i * 4 * M);
#pragma unroll
  for (int c = 0; c < 4; ++c)
    tma_load_bulk(stage + c * 16384, src + c * 16384, 16384, bar);
}

__device__ __forceinline__ void tmem_st16(u32 taddr, const u32 (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, "
      "%11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
      "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]),
      "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(u32 taddr, u32 (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
      "%12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_wait_st() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// release the stage; the last of the 2*n_active warps refills it with the next GGSW
__device__ __forceinline__ void release_stage(V2Ctrl *ctrl, unsigned char *stage,
                                              const double2 *bsk, int lane, u32 n_active, u32 i,
                                              u32 n, bool more_rounds) {
  __syncwarp();
  if (lane == 0) {
    __threadfence_block();
    const u32 prev = atomicAdd(&ctrl->released, 1u);
    if (prev + 1 == 2 * n_active) {
      ctrl->released = 0;
      __threadfence_block();
      if (i + 1 < n)
        issue_stage(stage, bsk, i + 1, &ctrl->full_bar);
      else if (more_rounds)
        issue_stage(stage, bsk, 0, &ctrl->full_bar);
    }
  }
}

__global__ void __launch_bounds__(NTHREADS2, 1) pbs_kernel_w32v2(const PbsArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  double2 *tws = reinterpret_cast<double2 *>(smem);
  unsigned char *stage = smem + V2_OFF_STAGE;
  V2Ctrl *ctrl = reinterpret_cast<V2Ctrl *>(smem + V2_OFF_CTRL);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ct = warp >> 1, r = warp & 1;
  // this warp's 16 KiB buffer: accumulator home between two CMUXes, transpose scratch inside one
  double2 *buf = reinterpret_cast<double2 *>(smem + V2_OFF_BUF + (size_t)warp * V2_BUF_BYTES);
  double2 *buf_oth =
      reinterpret_cast<double2 *>(smem + V2_OFF_BUF + (size_t)(warp ^ 1) * V2_BUF_BYTES);
  ulonglong2 *acc = reinterpret_cast<ulonglong2 *>(buf);
  const int bar_id = 1 + ct;
  const u32 n = a.n, base_log = a.base_log;
  const size_t out_size = (size_t)N + 1;

  load_tws(tws, a.tw, tid, NTHREADS2);
  if (tid == 0) {
    mbar_init(&ctrl->full_bar, 1);
    ctrl->released = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(
                     smem_u32(&ctrl->tmem_base))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const u32 tmem_own = ctrl->tmem_base + ((u32)((warp & 3) * 32) << 16) + (u32)(warp >> 2) * 128;

  const u32 CPBA = a.cpb_active;
  const u32 round_stride = gridDim.x * CPBA;
  const u32 first = blockIdx.x * CPBA;
  if (tid == 0 && first < a.num_samples) issue_stage(stage, a.bsk, 0, &ctrl->full_bar);
  u32 full_parity = 0;
  // de-phase the ciphertexts of the CTA so that their shared-memory and FP64 phases interleave
  for (int d = 0; d < ct; ++d) __nanosleep(a.skew_ns);

  const double2 *g_own = reinterpret_cast<const double2 *>(stage) + (size_t)(r * 2 + r) * M + lane;
  const double2 *g_oth =
      reinterpret_cast<const double2 *>(stage) + (size_t)(r * 2 + (1 - r)) * M + lane;

  for (u32 base_s = first; base_s < a.num_samples; base_s += round_stride) {
    const u32 sample = base_s + ct;
    if (sample >= a.num_samples || (u32)ct >= CPBA) break; // higher ct slots stay inactive for good
    const u32 n_active = min(CPBA, a.num_samples - base_s);
    const bool more_rounds = (base_s + round_stride) < a.num_samples;

    const u64 *lwe = a.lwe_in + a.in_idx[sample] * (u64)(n + 1);
    const u32 b_hat = modulus_switch(__ldg(&lwe[n]), LOGM + 2);
    const u64 *lut = a.lut_vector + a.lut_idx[sample] * (u64)(2 * N) + (size_t)r * N;

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
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      u32 v[16];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        v[4 * e + 0] = (u32)own_re[4 * c + e];
        v[4 * e + 1] = (u32)(own_re[4 * c + e] >> 32);
        v[4 * e + 2] = (u32)own_im[4 * c + e];
        v[4 * e + 3] = (u32)(own_im[4 * c + e] >> 32);
      }
      tmem_st16(tmem_own + 16 * c, v);
    }
    tmem_wait_st();
    __syncwarp();

    u32 a_next = modulus_switch(__ldg(&lwe[0]), LOGM + 2);
#pragma unroll 1
    for (u32 i = 0; i < n; ++i) {
      const u32 a_hat = a_next;
      if (i + 1 < n) a_next = modulus_switch(__ldg(&lwe[i + 1]), LOGM + 2);

      if (a_hat == 0) { // X^0 - 1 = 0: no CMUX (bootstrap.rs:287); still take part in the stage
        mbar_wait(&ctrl->full_bar, full_parity);
        full_parity ^= 1;
        release_stage(ctrl, stage, a.bsk, lane, n_active, i, n, more_rounds);
      } else {
        double2 x[32];
#pragma unroll
        for (int p = 0; p < 32; ++p) {
          const int j = lane + 32 * p;
          const u32 s = (u32)(j - (int)a_hat) & (2 * N - 1);
          const u32 q = s >> LOGM;
          const ulonglong2 src = acc[s & (M - 1)];
          const u64 ra = (q & 1) ? src.y : src.x;
          const u64 rb = (q & 1) ? src.x : src.y;
          const u64 m_re = (u64)0 - (u64)(q >> 1);
          const u64 m_im = (u64)0 - (u64)(((q + 1) >> 1) & 1);
          const u64 c_re = ((ra ^ m_re) - m_re) - own_re[p];
          const u64 c_im = ((rb ^ m_im) - m_im) - own_im[p];
          x[p].x = __int2double_rn(digit_l1((u32)(c_re >> 32), base_log));
          x[p].y = __int2double_rn(digit_l1((u32)(c_im >> 32), base_log));
        }
        forward(x, buf, tws, lane);

        // GGSW i is in the stage once the TMA transaction completes
        mbar_wait(&ctrl->full_bar, full_parity);
        full_parity ^= 1;

        __syncwarp(); // transpose reads of buf are done in every lane
#pragma unroll
        for (int p = 0; p < 32; ++p) {
          const double2 wo = g_oth[p * 32];
          double2 cr;
          cr.x = fma(x[p].x, wo.x, -x[p].y * wo.y);
          cr.y = fma(x[p].x, wo.y, x[p].y * wo.x);
          buf[swz(32 * lane + p, lane)] = cr;
          const double2 w = g_own[p * 32];
          const double re = fma(x[p].x, w.x, -x[p].y * w.y);
          const double im = fma(x[p].x, w.y, x[p].y * w.x);
          x[p].x = re;
          x[p].y = im;
        }
        release_stage(ctrl, stage, a.bsk, lane, n_active, i, n, more_rounds);

        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
#pragma unroll
        for (int p = 0; p < 32; ++p) {
          const double2 o = buf_oth[swz(32 * lane + p, lane)];
          x[p].x += o.x;
          x[p].y += o.y;
        }
        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory"); // partner done with buf

        inverse(x, buf, tws, lane);

        // ACC += round(x): own copy from TMEM, new value to TMEM, to the shared home (for the
        // rotated read of the next CMUX) and kept in registers as the next "own" operand
        __syncwarp(); // transpose reads of buf are done before it becomes the accumulator again
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          u32 v[16];
          tmem_ld16(tmem_own + 16 * c, v);
          tmem_wait_ld();
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int p = 4 * c + e;
            const u64 o_re = ((u64)v[4 * e + 1] << 32) | v[4 * e + 0];
            const u64 o_im = ((u64)v[4 * e + 3] << 32) | v[4 * e + 2];
            own_re[p] = o_re + from_torus_scaled(x[p].x);
            own_im[p] = o_im + from_torus_scaled(x[p].y);
            acc[lane + 32 * p] = make_ulonglong2(own_re[p], own_im[p]);
            v[4 * e + 0] = (u32)own_re[p];
            v[4 * e + 1] = (u32)(own_re[p] >> 32);
            v[4 * e + 2] = (u32)own_im[p];
            v[4 * e + 3] = (u32)(own_im[p] >> 32);
          }
          tmem_st16(tmem_own + 16 * c, v);
        }
        tmem_wait_st();
        __syncwarp();
      }
    }

    // sample extraction: warp 0 of the pair holds the mask polynomial, warp 1 the body
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

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(ctrl->tmem_base)
                 : "memory");
}

} // namespace w32
