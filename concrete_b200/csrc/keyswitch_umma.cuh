// keyswitch_umma.cuh — K1b: the batched keyswitch as an exact INT8 GEMM on the 5th-generation
// tensor cores.
//
// Same arithmetic as ks_kernel (keyswitch.cu) and tfhe lwe_keyswitch.rs:138-228:
//   out[s][o] = body[s]*[o == n_out] - sum_kappa D[s][kappa] * KSK[kappa][o]     (wrapping u64)
// with kappa = i*l + j and D = the balanced signed digits (|D| <= B/2).  The product is split by
// bytes of the key word:  KSK[kappa][o] = sum_limb 2^(8 limb) * byte_limb(KSK[kappa][o]),  so
//   sum_kappa D*KSK = sum_limb 2^(8 limb) * ( sum_kappa D[s][kappa] * byte_limb[kappa][o] )
// and every inner sum is a signed-8-bit x unsigned-8-bit dot product accumulated exactly in
// s32 (|sum| <= (B/2) * 255 * K < 2^31 is checked by the host before this path is chosen).  The
// recombination is wrapping u64 arithmetic, so the result is bit-identical to the IMAD kernel and
// to the CPU reference for every input.  A real dense contraction: S x K x 8(n+1) (P2048,
// S = 16 384: 2.0e12 byte multiply-adds).  Round 1 ran it on `mma.sync.m16n8k32` (3.45 ms, kept
// under experiments/keyswitch_imma.cuh); here it is issued as
// `tcgen05.mma.cta_group::1.kind::i8` (SASS UTCIMMA) with the accumulators in tensor memory
// (1.13 ms):
//   * CTA tile 128 samples x 256 byte columns (= 32 output words x 8 byte planes), K in stages of
//     128 bytes; one elected thread issues four M128 N256 K32 MMAs per stage, operands straight
//     from shared memory through matrix descriptors (K-major, no swizzle);
//   * the two pre-passes write their operands as TILE IMAGES — for every (tile, 16-byte K chunk)
//     the rows of the tile side by side, [tile][k16][row][16 B], which is exactly the
//     "interleaved" core-matrix layout the descriptors describe (8 rows x 16 B contiguous,
//     row-group stride 128 B, K-chunk stride rows x 16 B) — so a whole stage of an operand is ONE
//     contiguous `cp.async.bulk` (TMA bulk copy, SASS UBLKCP) behind an mbarrier, no tensor map;
//   * warp-specialised persistent CTAs: warp 0 = copy producer, warp 1 = MMA issuer, warps 2-5 =
//     epilogue (one TMEM lane quadrant each): `tcgen05.ld` the 8 s32 planes of an output word,
//     recombine, subtract from the body / zero, store; two accumulator buffers (2 x 256 TMEM
//     columns) so the epilogue of tile t overlaps the MMAs of tile t + 1; a 4-stage operand ring.
// Bit-exact with ks_kernel and the CPU reference (lwe_keyswitch.rs:138-228).
#pragma once

namespace umma {

__device__ __forceinline__ u64 closest_representable(u64 x, u32 base_log, u32 level) {
  const u32 shift = 64 - level * base_log - 1;
  u64 r = x >> shift;
  r += 1;
  r &= ~(u64)1;
  return r << shift;
}
__device__ __forceinline__ u64 decompose_one_level(u64 &state, u64 mask, u32 base_log) {
  u64 res = state & mask;
  state >>= base_log;
  u64 carry = ((res - 1) | state) & res;
  carry >>= (base_log - 1);
  state += carry;
  return res - (carry << base_log);
}

constexpr int UM = 128;            // samples per tile (MMA M)
constexpr int UN = 256;            // byte columns per tile (MMA N) = 32 output words
constexpr int UKT = 128;           // K bytes per stage
constexpr int UK = 32;             // K bytes per MMA
constexpr int STAGES = 4;
constexpr int A_STAGE = UM * UKT;  // 16 KiB
constexpr int B_STAGE = UN * UKT;  // 32 KiB
constexpr int NT = 192;            // 6 warps
constexpr u32 OFF_BAR = STAGES * (A_STAGE + B_STAGE);
constexpr size_t SMEM_BYTES = OFF_BAR + 256;

struct Bars {
  unsigned long long full[STAGES], empty[STAGES], acc_full[2], acc_empty[2];
  unsigned int tmem_base;
};

__device__ __forceinline__ u32 s32addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(unsigned long long *b, u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32addr(b)), "r"(count));
}
__device__ __forceinline__ void bar_wait(unsigned long long *b, u32 parity) {
  u32 done;
  do {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(done)
        : "r"(s32addr(b)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void bar_arrive(unsigned long long *b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32addr(b)) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(unsigned long long *b, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32addr(b)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, u32 bytes, unsigned long long *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                   "r"(s32addr(dst)),
               "l"(src), "r"(bytes), "r"(s32addr(b))
               : "memory");
}
// K-major, no swizzle: 8 rows x 16 B core matrices; SBO = stride between 8-row groups,
// LBO = stride between 16-byte K chunks (both >> 4); bits 46-47 = descriptor version 1 (sm_100)
__device__ __forceinline__ u64 smem_desc(u32 addr, u32 lbo_bytes, u32 sbo_bytes) {
  return (u64)((addr >> 4) & 0x3FFF) | ((u64)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((u64)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((u64)1 << 46);
}
// instruction descriptor: D = s32 (2 << 4), A = signed 8-bit (1 << 7), B = unsigned 8-bit (0 << 10),
// both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr u32 IDESC = (2u << 4) | (1u << 7) | (0u << 10) | ((u32)(UN >> 3) << 17) | ((u32)(UM >> 4) << 24);

__device__ __forceinline__ void mma_i8(u32 tmem_d, u64 a_desc, u64 b_desc, u32 accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
      "l"(a_desc), "l"(b_desc), "r"(IDESC), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(unsigned long long *b) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   s32addr(b))
               : "memory");
}

// ---- pre-pass 1: digits as tile images ---------------------------------------------------------
// Dimg[((tile * (Kp/16) + c) * 128 + r) * 16 + b] = digit kappa = 16c + b of row tile*128 + r;
// kappa = i * level + j (j-th emitted digit of a_i, level l first).  One thread = one 16-byte
// chunk of one row; lanes walk the chunks of a row (coalesced reads of the ciphertext).
__global__ void __launch_bounds__(256)
ks_digits_img_kernel(signed char *__restrict__ Dimg, const u64 *__restrict__ in,
                     const u64 *__restrict__ in_idx, u32 n_in, u32 base_log, u32 level, u32 s0,
                     u32 count, u32 Kp) {
  const u32 row = blockIdx.x, chunks = Kp / 16, K = n_in * level;
  const u32 tile = row / UM, r = row % UM;
  const u64 *src = row < count ? in + in_idx[s0 + row] * (u64)(n_in + 1) : nullptr;
  const u64 mask = ((u64)1 << base_log) - 1;
  const u32 state_shift = 64 - base_log * level;
  for (u32 c = blockIdx.y * blockDim.x + threadIdx.x; c < chunks; c += gridDim.y * blockDim.x) {
    u32 w[4] = {0, 0, 0, 0};
    if (src && 16 * c < K) {
      const u32 k0 = 16 * c, k1 = min(k0 + 16, K);
      for (u32 i = k0 / level; i * level < k1; ++i) {
        u64 state = closest_representable(src[i], base_log, level) >> state_shift;
        for (u32 j = 0; j < level; ++j) {
          const u32 d = (u32)decompose_one_level(state, mask, base_log) & 0xFFu;
          const u32 kappa = i * level + j;
          if (kappa >= k0 && kappa < k1) w[(kappa - k0) >> 2] |= d << (8 * ((kappa - k0) & 3));
        }
      }
    }
    *reinterpret_cast<uint4 *>(Dimg + (((size_t)tile * chunks + c) * UM + r) * 16) =
        make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// ---- pre-pass 2: key byte planes as tile images -------------------------------------------------
// Bimg[((ct * (Kp/16) + c) * 256 + col % 256) * 16 + b] = byte `limb` of KSK[16c + b][o],
// col = o * 8 + limb, ct = col / 256.  grid (ceil(words/32), Kp/64)
__global__ void __launch_bounds__(256)
ks_planes_img_kernel(unsigned char *__restrict__ Bimg, const u64 *__restrict__ ksk, u32 K,
                     u32 out_size, u32 Kp, u32 words_padded) {
  __shared__ u64 tile[64][33];
  const u32 o0 = blockIdx.x * 32, k0 = blockIdx.y * 64, tid = threadIdx.x, chunks = Kp / 16;
  for (u32 idx = tid; idx < 64 * 32; idx += 256) {
    const u32 kk = idx >> 5, oo = idx & 31;
    u64 v = 0;
    if (k0 + kk < K && o0 + oo < out_size) v = ksk[(u64)(k0 + kk) * out_size + o0 + oo];
    tile[kk][oo] = v;
  }
  __syncthreads();
  for (u32 c = tid; c < 32 * 8 * 4; c += 256) {
    const u32 rowi = c >> 2, q = c & 3, oo = rowi >> 3, limb = rowi & 7;
    if (o0 + oo >= words_padded) continue;
    u32 w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      u32 acc = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b)
        acc |= (u32)((tile[16 * q + 4 * e + b][oo] >> (8 * limb)) & 0xFF) << (8 * b);
      w[e] = acc;
    }
    const u32 col = (o0 + oo) * 8 + limb, ct = col / UN;
    *reinterpret_cast<uint4 *>(Bimg + (((size_t)ct * chunks + (k0 / 16 + q)) * UN + col % UN) * 16) =
        make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// ---- the GEMM ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 1)
ks_umma_kernel(u64 *__restrict__ out, const u64 *__restrict__ out_idx, const u64 *__restrict__ in,
               const u64 *__restrict__ in_idx, const signed char *__restrict__ Dimg,
               const unsigned char *__restrict__ Bimg, u32 n_in, u32 n_out, u32 s0, u32 count,
               u32 Kp, u32 m_tiles, u32 n_tiles) {
  extern __shared__ __align__(1024) unsigned char smem[];
  Bars *bars = reinterpret_cast<Bars *>(smem + OFF_BAR);
  const u32 tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const u32 k_iters = Kp / UKT, chunks = Kp / 16, tiles = m_tiles * n_tiles;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      bar_init(&bars->full[s], 1);
      bar_init(&bars->empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      bar_init(&bars->acc_full[b], 1);
      bar_init(&bars->acc_empty[b], 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     s32addr(&bars->tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const u32 tmem = bars->tmem_base;

  if (warp == 0) {
    // ---- producer: one bulk copy per operand per stage
    if (lane == 0) {
      u32 st = 0, ph = 0;
      for (u32 t = blockIdx.x; t < tiles; t += gridDim.x) {
        const u32 mt = t / n_tiles, nt = t % n_tiles;
        const signed char *a = Dimg + (size_t)mt * chunks * UM * 16;
        const unsigned char *b = Bimg + (size_t)nt * chunks * UN * 16;
        for (u32 k = 0; k < k_iters; ++k) {
          bar_wait(&bars->empty[st], ph ^ 1);
          bar_expect_tx(&bars->full[st], A_STAGE + B_STAGE);
          bulk_load(smem + st * A_STAGE, a + (size_t)k * A_STAGE, A_STAGE, &bars->full[st]);
          bulk_load(smem + STAGES * A_STAGE + st * B_STAGE, b + (size_t)k * B_STAGE, B_STAGE,
                    &bars->full[st]);
          if (++st == STAGES) {
            st = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer
    if (lane == 0) {
      u32 st = 0, ph = 0, buf = 0, bph = 0;
      for (u32 t = blockIdx.x; t < tiles; t += gridDim.x) {
        bar_wait(&bars->acc_empty[buf], bph ^ 1); // the epilogue has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const u32 d = tmem + buf * UN;
        for (u32 k = 0; k < k_iters; ++k) {
          bar_wait(&bars->full[st], ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const u32 a0 = s32addr(smem + st * A_STAGE), b0 = s32addr(smem + STAGES * A_STAGE + st * B_STAGE);
#pragma unroll
          for (u32 kk = 0; kk < UKT / UK; ++kk) {
            // K = 32 bytes = two 16-byte chunks: chunk stride = rows * 16 B
            const u64 ad = smem_desc(a0 + kk * 2 * (UM * 16), UM * 16, 128);
            const u64 bd = smem_desc(b0 + kk * 2 * (UN * 16), UN * 16, 128);
            mma_i8(d, ad, bd, (k | kk) ? 1u : 0u);
          }
          mma_commit(&bars->empty[st]); // the stage is free once these MMAs have read it
          if (++st == STAGES) {
            st = 0;
            ph ^= 1;
          }
        }
        mma_commit(&bars->acc_full[buf]);
        if (++buf == 2) {
          buf = 0;
          bph ^= 1;
        }
      }
    }
  } else {
    // ---- epilogue: warp w owns TMEM lanes 32 (w % 4) .. +31 = rows of the tile
    const u32 quad = warp & 3, row = quad * 32 + lane, out_size = n_out + 1;
    u32 buf = 0, bph = 0;
    for (u32 t = blockIdx.x; t < tiles; t += gridDim.x) {
      const u32 mt = t / n_tiles, nt = t % n_tiles;
      const u32 s = mt * UM + row;
      const bool live = s < count;
      u64 *dst = nullptr;
      u64 body = 0;
      if (live) {
        dst = out + out_idx[s0 + s] * (u64)out_size;
        body = in[in_idx[s0 + s] * (u64)(n_in + 1) + n_in];
      }
      bar_wait(&bars->acc_full[buf], bph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const u32 taddr = tmem + ((quad * 32) << 16) + buf * UN;
#pragma unroll 1
      for (u32 w4 = 0; w4 < UN / 32; ++w4) { // 4 output words (32 columns) per TMEM load
        u32 v[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, "
            "%11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, "
            "%28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
              "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
              "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
              "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]),
              "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr + w4 * 32)
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (live) {
#pragma unroll
          for (u32 e = 0; e < 4; ++e) {
            const u32 o = nt * (UN / 8) + w4 * 4 + e;
            if (o < out_size) {
              u64 acc = 0;
#pragma unroll
              for (u32 limb = 0; limb < 8; ++limb)
                acc += (u64)(i64)(int)v[8 * e + limb] << (8 * limb); // wrapping on purpose
              dst[o] = (o == n_out ? body : (u64)0) - acc;
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) bar_arrive(&bars->acc_empty[buf]);
      if (++buf == 2) {
        buf = 0;
        bph ^= 1;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u)
                 : "memory");
}

} // namespace umma
