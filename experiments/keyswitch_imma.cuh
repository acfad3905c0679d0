// (experiments/: round-1 tensor-core keyswitch on mma.sync, superseded by csrc/keyswitch_umma.cuh —
// 3.45 ms vs 1.13 ms at P2048 x 16 384; not compiled into the library)
// keyswitch_imma.cuh — K1b: the batched keyswitch as an exact INT8 tensor-core GEMM.
//
// Same arithmetic as ks_kernel (keyswitch.cu) and tfhe lwe_keyswitch.rs:138-228:
//   out[s][o] = body[s]*[o == n_out] - sum_kappa D[s][kappa] * KSK[kappa][o]     (wrapping u64)
// with kappa = i*l + j and D = the balanced signed digits (|D| <= B/2).  The product is split by
// bytes of the key word:  KSK[kappa][o] = sum_limb 2^(8 limb) * byte_limb(KSK[kappa][o]),  so
//   sum_kappa D*KSK = sum_limb 2^(8 limb) * ( sum_kappa D[s][kappa] * byte_limb[kappa][o] )
// and every inner sum is a signed-8-bit x unsigned-8-bit dot product accumulated exactly in
// s32 (|sum| <= (B/2) * 255 * K < 2^31 is checked by the host before this path is chosen).  The
// recombination is wrapping u64 arithmetic, so the result is bit-identical to the IMAD kernel and
// to the CPU reference for every input.  This is a real dense contraction: S x K x 8(n+1)
// (P2048, S = 16 384: 2.0e12 multiply-adds on bytes instead of 1.25e11 on 64-bit words, 34 ms
// on the integer pipe).
//
// Three launches per chunk of samples:
//   ks_digits_kernel  u64 ciphertext words -> s8 digits D[s][kappa]        (K contiguous)
//   ks_planes_kernel  KSK u64 [kappa][o]   -> u8 planes Bt[o*8+limb][kappa] (K contiguous)
//   ks_imma_kernel    TN GEMM, mma.sync.m16n8k32.s32.s8.u8 fed by ldmatrix from a 4-stage
//                     cp.async ring; an n8 MMA tile is exactly the 8 byte planes of ONE output
//                     word, so the epilogue recombines inside a thread quad (two shuffles).
// Both scratch buffers are padded to whole tiles and zero-filled: no guards in the main loop.
#pragma once

namespace imma {

constexpr int TM = 128;        // samples per CTA
constexpr int TN = 128;        // byte columns per CTA = 16 output words
constexpr int TK = 64;         // kappa bytes per stage
constexpr int ROWB = TK + 16;  // padded smem row (80 B): conflict-free ldmatrix, 16 B aligned
constexpr int STAGES = 4;
constexpr int NT = 256;        // 8 warps: 4 along samples x 2 along byte columns
constexpr int STAGE_BYTES = (TM + TN) * ROWB;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES; // 81 920

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

// D[(s - s0)][i*level + j] = j-th emitted digit of a_i (level l first), as s8.
// grid (rows, column blocks): the row index sits on blockIdx.x (limit 2^31 - 1), so batches of
// more than 65 535 padded rows launch too (blockIdx.y is capped at 65 535)
__global__ void ks_digits_kernel(signed char *__restrict__ D, const u64 *__restrict__ in,
                                 const u64 *__restrict__ in_idx, u32 n_in, u32 base_log,
                                 u32 level, u32 s0, u32 count, u32 Kp) {
  const u32 row = blockIdx.x; // grid.x = rows padded to whole tiles
  signed char *dst = D + (size_t)row * Kp;
  const u32 K = n_in * level;
  if (row >= count) { // padding row: zeros
    for (u32 c = (blockIdx.y * blockDim.x + threadIdx.x) * 16; c < Kp; c += gridDim.y * blockDim.x * 16)
      *reinterpret_cast<uint4 *>(dst + c) = make_uint4(0, 0, 0, 0);
    return;
  }
  if (blockIdx.y == 0)
    for (u32 c = K + threadIdx.x; c < Kp; c += blockDim.x) dst[c] = 0; // padding columns
  const u64 *src = in + in_idx[s0 + row] * (u64)(n_in + 1);
  const u64 mask = ((u64)1 << base_log) - 1;
  const u32 state_shift = 64 - base_log * level;
  for (u32 i = blockIdx.y * blockDim.x + threadIdx.x; i < n_in; i += gridDim.y * blockDim.x) {
    u64 state = closest_representable(src[i], base_log, level) >> state_shift;
    for (u32 j = 0; j < level; ++j)
      dst[i * level + j] = (signed char)(int)(u32)decompose_one_level(state, mask, base_log);
  }
}

// Bt[(o*8 + limb)][kappa] = byte `limb` of KSK[kappa][o].  grid (ceil(Np/8/32), Kp/64)
__global__ void __launch_bounds__(256)
ks_planes_kernel(unsigned char *__restrict__ Bt, const u64 *__restrict__ ksk, u32 K, u32 out_size,
                 u32 Kp, u32 words_padded /* Np / 8: padding words are written as zeros */) {
  __shared__ u64 tile[64][33];
  const u32 o0 = blockIdx.x * 32, k0 = blockIdx.y * 64, tid = threadIdx.x;
  for (u32 idx = tid; idx < 64 * 32; idx += 256) {
    const u32 kk = idx >> 5, oo = idx & 31;
    u64 v = 0;
    if (k0 + kk < K && o0 + oo < out_size) v = ksk[(u64)(k0 + kk) * out_size + o0 + oo];
    tile[kk][oo] = v;
  }
  __syncthreads();
  for (u32 c = tid; c < 32 * 8 * 4; c += 256) {
    const u32 row = c >> 2, q = c & 3, oo = row >> 3, limb = row & 7;
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
    *reinterpret_cast<uint4 *>(Bt + ((size_t)(o0 + oo) * 8 + limb) * Kp + k0 + 16 * q) =
        make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__device__ __forceinline__ void cp_async16(u32 dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void ldsm_x4(u32 addr, u32 (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_s8u8(int (&c)[4], const u32 (&a)[4], u32 b0, u32 b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, "
      "{%8, %9}, {%0, %1, %2, %3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// grid (Np / TN, rows_padded / TM).  D: [rows_padded][Kp] s8, Bt: [Np][Kp] u8, Kp % TK == 0.
__global__ void __launch_bounds__(NT, 2)
ks_imma_kernel(u64 *__restrict__ out, const u64 *__restrict__ out_idx, const u64 *__restrict__ in,
               const u64 *__restrict__ in_idx, const signed char *__restrict__ D,
               const unsigned char *__restrict__ Bt, u32 n_in, u32 n_out, u32 s0, u32 count,
               u32 Kp) {
  extern __shared__ __align__(128) unsigned char smem[];
  const u32 smem_base = (u32)__cvta_generic_to_shared(smem);
  const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 wm = warp & 3, wn = warp >> 2;
  const u32 row0 = blockIdx.y * TM, col0 = blockIdx.x * TN;
  const signed char *gA = D + (size_t)row0 * Kp;
  const unsigned char *gB = Bt + (size_t)col0 * Kp;
  const u32 ktiles = Kp / TK;

  // each thread copies 2 A chunks and 2 B chunks of 16 B per stage
  auto issue = [&](u32 kt, u32 stage) {
    const u32 sA = smem_base + stage * STAGE_BYTES, sB = sA + TM * ROWB;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const u32 c = tid + e * NT, r = c >> 2, q = c & 3;
      cp_async16(sA + r * ROWB + 16 * q, gA + (size_t)r * Kp + (size_t)kt * TK + 16 * q);
      cp_async16(sB + r * ROWB + 16 * q, gB + (size_t)r * Kp + (size_t)kt * TK + 16 * q);
    }
  };

  int acc[2][8][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mt][nt][e] = 0;

#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if ((u32)s < ktiles) issue(s, s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }

  // ldmatrix lane addressing (byte offsets inside a stage)
  const u32 a_off = (wm * 32 + (lane & 7) + 8 * ((lane >> 3) & 1)) * ROWB + 16 * (lane >> 4);
  const u32 b_off = TM * ROWB + (wn * 64 + 8 * (lane >> 4) + (lane & 7)) * ROWB +
                    16 * ((lane >> 3) & 1);

  for (u32 kt = 0; kt < ktiles; ++kt) {
    asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2) : "memory");
    __syncthreads(); // stage kt landed for every thread; stage kt-1 is free for the next issue
    if (kt + STAGES - 1 < ktiles) issue(kt + STAGES - 1, (kt + STAGES - 1) % STAGES);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const u32 sbase = smem_base + (kt % STAGES) * STAGE_BYTES;
#pragma unroll
    for (int kh = 0; kh < TK / 32; ++kh) {
      u32 a[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) ldsm_x4(sbase + a_off + mt * 16 * ROWB + kh * 32, a[mt]);
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        u32 b[4];
        ldsm_x4(sbase + b_off + np * 16 * ROWB + kh * 32, b);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          mma_s8u8(acc[mt][2 * np], a[mt], b[0], b[1]);
          mma_s8u8(acc[mt][2 * np + 1], a[mt], b[2], b[3]);
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");

  // epilogue: MMA tile nt of this warp = the 8 byte planes of output word o
  const u32 g = lane >> 2, t = lane & 3;
  const u32 out_size = n_out + 1;
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const u32 row = row0 + wm * 32 + mt * 16 + g + 8 * half;
      const bool live = row < count;
      u64 *dst = nullptr;
      u64 body = 0;
      if (live) {
        dst = out + out_idx[s0 + row] * (u64)out_size;
        body = in[in_idx[s0 + row] * (u64)(n_in + 1) + n_in];
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int c0 = acc[mt][nt][2 * half], c1 = acc[mt][nt][2 * half + 1];
        u64 p = ((u64)(long long)c0 << (16 * t)) + ((u64)(long long)c1 << (16 * t + 8));
        p += __shfl_xor_sync(0xffffffffu, p, 1);
        p += __shfl_xor_sync(0xffffffffu, p, 2);
        const u32 o = (col0 >> 3) + wn * 8 + nt;
        if (live && o < out_size && (u32)(nt & 3) == t) dst[o] = ((o == n_out) ? body : (u64)0) - p;
      }
    }
  }
}

} // namespace imma
