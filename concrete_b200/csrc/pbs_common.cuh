// pbs_common.cuh — what every PBS translation unit shares: integer helpers (restating the
// reference's decomposer / modulus switch / torus rounding), the kernel argument block and the
// per-device twiddle tables.  Everything has internal linkage (anonymous namespace): each
// translation unit owns its copy of the __constant__ twiddle bank and fills it itself.
#pragma once
#include "common.cuh"
#include "fft.cuh"
#include "../../include/concrete_b200.h"
#include <algorithm>
#include <cmath>
#include <map>
#include <mutex>
#include <type_traits>
#include <vector>

namespace {

// ------------------------------------------------------------------------------------------
// Twiddle tables: tw[2^s + b] = exp(i*pi*(1 + 4*bitrev_s(b)) / 2^(s+2)), one per (device, LOGM)
// ------------------------------------------------------------------------------------------
// optional per-TU hook: called with the host table of LOGM = 10 (pbs.cu: w32 constant bank)
void (*g_tw_hook_logm10)(const double2 *host_tw) = nullptr;
std::mutex g_tw_mutex;
std::map<std::pair<int, int>, double2 *> g_tw_cache;

u32 bitrev(u32 x, int bits) {
  u32 r = 0;
  for (int i = 0; i < bits; ++i) r = (r << 1) | ((x >> i) & 1);
  return r;
}

const double2 *get_twiddles(int gpu, int logm) {
  std::lock_guard<std::mutex> g(g_tw_mutex);
  auto key = std::make_pair(gpu, logm);
  auto it = g_tw_cache.find(key);
  if (it != g_tw_cache.end()) return it->second;
  const size_t M = (size_t)1 << logm;
  std::vector<double2> h(M);
  h[0] = make_double2(1.0, 0.0);
  const long double pi = 3.14159265358979323846264338327950288L;
  for (int s = 0; s < logm; ++s)
    for (u32 b = 0; b < (1u << s); ++b) {
      long double ang = pi * (long double)(1 + 4 * (uint64_t)bitrev(b, s)) /
                        (long double)((uint64_t)1 << (s + 2));
      h[((size_t)1 << s) + b] = make_double2((double)cosl(ang), (double)sinl(ang));
    }
  double2 *d = nullptr;
  CB_CHECK(cudaMalloc(&d, M * sizeof(double2)));
  CB_CHECK(cudaMemcpy(d, h.data(), M * sizeof(double2), cudaMemcpyHostToDevice));
  if (logm == 10 && g_tw_hook_logm10) g_tw_hook_logm10(h.data()); // current device == gpu
  { // generic constant table: entries [0, min(M, 256)) are the same for every size
    static std::map<int, size_t> filled; // per device: how many entries are valid
    const size_t want = std::min<size_t>(M, (size_t)cbfft::CONST_TW);
    if (filled[gpu] < want) {
      CB_CHECK(cudaMemcpyToSymbol(cbfft::c_twg, h.data(), want * sizeof(double2)));
      filled[gpu] = want;
    }
  }
  g_tw_cache[key] = d;
  return d;
}

// ------------------------------------------------------------------------------------------
// Integer helpers (restate decomposer.rs:23-47, iter.rs:121-133, fft_impl/common.rs:17-25,
// commons/math/torus/mod.rs:72-78)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 closest_representable(u64 x, u32 shift_m1) {
  u64 r = x >> shift_m1;
  r += 1;
  r &= ~(u64)1;
  return r << shift_m1;
}
__device__ __forceinline__ u64 decompose_one_level(u64 &state, u64 mask, u32 base_log) {
  u64 res = state & mask;
  state >>= base_log;
  u64 carry = ((res - 1) | state) & res;
  carry >>= (base_log - 1);
  state += carry;
  return res - (carry << base_log);
}
__device__ __forceinline__ u32 modulus_switch(u64 x, u32 log_modulus) {
  return (u32)((x + ((u64)1 << (63 - log_modulus))) >> (64 - log_modulus));
}
__device__ __forceinline__ u64 from_torus(double x) {
  double fr = x - rint(x);
  return (u64)__double2ll_rn(fr * 18446744073709551616.0);
}


// signed digit of one level on a state of type S (iter.rs:121-133).  S = u32 is exact whenever
// base_log * level <= 31 (the state has base_log * level bits) and halves the registers the
// decomposition states hold across the transforms of a multi-level CMUX.
template <class S> __device__ __forceinline__ S grp_decompose(S &state, S mask, u32 base_log) {
  S res = state & mask;
  state >>= base_log;
  S carry = ((res - 1) | state) & res;
  carry >>= (base_log - 1);
  state += carry;
  return res - (carry << base_log);
}

struct PbsArgs {
  u64 *lwe_out;
  const u64 *out_idx;
  const u64 *lut_vector;
  const u64 *lut_idx;
  const u64 *lwe_in;
  const u64 *in_idx;
  const double2 *bsk;
  const double2 *tw;
  ulonglong2 *global_acc; // [gridDim.x][(K+1)*M] pairs, only when ACC_GLOBAL
  u32 n, base_log, level, num_samples, lut_count, lut_stride;
};

} // namespace
