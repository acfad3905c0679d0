// pbs_v4.cu — host side of the v4 fused PBS kernels (pbs_v4.cuh): which (k, N) shapes they cover,
// the BSK -> Fourier conversion into their key layout, and the launch.  Own translation unit so
// that the fifteen kernel instantiations compile in parallel with pbs.cu.
#include "pbs_common.cuh"
#include "sm100_prims.cuh"
#include "pbs_v4.cuh"
#include <atomic>

namespace {

// tuning hook (tools/ and tests only): alternative instantiations of the N = 16384 two-level kernel
std::atomic<int> g_v4_alt{0};

struct V4Variant {
  void (*one_level)(const PbsArgs);    // l = 1
  void (*multi_narrow)(const PbsArgs); // l >= 2, base_log * l <= 31 (32-bit decomposition states)
  void (*two_narrow)(const PbsArgs);   // l == 2 known at compile time (P8), 32-bit states
  void (*multi_wide)(const PbsArgs);   // l >= 2, base_log * l  > 31
  void (*convert)(double2 *, const double2 *, size_t);
  int threads, T, cpb, logm, logp;
  int cluster; // CTAs per ciphertext (k+1 in the cluster variant, else 1)
  size_t smem, smem_stash, convert_smem;
};

template <int K, int LOGN, int CPB, int RB, int NB, bool CLUSTER = false, int LOGP = 4>
V4Variant make_v4() {
  using C1 = v4::Cfg<K, LOGN, LOGP, CPB, false, RB, NB, CLUSTER>;
  using CL = v4::Cfg<K, LOGN, LOGP, CPB, true, RB, NB, CLUSTER>;
  V4Variant v;
  v.one_level = v4::pbs_kernel_v4<C1, true, 1>;
  v.multi_narrow = v4::pbs_kernel_v4<CL, true, 0>;
  v.two_narrow = v4::pbs_kernel_v4<CL, true, 2>;
  v.multi_wide = v4::pbs_kernel_v4<CL, false, 0>;
  v.convert = v4::bsk_convert_kernel_v4<C1>;
  v.threads = C1::THREADS;
  v.T = C1::T;
  v.cpb = CPB;
  v.cluster = CLUSTER ? K + 1 : 1;
  v.logm = C1::LOGM;
  v.logp = LOGP;
  v.smem = C1::SMEM;
  v.smem_stash = CL::SMEM;
  v.convert_smem = C1::BUF_BYTES;
  return v;
}

// shapes: every k = 1 polynomial size of the BASELINE configs from 1024 up to 16384, and P4's
// (k = 2, N = 1024).  At N = 16384 one polynomial buffer is 128 KiB: the two polynomials of a
// ciphertext live on the two SMs of a thread-block cluster (Cfg::CL).
bool v4_variant(u32 k, u32 N, V4Variant &v) {
  // <k, log2 N, ciphertexts per CTA, rows per ring block, ring blocks>: RB * NB of the 16 rows of a
  // GGSW level are resident; the whole level where the shared memory left by the buffers allows
  if (k == 1 && N == 1024) { v = make_v4<1, 10, 8, 8, 2>(); return true; }  // 128 + 32 KiB
  if (k == 1 && N == 2048) { v = make_v4<1, 11, 4, 8, 2>(); return true; }  // 128 + 64 KiB
  // (finer ring blocks - RB x NB = 2 x 6 at N = 4096, 1 x 6 at N = 8192 - measured 165.4 against 150.5 ms
  // and 482.2 against 407.3 ms: every block costs a barrier wait and a counted hand-back per warp)
  if (k == 1 && N == 4096) { v = make_v4<1, 12, 2, 4, 3>(); return true; }  // 128 + 96 KiB
  // N = 8192: ring blocks of 3 rows in 2 slots of 48 KiB - six blocks per CMUX (5 x 3 + 1 rows; the block
  // size need not divide the 16 rows) instead of eight blocks of 2 rows in 3 slots of 32 KiB: P7 x 4096
  // 407.2 -> 389.1 ms.  The same 2 x 48 KiB geometry measured neutral at N = 4096 (6 + 6 + 4 rows: 150.4 ms
  // either way) and at N = 16384 (5 x 6 + 2 rows per level: 581.1 against 579.5 ms), which keep 3 x 32 KiB.
  if (k == 1 && N == 8192) { v = make_v4<1, 13, 1, 3, 2>(); return true; }  // 128 + 96 KiB
  if (k == 1 && N == 16384) { // 128 + 96 KiB
    // 32 points per thread (8 warps, up to 255 registers, 5 + 5 + 3 stages: two exchanges per
    // transform): with the decomposition states parked in tensor memory it is spill-free and
    // measured 579.3 ms at P8 x 1024 against 602.3 ms for 16 points per thread (16 warps, 128
    // registers, 4 + 4 + 4 + 1 stages with the last exchange by shuffle), which stays selectable:
    // tuning hook 2 = 16 points per thread, 1 = the same with packed decomposition states.
    const int alt = g_v4_alt.load();
    if (alt == 1 || alt == 2) {
      v = make_v4<1, 14, 1, 2, 3, true>();
      if (alt == 1) v.two_narrow = v4::pbs_kernel_v4<v4::Cfg<1, 14, 4, 1, true, 2, 3, true, true>, true, 2>;
    } else {
      v = make_v4<1, 14, 1, 4, 3, true, 5>();
    }
    return true;
  }
  if (k == 2 && N == 1024) {
    // 16 points per thread = one warp per polynomial (no named barrier inside a transform), 5
    // ciphertexts per CTA, the whole 72 KiB GGSW resident as ONE ring block (120 + 72 KiB): one
    // barrier wait and one counted hand-back per CMUX, refilled while the CTA is busy with the
    // inverse transform, phase A and the next forward transform.  P4 x 16 384: 179.5 ms; as two
    // half-GGSW blocks 183.8 ms; the first configuration (8 points per thread, two warps per
    // polynomial, 3 ciphertexts per CTA, 72 + 144 KiB), which stays behind tuning hook 3: 223.5 ms.
    // pbs.cu's group kernel reads this key layout too (FOLD = 1) and takes the small batches.
    v = make_v4<2, 10, 5, 16, 1, false, 4>();
    if (g_v4_alt.load() == 3) v = make_v4<2, 10, 3, 4, 4, false, 3>();
    return true;
  }
  return false;
}

} // namespace

extern "C" void concrete_b200_set_v4_alt(int alt) { g_v4_alt.store(alt); }

bool cb_pbs_v4_shape(u32 k, u32 N) {
  return (k == 1 && (N == 1024 || N == 2048 || N == 4096 || N == 8192 || N == 16384)) ||
         (k == 2 && N == 1024);
}

// log2(points per thread) of the key layout in force (pbs.cu: which group-kernel FOLD matches it)
int cb_pbs_v4_logp(u32 k, u32 N) {
  V4Variant v;
  return v4_variant(k, N, v) ? v.logp : 0;
}

int cb_pbs_v4_cpb(u32 k, u32 N) {
  V4Variant v;
  return v4_variant(k, N, v) ? v.cpb : 0;
}

// the standard-domain key is already in `dest` (device); transform every polynomial in place
void cb_pbs_v4_convert(cudaStream_t stream, int gpu, void *dest, size_t polys, u32 k, u32 N) {
  V4Variant v;
  if (!v4_variant(k, N, v)) CB_PANIC("Cuda error (convert BSK, v4): unsupported shape");
  const double2 *tw = get_twiddles(gpu, v.logm);
  int sms = 0;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  CB_CHECK(cudaFuncSetAttribute(v.convert, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)v.convert_smem));
  const u32 grid = (u32)std::min<size_t>(polys, (size_t)sms * 8);
  v.convert<<<grid, v.T, v.convert_smem, stream>>>(reinterpret_cast<double2 *>(dest), tw, polys);
  CB_CHECK(cudaGetLastError());
  CB_LAUNCHED();
}

void cb_pbs_v4_launch(cudaStream_t stream, int gpu, u64 *lwe_out, const u64 *out_idx,
                      const u64 *lut_vector, const u64 *lut_idx, const u64 *lwe_in,
                      const u64 *in_idx, const double2 *bsk, u32 n, u32 k, u32 N, u32 base_log,
                      u32 level, u32 num_samples, u32 lut_count, u32 lut_stride) {
  V4Variant v;
  if (!v4_variant(k, N, v)) CB_PANIC("Cuda error (classical PBS, v4): unsupported shape");
  PbsArgs a;
  a.lwe_out = lwe_out;
  a.out_idx = out_idx;
  a.lut_vector = lut_vector;
  a.lut_idx = lut_idx;
  a.lwe_in = lwe_in;
  a.in_idx = in_idx;
  a.bsk = bsk;
  a.tw = get_twiddles(gpu, v.logm);
  a.global_acc = nullptr;
  a.n = n;
  a.base_log = base_log;
  a.level = level;
  a.num_samples = num_samples;
  a.lut_count = lut_count;
  a.lut_stride = lut_stride;
  void (*kern)(const PbsArgs) =
      level == 1 ? v.one_level
                 : (base_log * level <= 31 ? (level == 2 ? v.two_narrow : v.multi_narrow)
                                           : v.multi_wide);
  const size_t smem = level == 1 ? v.smem : v.smem_stash;
  if (!kern) CB_PANIC("Cuda error (classical PBS, v4): no kernel for this shape in the selected variant");
  CB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int sms = 0;
  CB_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu));
  if (v.cluster == 1) {
    const u32 grid = std::min<u32>((u32)sms, num_samples); // the kernel balances the tail
    kern<<<grid, v.threads, smem, stream>>>(a);
    CB_CHECK(cudaGetLastError());
    CB_LAUNCHED();
    return;
  }
  // cluster variant: one cluster of k+1 CTAs per ciphertext, as many clusters as are co-resident
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3((u32)v.threads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = (u32)v.cluster;
  attr.val.clusterDim.y = 1;
  attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int clusters = 0;
  {
    static std::mutex mu;
    static std::map<std::pair<int, const void *>, int> cache; // (device, kernel) -> clusters
    std::lock_guard<std::mutex> g(mu);
    auto key = std::make_pair(gpu, (const void *)kern);
    auto it = cache.find(key);
    if (it == cache.end()) {
      cfg.gridDim = dim3((u32)(sms / v.cluster) * v.cluster, 1, 1);
      int c = 0;
      if (cudaOccupancyMaxActiveClusters(&c, kern, &cfg) != cudaSuccess || c < 1) {
        (void)cudaGetLastError();
        c = sms / v.cluster / 2; // conservative
      }
      it = cache.emplace(key, c).first;
    }
    clusters = it->second;
  }
  clusters = (int)std::min<u32>((u32)clusters, num_samples);
  cfg.gridDim = dim3((u32)clusters * v.cluster, 1, 1);
  CB_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
  CB_LAUNCHED();
}
