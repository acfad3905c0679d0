#include <cstdint>
#include <cstdio>
__device__ __forceinline__ void tmem_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
}
__global__ void k(uint32_t *out) {
  __shared__ uint32_t tbase;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbase)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t base = tbase;
  uint32_t taddr = base + (((uint32_t)(warp & 3) * 32) << 16) + (warp >> 2) * 128;
  for (int c = 0; c < 128; c += 4) tmem_st4(taddr + c, threadIdx.x * 1000 + c, threadIdx.x * 1000 + c + 1, threadIdx.x * 1000 + c + 2, threadIdx.x * 1000 + c + 3);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  uint32_t bad = 0;
  for (int c = 0; c < 128; c += 4) {
    uint32_t a, b, cc, d;
    tmem_ld4(taddr + c, a, b, cc, d);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    bad += (a != threadIdx.x * 1000 + c) + (b != threadIdx.x * 1000 + c + 1) + (cc != threadIdx.x * 1000 + c + 2) + (d != threadIdx.x * 1000 + c + 3);
  }
  out[threadIdx.x] = bad;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(base) : "memory");
}
int main() {
  uint32_t *d; cudaMalloc(&d, 256 * 4);
  k<<<2, 256>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  uint32_t h[256]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  uint32_t tot = 0; for (int i = 0; i < 256; ++i) tot += h[i];
  printf("err=%s bad=%u\n", cudaGetErrorString(e), tot);
  return 0;
}
