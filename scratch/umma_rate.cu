// Cycles per tcgen05.mma (bf16, M = 128, K = 16) by operand layout and N: K-major vs MN-major SWIZZLE_128B operands.
// Question: do the MN-major ("transposed") operands the weight-gradient kernels use run at the same rate as K-major ones?
// One CTA per SM (grid = argv[1], default 1), one thread issues REP back-to-back MMAs on zero-filled shared memory and
// waits for the commit; cycles = clock64 difference / REP.
// argv[2] = 1: while the MMAs run, a second warp streams bulk copies (global, L2-resident -> shared memory, 16 KB each,
// 8 in flight) into a region the MMAs do not read: do TMA writes and MMA operand reads share shared-memory bandwidth?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scratch/umma_rate scratch/umma_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (!done && spins > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ void umma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo_bytes) {   // SWIZZLE_128B, SBO = 1024
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

constexpr int NCFG = 8;
__global__ void __launch_bounds__(128) k(float* out, const uint8_t* src, int copy) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 192 * 1024);
  uint64_t* cbar = bar + 1;                       // [8] copy barriers
  volatile int* stop = reinterpret_cast<volatile int*>(bar + 9);
  unsigned long long* copied = reinterpret_cast<unsigned long long*>(bar + 10);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 11);
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&cbar[i], 1);
    *stop = 0; *copied = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  uint32_t ph = 0;
  uint32_t cph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  constexpr int REP = 2048;
  for (int cfg = 0; cfg < NCFG; ++cfg) {
    const int mn = cfg >> 2;                       // 0: K-major A and B; 1: MN-major A and B
    const int N = 64 << (cfg & 3) > 256 ? 192 : 64 << (cfg & 3);   // 64, 128, 256, 192
    if (copy && threadIdx.x == 32) {
      // stream 16 KB bulk copies into smem [64 KB, 192 KB) until the MMA thread says stop
      unsigned long long n = 0;
      for (int i = 0; i < 8; ++i) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&cbar[i])), "r"(16384) : "memory");
        asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + 64 * 1024 + i * 16384)), "l"(src + ((blockIdx.x * 8 + i) & 255) * 16384), "r"(16384),
                       "r"(smem_u32(&cbar[i])) : "memory");
      }
      const long long c0 = clock64();
      for (int i = 0; !*stop; i = (i + 1) & 7) {
        mbar_wait(&cbar[i], cph[i]); cph[i] ^= 1; ++n;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&cbar[i])), "r"(16384) : "memory");
        asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + 64 * 1024 + i * 16384)), "l"(src + ((blockIdx.x * 8 + i + n) & 255) * 16384), "r"(16384),
                       "r"(smem_u32(&cbar[i])) : "memory");
      }
      const long long c1 = clock64();
      for (int i = 0; i < 8; ++i) { mbar_wait(&cbar[i], cph[i]); cph[i] ^= 1; }      // drain
      if (blockIdx.x == 0) out[NCFG + cfg] = (float)(n * 16384) / (float)(c1 - c0);
      *stop = 0;
    }
    if (threadIdx.x == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (mn ? (1u << 15) | (1u << 16) : 0u) |
                             ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 16 * 1024;
      uint64_t ad[4], bd[4];
      for (int i = 0; i < 4; ++i) {
        // K-major: 128-byte rows, K step = 32 bytes inside the row. MN-major: rows are K (pixels), 16 rows = 2048 bytes per
        // K step; 64-element M/N blocks 8192 bytes apart (LBO)
        ad[i] = mn ? desc(a0 + i * 2048, 8192) : desc(a0 + i * 32, 16);
        bd[i] = mn ? desc(b0 + i * 2048, 8192) : desc(b0 + i * 32, 16);
      }
      const long long t0 = clock64();
      for (int rep = 0; rep < REP / 4; ++rep) {
#pragma unroll
        for (int i = 0; i < 4; ++i) umma(tmem, ad[i], bd[i], idesc, 1u);
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
      mbar_wait(bar, ph);
      const long long t1 = clock64();
      if (blockIdx.x == 0) out[cfg] = (float)(t1 - t0) / REP;
      if (copy) { *stop = 1; while (*stop) {} }
    }
    ph ^= 1;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
}

int main(int argc, char** argv) {
  const int grid = argc > 1 ? atoi(argv[1]) : 1;
  const int copy = argc > 2 ? atoi(argv[2]) : 0;
  float* out;
  uint8_t* src;
  cudaMalloc(&out, 2 * NCFG * sizeof(float));
  cudaMemset(out, 0, 2 * NCFG * sizeof(float));
  cudaMalloc(&src, 256 * 16384);
  cudaMemset(src, 0, 256 * 16384);
  const int smem = 192 * 1024 + 1024 + 128;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k<<<grid, 128, smem>>>(out, src, copy);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
  float h[2 * NCFG];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  const int Ns[4] = {64, 128, 256, 192};
  for (int c = 0; c < NCFG; ++c)
    printf("grid %3d  %s  M=128 N=%3d K=16: %6.1f cycles per MMA (ideal %d at 4096 MAC/clk)  concurrent bulk copy %5.1f B/clk\n", grid,
           c >> 2 ? "MN-major" : "K-major ", Ns[c & 3], h[c], 128 * Ns[c & 3] * 16 / 4096, h[NCFG + c]);
  return 0;
}
