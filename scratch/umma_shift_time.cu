// Timing variant of umma_shift_test.cu: cycles per MMA (M=128, N=64, K=16) as a function of the A operand's start row.
// Experiment: can a K-major SWIZZLE_128B UMMA operand start at an arbitrary 128-byte row of a TMA-written tile
// (start address not 1024-byte aligned), and if so with which descriptor base_offset?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scratch/umma_shift_test scratch/umma_shift_test.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (!done && spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ uint64_t desc_k_sw128(uint32_t addr, uint32_t base_off) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         ((uint64_t)(base_off & 7) << 49) | (2ull << 61);
}
__device__ __forceinline__ void umma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

constexpr int NSHIFT = 20;

// mode 0: base_offset = 0; mode 1: base_offset = shift & 7; mode 2: base_offset = (8 - shift) & 7
__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                                         float* out, int mode) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // 512 rows x 128 B
  uint8_t* sB = smem + 512 * 128;     // 64 rows x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 64 * 128);
  uint64_t* mma_bar = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1); mbar_init(mma_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, 512 * 128 + 64 * 128);
    tma_load_2d(sA, &map_a, bar, 0, 0);
    tma_load_2d(sA + 256 * 128, &map_a, bar, 0, 256);
    tma_load_2d(sB, &map_b, bar, 0, 0);
  }
  mbar_wait(bar, 0);
  constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  uint32_t ph = 0;
  constexpr int REP = 1200;
  for (int sh = 0; sh < NSHIFT; ++sh) {
    const int shift = sh < 18 ? sh : (sh == 18 ? 131 : 77);
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a0 = smem_u32(sA) + shift * 128, b0 = smem_u32(sB);
      // descriptors built before the clock starts: the timed loop is MMA issue only
      uint64_t ad[12], bd[4];
      for (int i = 0; i < 12; ++i) {
        const uint32_t extra = mode == 1 ? (uint32_t)(i >> 2) * 128 : (mode == 2 ? (uint32_t)(i >> 2) * 1024 : 0u);
        ad[i] = desc_k_sw128(a0 + extra + (i & 3) * 32, 0);
      }
      for (int i = 0; i < 4; ++i) bd[i] = desc_k_sw128(b0 + i * 32, 0);
      t0 = clock64();
      // shift index reused as experiment selector: sh % 4 == 0: plain; 1: commit to a spare barrier every 36 MMAs;
      // 2: also rotate the accumulator (4 x 64 columns need 256 TMEM columns: only 64 allocated here -> 2 x 32 instead);
      // 3: accumulate flag 0 on the first MMA of every group of 36
      for (int rep = 0; rep < REP / 12; ++rep) {
        const uint32_t acc0 = (sh % 4 == 3 && rep % 3 == 0) ? 0u : 1u;
        umma(tmem, ad[0], bd[0], idesc, acc0);
#pragma unroll
        for (int i = 1; i < 12; ++i) umma(tmem, ad[i], bd[i & 3], idesc, 1u);
        if ((sh % 4 == 1 || sh % 4 == 2) && rep % 3 == 2)
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(mma_bar)) : "memory");
    }
    mbar_wait(mma_bar, ph);
    ph ^= 1;
    if (threadIdx.x == 0) {
      t1 = clock64();
      out[sh] = (float)(t1 - t0) / (REP / 12 * 12);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
  }
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int R = 1024;
  std::vector<__nv_bfloat16> hA(R * 64), hB(64 * 64);
  std::vector<float> fA(R * 64), fB(64 * 64);
  srand(1);
  for (int i = 0; i < R * 64; ++i) { float v = (rand() % 17 - 8) / 8.f; hA[i] = __float2bfloat16(v); fA[i] = __bfloat162float(hA[i]); }
  for (int i = 0; i < 64 * 64; ++i) { float v = (rand() % 13 - 6) / 8.f; hB[i] = __float2bfloat16(v); fB[i] = __bfloat162float(hB[i]); }
  __nv_bfloat16 *dA, *dB; float* dOut;
  cudaMalloc(&dA, R * 64 * 2); cudaMalloc(&dB, 64 * 64 * 2); cudaMalloc(&dOut, NSHIFT * 128 * 64 * 4);
  cudaMemcpy(dA, hA.data(), R * 64 * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), 64 * 64 * 2, cudaMemcpyHostToDevice);
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fp;
  CUtensorMap ma, mb;
  cuuint64_t dimsA[2] = {64, (cuuint64_t)R}, strA[1] = {128};
  cuuint32_t boxA[2] = {64, 256}, es[2] = {1, 1};
  CUresult r1 = enc(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dA, dimsA, strA, boxA, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  cuuint64_t dimsB[2] = {64, 64};
  cuuint32_t boxB[2] = {64, 64};
  CUresult r2 = enc(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dB, dimsB, strA, boxB, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r1 || r2) { printf("encode failed %d %d\n", (int)r1, (int)r2); return 1; }
  const size_t smem = 512 * 128 + 64 * 128 + 64 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  std::vector<float> hOut(NSHIFT * 128 * 64);
  for (int mode = 0; mode < 3; ++mode) {
    cudaMemset(dOut, 0, hOut.size() * 4);
    k<<<1, 128, smem>>>(ma, mb, dOut, mode);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mode %d: CUDA error %s\n", mode, cudaGetErrorString(e)); return 2; }
    cudaMemcpy(hOut.data(), dOut, hOut.size() * 4, cudaMemcpyDeviceToHost);
    printf("mode %d cycles/MMA:", mode);
    for (int sh = 0; sh < NSHIFT; ++sh) {
      const int shift = sh < 18 ? sh : (sh == 18 ? 131 : 77);
      printf(" s%d:%.0f", shift, hOut[sh]);
    }
    printf("\n");
  }
  return 0;
}
