// Experiment: MN-major SWIZZLE_128B operands with (i) arbitrary start row, (ii) LBO (stride between 64-element
// M/N blocks) set to a small row offset so that the blocks are overlapping shifted windows of ONE tile.
//   D[mb*64 + k][nb*64 + c] = sum_{p < KP} A[a0 + mb*DA + p][k] * B[b0 + nb*DB + p][c]
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scratch/umma_mn_test scratch/umma_mn_test.cu -lcuda
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
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t addr, uint32_t lbo_bytes) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
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

struct Case { int a0, b0, DA, DB, KP; };
constexpr int NCASE = 8;
__constant__ Case cases[NCASE];

__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                                         float* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // 512 rows x 128 B
  uint8_t* sB = smem + 512 * 128;     // 512 rows x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 512 * 128);
  uint64_t* mma_bar = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1); mbar_init(mma_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, 2 * 512 * 128);
    tma_load_2d(sA, &map_a, bar, 0, 0);
    tma_load_2d(sA + 256 * 128, &map_a, bar, 0, 256);
    tma_load_2d(sB, &map_b, bar, 0, 0);
    tma_load_2d(sB + 256 * 128, &map_b, bar, 0, 256);
  }
  mbar_wait(bar, 0);
  // D=f32, A=B=bf16, both MN-major (bits 15, 16), N=192, M=128
  constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(192 >> 3) << 17) |
                             ((uint32_t)(128 >> 4) << 24);
  uint32_t ph = 0;
  for (int ci = 0; ci < NCASE; ++ci) {
    const Case c = cases[ci];
    if (threadIdx.x == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a0 = smem_u32(sA) + c.a0 * 128, b0 = smem_u32(sB) + c.b0 * 128;
      for (int kk = 0; kk < c.KP / 16; ++kk)
        umma(tmem, desc_mn_sw128(a0 + kk * 2048, c.DA * 128), desc_mn_sw128(b0 + kk * 2048, c.DB * 128), idesc, kk != 0);
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(mma_bar)) : "memory");
    }
    mbar_wait(mma_bar, ph);
    ph ^= 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int m = warp * 32 + lane;
    for (int cc = 0; cc < 192; cc += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cc, v);
      for (int j = 0; j < 32; ++j) out[((size_t)ci * 128 + m) * 192 + cc + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
  }
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int R = 512;
  std::vector<__nv_bfloat16> hA(R * 64), hB(R * 64);
  std::vector<float> fA(R * 64), fB(R * 64);
  srand(1);
  for (int i = 0; i < R * 64; ++i) { float v = (rand() % 17 - 8) / 8.f; hA[i] = __float2bfloat16(v); fA[i] = __bfloat162float(hA[i]); }
  for (int i = 0; i < R * 64; ++i) { float v = (rand() % 13 - 6) / 8.f; hB[i] = __float2bfloat16(v); fB[i] = __bfloat162float(hB[i]); }
  __nv_bfloat16 *dA, *dB; float* dOut;
  cudaMalloc(&dA, R * 64 * 2); cudaMalloc(&dB, R * 64 * 2); cudaMalloc(&dOut, NCASE * 128 * 192 * 4);
  cudaMemcpy(dA, hA.data(), R * 64 * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), R * 64 * 2, cudaMemcpyHostToDevice);
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fp;
  CUtensorMap ma, mb;
  cuuint64_t dims[2] = {64, (cuuint64_t)R}, str[1] = {128};
  cuuint32_t box[2] = {64, 256}, es[2] = {1, 1};
  CUresult r1 = enc(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dA, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUresult r2 = enc(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dB, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r1 || r2) { printf("encode failed %d %d\n", (int)r1, (int)r2); return 1; }
  // {a0, b0, DA, DB, KP}: aligned blocks first (DA = DB = 64 rows = the normal 8 KB box stride), then shifted windows
  Case hc[NCASE] = {{0, 0, 64, 64, 64},  {0, 0, 64, 64, 16}, {8, 16, 64, 64, 32}, {3, 5, 64, 64, 32},
                    {0, 0, 16, 1, 16},   {17, 0, 16, 1, 64}, {33, 2, 16, 1, 208}, {5, 9, 31, 1, 48}};
  cudaMemcpyToSymbol(cases, hc, sizeof(hc));
  const size_t smem = 2 * 512 * 128 + 64 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  std::vector<float> hOut(NCASE * 128 * 192);
  cudaMemset(dOut, 0, hOut.size() * 4);
  k<<<1, 128, smem>>>(ma, mb, dOut);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 2; }
  cudaMemcpy(hOut.data(), dOut, hOut.size() * 4, cudaMemcpyDeviceToHost);
  for (int ci = 0; ci < NCASE; ++ci) {
    const Case c = hc[ci];
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 192; ++n) {
        const int mb_ = m / 64, kk = m % 64, nb = n / 64, cc = n % 64;
        double ref = 0;
        for (int p = 0; p < c.KP; ++p) ref += (double)fA[(c.a0 + mb_ * c.DA + p) * 64 + kk] * fB[(c.b0 + nb * c.DB + p) * 64 + cc];
        double err = fabs(ref - hOut[((size_t)ci * 128 + m) * 192 + n]);
        if (err > maxerr) maxerr = err;
      }
    printf("case a0=%d b0=%d DA=%d DB=%d KP=%d: %s (max err %.3g)\n", c.a0, c.b0, c.DA, c.DB, c.KP, maxerr < 1e-2 ? "ok" : "BAD", maxerr);
  }
  return 0;
}
