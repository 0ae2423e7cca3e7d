// Shared device/host helpers for the tcgen05 / TMA kernels (conv_tc.cu, conv_win.cu): mbarrier, TMA loads, UMMA
// descriptors and issue, TMEM loads, and the driver entry point that encodes tensor maps.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace iit {

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
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (!done && spins > (1u << 24)) __trap();  // protocol bug: fail loudly, never hang the GPU
  }
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// K-major, SWIZZLE_128B operand descriptor: 128-byte rows, 8-row (1024 B) swizzle atoms stacked along M/N
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) /* LBO (unused) */ | ((uint64_t)(1024 >> 4) << 32) /* SBO */ |
         (1ull << 46) /* descriptor version: Blackwell */ | (2ull << 61) /* SWIZZLE_128B */;
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The MMA-issuing thread is a serial instruction stream: at 32-64 tensor-pipe cycles per instruction there is room for
// only a handful of ALU instructions per MMA. A descriptor's high word (SBO, version, swizzle) is constant and its low
// word is (LBO << 16) | (smem address >> 4), so stepping an operand by `bytes` is one 32-bit add of bytes/16
// (addresses stay below 256 KB, so the 14-bit field cannot carry out).
__device__ __forceinline__ void umma_bf16_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One lane of a converged warp. The issuing loops run warp-uniform (every lane waits on the barriers and steps the
// descriptors) and only the tcgen05 / TMA instruction itself sits under this predicate: operands then live in uniform
// registers. Inside an `if (lane == 0)` region the compiler must instead move every operand from vector to uniform
// registers with an ELECT / R2UR.BROADCAST loop per instruction (~100 issue cycles per MMA, measured with ncu sampling).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
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

// ---- CTA pairs (cta_group::2): two SMs of one cluster compute one 256-row tile; each holds its 128 rows of A and half
// of B. TMA loads of both CTAs signal the LEADER's barrier, MMA completion is multicast to both, and the peer's epilogue
// hands the accumulator back by arriving on the leader's barrier through the cluster address space. -----------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of THIS CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {   // acquire at cluster scope
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (!done && spins > (1u << 24)) __trap();
  }
}
// TMA loads whose completion bytes are credited to a barrier given by its shared::cluster address (the leader's)
__device__ __forceinline__ void tma2_load_4d(void* dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma2_load_2d(void* dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives (once the MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// ---- epilogue stores -------------------------------------------------------------------------------
// A lane of an epilogue warp holds one pixel x 32 channels (64 bytes of bf16). Stored as such, one warp store instruction
// touches 32 lines with 16 bytes each - half a sector - and costs the tiled convolutions 10-15 % of their time. Each warp
// passes the chunk through its own 32 staging rows instead and stores 64 contiguous bytes per pixel: four lanes per pixel
// (`piece` = lane & 3), pixels `sub` + 8i (`sub` = lane >> 2) of the warp's 32.
constexpr int TC_STAGE_PITCH = 80;   // bytes per staged pixel row: 64 of data, padded so that 16-byte accesses spread over the banks

// pk = this lane's 32 channels as 16 packed bf16 pairs; sdst[i] points at piece `piece` of pixel sub + 8i, channel 0 of the
// filter block (nullptr: pixel outside the tensor); c = first channel of the chunk.
// `rs` (four registers of the calling lane, used when `stats`): running per-channel sum and sum of squares of the staged values,
// for the BatchNorm that follows. The staged rows make this a column sum: lane (row parity `lane >> 4`, channel pair
// `lane & 15`) reads 16 words and keeps {sum, sum, squares, squares} of its two channels over its 16 rows IN REGISTERS, across
// chunks of the same columns and across tiles; tc_flush_stats combines parities, warps and CTAs when the filter block
// changes. (Per chunk this used to end in 4 shuffles and 4 shared-memory atomics per lane -- compare-and-swap loops that
// four warps fought over -- and two CTA barriers per tile: 40 % of a short-reduction layer's time.) Rows outside the
// tensor must be staged as zeros when statistics are taken.
// 8 bf16 values: dst[0..7] = v (plain 16-byte store) or dst[0..7] += v (ONE 16-byte reduction instruction,
// red.global.add.noftz.v4.bf16x2 = REDG.E.ADD.BF16x8: the addition happens in L2, the SM neither loads the old values nor
// waits for them). Every element receives exactly one addend per launch, so the result does not depend on any order.
__device__ __forceinline__ void st_or_add_bf16x8(__nv_bfloat16* dst, const uint4& v, bool accumulate) {
  if (accumulate) {
    asm volatile("red.global.add.noftz.v4.bf16x2 [%0], {%1, %2, %3, %4};"
                 ::"l"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  } else {
    *reinterpret_cast<uint4*>(dst) = v;
  }
}

__device__ __forceinline__ void tc_store_packed(const uint32_t (&pk)[16], uint8_t* stage, int lane, int sub, int piece,
                                                __nv_bfloat16* const (&sdst)[4], int c, float (&rs)[4], bool stats,
                                                bool accumulate = false) {
#pragma unroll
  for (int j = 0; j < 16; j += 4)
    *reinterpret_cast<uint4*>(stage + lane * TC_STAGE_PITCH + j * 4) = make_uint4(pk[j], pk[j + 1], pk[j + 2], pk[j + 3]);
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 o = *reinterpret_cast<const uint4*>(stage + (sub + 8 * i) * TC_STAGE_PITCH + piece * 16);
    if (sdst[i] != nullptr) st_or_add_bf16x8(sdst[i] + c, o, accumulate);
  }
  if (stats) {
    const uint8_t* col = stage + (lane >> 4) * TC_STAGE_PITCH + (lane & 15) * 4;       // channels 2cp, 2cp+1 of rows par, par+2, ...
    float s0 = rs[0], s1 = rs[1], q0 = rs[2], q1 = rs[3];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(col + 2 * r * TC_STAGE_PITCH);
      const float f0 = __uint_as_float(w << 16), f1 = __uint_as_float(w & 0xffff0000u);
      s0 += f0; s1 += f1; q0 = fmaf(f0, f0, q0); q1 = fmaf(f1, f1, q1);
    }
    rs[0] = s0; rs[1] = s1; rs[2] = q0; rs[3] = q1;
  }
  __syncwarp();   // the rows are rewritten by the next chunk
}
// the same from the fp32 accumulator values; `keep` = this lane's pixel lies inside the tensor
__device__ __forceinline__ void tc_store_chunk(const uint32_t (&v)[32], uint8_t* stage, int lane, int sub, int piece,
                                               __nv_bfloat16* const (&sdst)[4], int c, bool keep, float (&rs)[4], bool stats,
                                               bool accumulate = false) {
  uint32_t pk[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
    pk[j] = keep ? *reinterpret_cast<uint32_t*>(&h) : 0u;
  }
  tc_store_packed(pk, stage, lane, sub, piece, sdst, c, rs, stats, accumulate);
}
__device__ __forceinline__ void tc_store_chunk(const uint32_t (&v)[32], uint8_t* stage, int lane, int sub, int piece,
                                               __nv_bfloat16* const (&sdst)[4], int c, bool keep = true) {
  float none[4] = {0.f, 0.f, 0.f, 0.f};
  tc_store_chunk(v, stage, lane, sub, piece, sdst, c, keep, none, false, false);
}

// The running statistics of the eight epilogue warps (tc_store_packed: rs[cc] = chunk cc * 64 + half * 32 of the filter block
// at channel `co`) added to the global fp64 sums: row parities by one shuffle, the four lane quarters through s_stat
// ([4 quarters][2][BN] floats), one fp64 atomic per channel and CTA; the registers restart at zero. Every epilogue thread
// calls it (two named barriers over the 256 of them); e = its index among them.
template <int BN>
__device__ __forceinline__ void tc_flush_stats(float (&rs)[BN / 64][4], float* s_stat, int quarter, int half, int lane, int e,
                                               int co, double* bn_sum, double* bn_sq) {
#pragma unroll
  for (int cc = 0; cc < BN / 64; ++cc) {
#pragma unroll
    for (int k = 0; k < 4; ++k) rs[cc][k] += __shfl_xor_sync(0xffffffffu, rs[cc][k], 16);
    if (lane < 16) {
      const int ch = cc * 64 + half * 32 + 2 * lane;
      *reinterpret_cast<float2*>(s_stat + (quarter * 2 + 0) * BN + ch) = make_float2(rs[cc][0], rs[cc][1]);
      *reinterpret_cast<float2*>(s_stat + (quarter * 2 + 1) * BN + ch) = make_float2(rs[cc][2], rs[cc][3]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) rs[cc][k] = 0.f;
  }
  asm volatile("bar.sync 1, 256;" ::: "memory");
  for (int ch = e; ch < BN; ch += 256) {
    double s = 0.0, q = 0.0;
#pragma unroll
    for (int qd = 0; qd < 4; ++qd) { s += (double)s_stat[(qd * 2 + 0) * BN + ch]; q += (double)s_stat[(qd * 2 + 1) * BN + ch]; }
    atomicAdd(&bn_sum[co + ch], s);
    atomicAdd(&bn_sq[co + ch], q);
  }
  asm volatile("bar.sync 1, 256;" ::: "memory");   // s_stat may be rewritten
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

}  // namespace iit
