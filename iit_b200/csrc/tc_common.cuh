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

// ---- epilogue arithmetic on a lane's 32 fp32 accumulator columns -----------------------------------------------------
__device__ __forceinline__ void tc_scale_shift(uint32_t (&v)[32], const float* __restrict__ scale, const float* __restrict__ shift) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 a = *reinterpret_cast<const float4*>(scale + j), b = *reinterpret_cast<const float4*>(shift + j);
    v[j] = __float_as_uint(fmaf(__uint_as_float(v[j]), a.x, b.x)); v[j + 1] = __float_as_uint(fmaf(__uint_as_float(v[j + 1]), a.y, b.y));
    v[j + 2] = __float_as_uint(fmaf(__uint_as_float(v[j + 2]), a.z, b.z)); v[j + 3] = __float_as_uint(fmaf(__uint_as_float(v[j + 3]), a.w, b.w));
  }
}
__device__ __forceinline__ void tc_add_bf16_row(uint32_t (&v)[32], const __nv_bfloat16* __restrict__ row) {
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const uint4 a4 = *reinterpret_cast<const uint4*>(row + j);
    const uint32_t w4[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      v[j + 2 * h] = __float_as_uint(__uint_as_float(v[j + 2 * h]) + __uint_as_float(w4[h] << 16));
      v[j + 2 * h + 1] = __float_as_uint(__uint_as_float(v[j + 2 * h + 1]) + __uint_as_float(w4[h] & 0xffff0000u));
    }
  }
}
__device__ __forceinline__ void tc_relu(uint32_t (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(fmaxf(__uint_as_float(v[j]), 0.f));
}

// ---- epilogue stores -------------------------------------------------------------------------------
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


// The tiled kernels' output path: TMA stores from a staged tile.
// A lane of an epilogue warp holds one pixel (accumulator row m) x 32 channels. The eight warps of a CTA -- two per TMEM lane
// quarter, `half` = which 32 of a 64-channel round -- write their rows into ONE [128 rows][128 bytes] tile in the
// SWIZZLE_128B layout (16-byte unit u of row m lives at unit u ^ (m & 7): conflict-free for a quarter-warp), and one thread
// hands the tile to the TMA unit: cp.async.bulk.tensor.4d (UTMASTG) with the same (channels, BW, BH, BNI) box the loads
// use, so rows outside the tensor are clipped by the hardware, or cp.reduce.async.bulk.tensor...add (UTMAREDG.ADD) when
// the result is to be ADDED to the tensor (bf16 addition in L2, one addend per element and launch: order-independent).
// This replaced per-lane stores through warp-private staging rows (64 contiguous bytes per pixel from four lanes): that
// epilogue spent ~270 instructions per 32-column chunk on addresses, bounds and the transposition, ran at 1.3 IPC on a
// dependent chain and WAS the tile time of every layer with a short reduction (ncu: profiles/r02_ncu_single_kernels.md).
// Per round now: 16 conversions, 4 shared stores, a proxy fence and two named barriers per thread.
constexpr uint32_t TC_OUT_TILE_BYTES = 128 * 128;

__device__ __forceinline__ void tma_store_4d(const void* src, const CUtensorMap* map, int c0, int c1, int c2, int c3, bool add) {
  if (add)
    asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2, %3, %4}], [%5];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(src)) : "memory");
  else
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3, %4}], [%5];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(src)) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// PENDING = 0: every store issued by this thread has read its shared-memory tile; 1: all but the most recent one
template <int PENDING>
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(PENDING) : "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// One 64-channel round of the CTA's eight epilogue warps (all 256 threads call it, in the same order).
//   v          this lane's row m = quarter * 32 + lane, fp32 columns [half * 32, +32) of the round
//   keep       the row lies inside the tensor (rows outside are staged as zeros: the statistics read them; TMA clips them)
//   leader     exactly one of the 256 threads; it issues the store of the round when `do_store`
//   (c0..c3)   tensor coordinates of the tile's box: first channel of the round, q, p, image
//   rs, stats  running BatchNorm statistics of the staged (bf16-rounded) values: lane (par = lane >> 4, cp = lane & 15)
//              owns channels half * 32 + 2cp, +1 and 16 of the warp's rows, chosen so that the two parities read different
//              64-byte halves of the swizzled rows (no bank conflicts); sums stay in registers across rounds of the same
//              columns and across tiles (tc_flush_stats)
// OBUF = number of output tiles the caller alternates between (s_out is the one of THIS round): with two, the leader waits
// only for the store before the one it just issued, so the TMA unit reads one tile while the warps fill the other; with one,
// every round ends with the store's read latency on the critical path of all eight warps.
template <int OBUF>
__device__ __forceinline__ void tc_round_store(const uint32_t (&v)[32], bool keep, uint8_t* s_out, int quarter, int half, int lane,
                                               bool leader, bool do_store, const CUtensorMap* map, int c0, int c1, int c2, int c3,
                                               bool accumulate, float (&rs)[4], bool stats) {
  uint32_t pk[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
    pk[j] = *reinterpret_cast<uint32_t*>(&h);
  }
  if (stats && !keep) {
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = 0u;
  }
  const int m = quarter * 32 + lane;
  const uint32_t row = smem_u32(s_out) + m * 128;
  asm volatile("bar.sync 1, 256;" ::: "memory");      // the previous round's tile has been read (TMA and statistics)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t a = row + ((uint32_t)((half * 4 + j) ^ (m & 7)) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pk[4 * j]), "r"(pk[4 * j + 1]), "r"(pk[4 * j + 2]),
                 "r"(pk[4 * j + 3]) : "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the TMA unit
  asm volatile("bar.sync 1, 256;" ::: "memory");      // the tile is complete
  if (leader && do_store) tma_store_4d(s_out, map, c0, c1, c2, c3, accumulate);
  if (stats) {
    const int par = lane >> 4, cp = lane & 15;
    const uint32_t u = (uint32_t)(half * 4 + (cp >> 2)), w4 = (uint32_t)(cp & 3) * 4;
    // both channels of a word at once: add.rn.f32x2 / fma.rn.f32x2 (SASS FADD2 / FFMA2), the same roundings as the scalar forms
    uint64_t s2, q2;
    asm("mov.b64 %0, {%1, %2};" : "=l"(s2) : "f"(rs[0]), "f"(rs[1]));
    asm("mov.b64 %0, {%1, %2};" : "=l"(q2) : "f"(rs[2]), "f"(rs[3]));
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int gr = quarter * 32 + (r >> 2) * 8 + par * 4 + (r & 3);
      uint32_t w;
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(smem_u32(s_out) + gr * 128 + ((u ^ (uint32_t)(gr & 7)) << 4) + w4) : "memory");
      uint64_t f2;
      asm("mov.b64 %0, {%1, %2};" : "=l"(f2) : "r"(w << 16), "r"(w & 0xffff0000u));
      asm("add.rn.f32x2 %0, %0, %1;" : "+l"(s2) : "l"(f2));
      asm("fma.rn.f32x2 %0, %1, %1, %0;" : "+l"(q2) : "l"(f2));
    }
    asm("mov.b64 {%0, %1}, %2;" : "=f"(rs[0]), "=f"(rs[1]) : "l"(s2));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(rs[2]), "=f"(rs[3]) : "l"(q2));
  }
  if (leader) tma_store_wait_read<OBUF - 1>();         // before this thread arrives at the next round's first barrier
}

// The running statistics of the eight epilogue warps (tc_round_store: rs[cc] = chunk cc * 64 + half * 32 of the filter block
// at channel `co`) added to the global fp64 sums: row parities by one shuffle, the four lane quarters through s_stat
// ([4 quarters][2][BN] floats), one fp64 atomic per channel and CTA; the registers restart at zero. Every epilogue thread
// calls it (named barriers over the 256 of them); e = its index among them.
// s_stat may be the FIRST OUTPUT TILE of the CTA (the kernels alias it: a flush is rare and the 4-8 KB are what lets two
// output tiles fit next to the operand ring): the leader first waits until the TMA unit has read every tile it was given.
template <int BN>
__device__ __forceinline__ void tc_flush_stats(float (&rs)[BN / 64][4], float* s_stat, int quarter, int half, int lane, int e,
                                               int co, double* bn_sum, double* bn_sq, bool leader) {
  if (leader) tma_store_wait_read<0>();
  asm volatile("bar.sync 1, 256;" ::: "memory");   // no tile is being read, every warp is past its statistics reads
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
  asm volatile("bar.sync 1, 256;" ::: "memory");   // s_stat may be rewritten (as an output tile)
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
