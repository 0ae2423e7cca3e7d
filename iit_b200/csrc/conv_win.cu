// (b) "Shifted-window" convolutions on tcgen05: stride-1 convolutions whose input rows are short enough that a
// whole zero-padded image (or row band) fits in shared memory as ONE TMA box, loaded ONCE.
//
// The box lands as a list of padded pixels, one swizzled row of RB bytes (the pixel's channels) per pixel, in
// raster order with PW pixels per image row. Output pixel j (raster index in the same padded grid) needs, for
// filter tap (dh, dw), input pixel j + dh*PW + dw: a constant row offset. UMMA's K-major swizzled operand may start
// at ANY row of such a tile - the 128B/32B swizzle is a function of the absolute shared-memory address
// (scratch/umma_shift_test*.cu measured it on B200) - so every tap's A operand is the same tile addressed
// `doff[tap]` rows further on: no im2col, no per-tap reload. The implicit GEMM
//     D[128 padded positions, 64 filters] += A_tile[rows j0+128t+doff[tap] ...][RB/2] * W_tap[64][RB/2]^T
// computes the halo columns too (they are discarded in the epilogue): 196/256 of the MMA work is useful for a
// 14x14 image, in exchange for 7x less L2->SMEM traffic than nine shifted loads (the previous bound of these layers).
//
// Used for: the four 3x3 / 64->64 convolutions of layer1 and their dgrads (RB = 128), and the 7x7/2 stem as a
// 4x4 stride-1 convolution over the 2x2 space-to-depth image (12 -> 16 channels, RB = 32).
//
// Persistent CTAs; all filter taps resident in shared memory; STAGES image tiles in flight; 4 TMEM accumulators of
// 64 columns so that the epilogue (TMEM -> bf16 NHWC + BatchNorm statistics) overlaps the next tiles' MMAs.
#include <stdlib.h>

#include "tc_common.cuh"

namespace iit {

constexpr int WIN_THREADS = 192;
constexpr int WIN_CONV_THREADS = 320;   // conv_win_kernel: TMA warp, MMA warp, eight epilogue warps
constexpr int WIN_BN = 64;        // filters per tile (all of them: these layers have 64)
constexpr int WIN_NACC = 4;       // TMEM accumulators in the ring
constexpr int WIN_MAX_TAPS = 16;

struct WinParams {
  int N, OH, OW;            // images, output size
  int PW;                   // pixels per padded image row (= shared-memory rows per image row)
  int bands, band_h;        // row bands per image, output rows per band
  int pad_top, pad_left;    // position of output (0,0) inside the box
  int y_step, y0, x0;       // TMA start row of band b = y0 + b*y_step; start column x0
  int mtiles;               // 128-position M tiles per band
  int ntaps;
  int stage_bytes;          // shared-memory bytes reserved per tile (multiple of 1024, includes the overrun rows)
  int tile_bytes;           // bytes TMA delivers per tile
  int16_t doff[WIN_MAX_TAPS];  // tap -> row offset dh*PW + dw (relative to the output position)
  int8_t wk[WIN_MAX_TAPS];     // tap -> block column of the filter matrix
  int accumulate;              // != 0: results are added to `out` (one 16-byte reduction per piece) instead of stored
  // folded eval-mode BatchNorm (iit_conv2d_fprop_tc_affine): y = act(acc * scale[k] + shift[k] (+ residual)); scale == null: off
  const float* scale;
  const float* shift;
  const __nv_bfloat16* residual;   // laid out like `out`, or null
  int relu;
};

// this lane's pixel (n, oh, ww), filters c .. c + 31 of the accumulator row, before rounding
__device__ __forceinline__ void win_affine(uint32_t (&v)[32], const WinParams& p, int n, int oh, int ww, int c, bool ok) {
  tc_scale_shift(v, p.scale + c, p.shift + c);
  if (p.residual != nullptr && ok) tc_add_bf16_row(v, p.residual + (((int64_t)n * p.OH + oh) * p.OW + ww) * WIN_BN + c);
  if (p.relu) tc_relu(v);
}

// K-major operand descriptor for rows of RB bytes: SWIZZLE_128B (RB = 128) or SWIZZLE_32B (RB = 32)
template <int RB>
__device__ __forceinline__ uint64_t win_desc(uint32_t smem_addr) {
  constexpr uint64_t swz = RB == 128 ? 2ull : 6ull;
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)((8 * RB) >> 4) << 32) | (1ull << 46) | (swz << 61);
}

// OBUF = 0: the epilogue stores from registers, one pixel row per lane. OBUF = 1 / 2: it stores by TMA from that many
// [128 positions][64 filters] tiles (tc_round_store) -- possible when an M tile is a rectangle of the output, i.e. the padded
// row length divides 128 and the image is one band: then the positions of the padding columns and of rows below the image
// fall outside the tensor map and the hardware clips them.
template <int RB, int NTAPS, int STAGES, int OBUF>
__global__ void __launch_bounds__(WIN_CONV_THREADS) conv_win_kernel(const __grid_constant__ CUtensorMap map_a,
                                                               const __grid_constant__ CUtensorMap map_b,
                                                               const __grid_constant__ CUtensorMap map_o,
                                                               __nv_bfloat16* __restrict__ out, WinParams p,
                                                               double* __restrict__ bn_sum, double* __restrict__ bn_sq) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int BN = WIN_BN;
  constexpr uint32_t B_TAP = BN * RB;                 // one tap's filter block
  constexpr int KSTEPS = RB / 32;                     // UMMA K = 16 bf16 = 32 bytes
  uint8_t* sB = smem;                                 // [ntaps][64][RB]
  uint8_t* sA = smem + ((p.ntaps * B_TAP + 1023) & ~1023u);
  uint8_t* s_out = sA + (size_t)STAGES * p.stage_bytes;   // [OBUF][128][128 bytes], SWIZZLE_128B (1024-byte aligned: stage_bytes is)
  uint64_t* full = reinterpret_cast<uint64_t*>(s_out + OBUF * TC_OUT_TILE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;               // [NACC]
  uint64_t* tmem_empty = tmem_full + WIN_NACC;        // [NACC]
  uint64_t* b_full = tmem_empty + WIN_NACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_full + 1);
  float* s_stat = reinterpret_cast<float*>(tmem_slot + 2);   // [2][BN]; with TMA stores [4 lane quarters][2][BN] (tc_flush_stats)
  if (OBUF == 0 && bn_sum != nullptr && threadIdx.x < BN) { s_stat[threadIdx.x] = 0.f; s_stat[threadIdx.x + BN] = 0.f; }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.N * p.bands;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < WIN_NACC; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 8); }
    mbar_init(b_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(WIN_NACC * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    {
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        // all filter taps, once
        mbar_expect_tx(b_full, p.ntaps * B_TAP);
        for (int t = 0; t < p.ntaps; ++t) tma_load_2d(sB + t * B_TAP, &map_b, b_full, p.wk[t] * (RB / 2), 0);
      }
      __syncwarp();
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n = tile / p.bands, band = tile - n * p.bands;
        mbar_wait(&empty[stage], phase ^ 1);
        if (issuer) {
          mbar_expect_tx(&full[stage], p.tile_bytes);
          tma_load_4d(sA + (size_t)stage * p.stage_bytes, &map_a, &full[stage], 0, p.x0, p.y0 + band * p.y_step, n);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    {   // warp-uniform loop, the elected lane issues (tc_common.cuh: elect_one)
      const bool issuer = elect_one();
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      mbar_wait(b_full, 0);
      int stage = 0;
      uint32_t phase = 0, it = 0;
      const int j0 = p.pad_top * p.PW + p.pad_left;
      // descriptor words: high word constant, low word = (1 << 16) | (address >> 4); per-tap steps in registers
      const uint64_t proto = win_desc<RB>(0);
      const uint32_t hi = (uint32_t)(proto >> 32), lo_flags = (uint32_t)proto;
      const uint32_t b_lo = lo_flags + (smem_u32(sB) >> 4);
      uint32_t a_step[NTAPS];
#pragma unroll
      for (int tap = 0; tap < NTAPS; ++tap) a_step[tap] = (uint32_t)((j0 + p.doff[tap]) * (RB / 16));
      const uint32_t a_lo0 = lo_flags + (smem_u32(sA) >> 4), stage_step = (uint32_t)p.stage_bytes >> 4;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&full[stage], phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_tile = a_lo0 + stage * stage_step;
        for (int t = 0; t < p.mtiles; ++t, ++it) {
          const uint32_t acc = it % WIN_NACC, acc_phase = (it / WIN_NACC) & 1;
          mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t tmem_d = tmem_base + acc * BN;
          const uint32_t a_m = a_tile + t * (128 * RB / 16);
          if (issuer) {
#pragma unroll
            for (int tap = 0; tap < NTAPS; ++tap) {
#pragma unroll
              for (int k = 0; k < KSTEPS; ++k)
                umma_bf16_lohi(tmem_d, a_m + a_step[tap] + 2 * k, hi, b_lo + tap * (B_TAP / 16) + 2 * k, hi, idesc,
                               (uint32_t)((tap | k) != 0));
            }
            umma_commit(&tmem_full[acc]);
          }
          __syncwarp();
        }
        if (issuer) umma_commit(&empty[stage]);   // the tile may be overwritten once every MMA above has read it
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // Epilogue, eight warps: warp w reads TMEM lanes [32 (w % 4), +32), so two warps share each lane quarter and each takes
    // one 32-column half of the 64 filters (four warps doing both halves, with 128 statistics registers per lane: layer1
    // 80 -> 77.5 us; the stem did not move -- and routing the stores through staging rows as conv_tc.cu does made both slower:
    // 224 and 80.6 us).
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int c = half * 32;              // this warp's filters
    const int m = quarter * 32 + lane;
    const int e = threadIdx.x - 64;       // 0..255
    if constexpr (OBUF > 0) {
      // TMA-store epilogue. ncu on the register-store form (profiles/r02_ncu_single_kernels.md): its 16-byte stores to 32
      // different lines per instruction are one L1 wavefront each -- 1024 per tile, 822 MB of half-filled sectors for the
      // stem's 411 MB -- on the SAME data pipe the MMAs read their operands through: stem 54 % store + 44 % tensor-core
      // wavefronts, layer1 69 % tensor-core. A TMA store reads the tile in 128 full-width wavefronts.
      const bool leader = e == 0;
      float rstat[1][4] = {{0.f, 0.f, 0.f, 0.f}};
      uint32_t it = 0, round = 0;
      const int j0 = p.pad_top * p.PW + p.pad_left;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n = tile;                                    // one band per image
        for (int t = 0; t < p.mtiles; ++t, ++it) {
          const int jj = j0 + 128 * t + m;
          const int hh = jj / p.PW - p.pad_top, ww = jj % p.PW - p.pad_left;
          const bool ok = ww >= 0 && ww < p.OW && hh >= 0 && hh < p.OH;
          const uint32_t acc = it % WIN_NACC, acc_phase = (it / WIN_NACC) & 1;
          mbar_wait(&tmem_full[acc], acc_phase);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + c), v);
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty[acc])) : "memory");
          if (p.scale != nullptr) win_affine(v, p, n, hh, ww, c, ok);
          // the tile's 128 positions are rows hh0 .. of the image, all PW columns from output column 0
          const int hh0 = (j0 + 128 * t) / p.PW - p.pad_top;
          tc_round_store<(OBUF > 0 ? OBUF : 1)>(v, ok, s_out + (round++ % (OBUF > 0 ? OBUF : 1)) * TC_OUT_TILE_BYTES, quarter, half, lane,
                                                leader, true, &map_o, 0, 0, hh0, n, p.accumulate != 0, rstat[0], bn_sum != nullptr);
        }
      }
      if (leader) tma_store_wait_all();
      if (bn_sum != nullptr) tc_flush_stats<BN>(rstat, s_stat, quarter, half, lane, e, 0, bn_sum, bn_sq, leader);
    } else {
    float st_s[32], st_q[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) { st_s[j] = 0.f; st_q[j] = 0.f; }
    uint32_t it = 0;
    const int j0 = p.pad_top * p.PW + p.pad_left;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n = tile / p.bands, band = tile - n * p.bands;
      for (int t = 0; t < p.mtiles; ++t, ++it) {
        const int jj = j0 + 128 * t + m;
        const int hh = jj / p.PW - p.pad_top, ww = jj % p.PW - p.pad_left;
        const int oh = band * p.band_h + hh;
        const bool ok = ww >= 0 && ww < p.OW && hh >= 0 && hh < p.band_h && oh < p.OH;
        __nv_bfloat16* dst = out + (((int64_t)n * p.OH + oh) * p.OW + ww) * BN + c;
        const uint32_t acc = it % WIN_NACC, acc_phase = (it / WIN_NACC) & 1;
        mbar_wait(&tmem_full[acc], acc_phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + c), v);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty[acc])) : "memory");
        if (p.scale != nullptr) win_affine(v, p, n, oh, ww, c, ok);
        // bf16 pairs as stored; the BatchNorm statistics are taken from these rounded values. Each lane keeps running sums
        // for its own accumulator row across all the tiles of this CTA (2 x 32 fp32 registers); the cross-lane reduction
        // happens once, after the tile loop.
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          pk[j] = *reinterpret_cast<uint32_t*>(&h);
        }
        if (ok) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            st_or_add_bf16x8(dst + 2 * j, make_uint4(pk[j], pk[j + 1], pk[j + 2], pk[j + 3]), p.accumulate != 0);
          if (bn_sum != nullptr) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float f0 = __uint_as_float(pk[j] << 16), f1 = __uint_as_float(pk[j] & 0xffff0000u);
              st_s[2 * j] += f0; st_q[2 * j] = fmaf(f0, f0, st_q[2 * j]);
              st_s[2 * j + 1] += f1; st_q[2 * j + 1] = fmaf(f1, f1, st_q[2 * j + 1]);
            }
          }
        }
      }
    }
    if (bn_sum != nullptr) {
      // once per CTA: butterfly over the 32 lanes (lane l ends with the warp total of channel c + l), shared-memory
      // atomics over the four warps of a half, one fp64 atomic per channel
#pragma unroll
      for (int h = 16; h >= 1; h >>= 1) {
        const bool up = (lane & h) != 0;
#pragma unroll
        for (int i = 0; i < h; ++i) {
          float sa = up ? st_s[i] : st_s[i + h], ka = up ? st_s[i + h] : st_s[i];
          float sb = up ? st_q[i] : st_q[i + h], kb = up ? st_q[i + h] : st_q[i];
          st_s[i] = ka + __shfl_xor_sync(0xffffffffu, sa, h);
          st_q[i] = kb + __shfl_xor_sync(0xffffffffu, sb, h);
        }
      }
      atomicAdd(&s_stat[c + lane], st_s[0]);
      atomicAdd(&s_stat[BN + c + lane], st_q[0]);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (e < BN) { atomicAdd(&bn_sum[e], (double)s_stat[e]); atomicAdd(&bn_sq[e], (double)s_stat[BN + e]); }
    }
    }  // OBUF == 0
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(WIN_NACC * BN));
  }
}

// ---- host side -----------------------------------------------------------------------------------
constexpr size_t WIN_SMEM_MAX = 227 * 1024;

static size_t win_smem_bytes(const WinParams& p, int RB, int stages, int obuf = 0) {
  const size_t b = ((size_t)p.ntaps * WIN_BN * RB + 1023) & ~(size_t)1023;
  return b + (size_t)stages * p.stage_bytes + (size_t)obuf * TC_OUT_TILE_BYTES + (2 * stages + 2 * WIN_NACC + 1) * 8 + 16 +
         (obuf ? 8 : 2) * WIN_BN * 4 + 1024;
}

// Fills the geometry for an image of IH x IW input pixels whose TMA box is (PW x box_h) pixels per band.
static bool win_plan(WinParams* p, int RB, int OH, int OW, int PW, int halo_h, int pad_top, int pad_left, int max_doff,
                     int min_doff) {
  // band height: the whole image when its box fits 256 rows of TMA box height and the shared-memory budget
  for (int band_h = OH; band_h >= 1; --band_h) {
    const int box_h = band_h + halo_h;
    if (box_h > 256 || PW > 256) continue;
    const int j0 = pad_top * PW + pad_left;
    const int span = (band_h - 1) * PW + OW;                  // padded positions from the first to the last output
    const int mtiles = (span + 127) / 128;
    if (j0 + min_doff < 0) return false;
    const int rows_touched = j0 + 128 * mtiles + max_doff;    // highest row an MMA reads, + 1
    const int rows = rows_touched > box_h * PW ? rows_touched : box_h * PW;
    const int stage_bytes = (rows * RB + 1023) & ~1023;
    p->OH = OH; p->OW = OW; p->PW = PW; p->band_h = band_h; p->bands = (OH + band_h - 1) / band_h;
    p->pad_top = pad_top; p->pad_left = pad_left; p->mtiles = mtiles; p->stage_bytes = stage_bytes;
    p->tile_bytes = box_h * PW * RB;
    if (win_smem_bytes(*p, RB, 2) <= WIN_SMEM_MAX) return true;
  }
  return false;
}

static int win_encode_a(CUtensorMap* m, const void* ptr, int C, int W, int H, int N, int box_w, int box_h, int RB, const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)(RB / 2), (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

static int win_encode_b(CUtensorMap* m, const void* wmat, int Kg, int RB, const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  cuuint64_t dims[2] = {(cuuint64_t)Kg, (cuuint64_t)WIN_BN};
  cuuint64_t strides[1] = {(cuuint64_t)Kg * 2};
  cuuint32_t box[2] = {(cuuint32_t)(RB / 2), (cuuint32_t)WIN_BN};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(wmat), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, RB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: filter tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

// the output as the TMA-store epilogue sees it: [N][OH][OW][64] in boxes of 64 filters x PW columns x 128 / PW rows
static int win_encode_out(CUtensorMap* m, void* out, const WinParams& p, const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  cuuint64_t dims[4] = {(cuuint64_t)WIN_BN, (cuuint64_t)p.OW, (cuuint64_t)p.OH, (cuuint64_t)p.N};
  cuuint64_t strides[3] = {(cuuint64_t)WIN_BN * 2, (cuuint64_t)p.OW * WIN_BN * 2, (cuuint64_t)p.OH * p.OW * WIN_BN * 2};
  cuuint32_t box[4] = {(cuuint32_t)WIN_BN, (cuuint32_t)p.PW, (cuuint32_t)(128 / p.PW), 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: output tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

static int win_tma_default() {
  const char* e = getenv("IIT_WIN_TMA_STORE");
  return (e != nullptr && e[0] == '0') ? 0 : 1;
}
static int g_win_tma_out = win_tma_default();     // iit_conv_win_tma_store(): 0 = always the register-store epilogue (A/B measurements)

template <int RB, int NTAPS>
static int win_launch(const CUtensorMap& ma, const CUtensorMap& mb, void* out, const WinParams& p, cudaStream_t st,
                      double* bn_sum, double* bn_sq, const char* name) {
  int grid = kNumSMs;
  if (grid > p.N * p.bands) grid = p.N * p.bands;
  const int dev = current_device_slot();
  // TMA stores when an M tile is a rectangle of the output (see the kernel): two output tiles and at least three operand
  // stages if that fits, else one tile
  const bool rect = g_win_tma_out && 128 % p.PW == 0 && p.bands == 1;
  if (rect) {
    int obuf = 0, stages = 0;
    if (win_smem_bytes(p, RB, 4, 2) <= WIN_SMEM_MAX) { obuf = 2; stages = 4; }
    else if (win_smem_bytes(p, RB, 3, 2) <= WIN_SMEM_MAX) { obuf = 2; stages = 3; }
    else if (win_smem_bytes(p, RB, 4, 1) <= WIN_SMEM_MAX) { obuf = 1; stages = 4; }
    else if (win_smem_bytes(p, RB, 3, 1) <= WIN_SMEM_MAX) { obuf = 1; stages = 3; }
    if (obuf) {
      CUtensorMap mo;
      if (win_encode_out(&mo, out, p, name)) return 1;
      const size_t smem = win_smem_bytes(p, RB, stages, obuf);
      static bool attr_t[kMaxDevices][5][3] = {};
#define IIT_WIN_LAUNCH_T(S, O)                                                                                           \
  if (stages == S && obuf == O) {                                                                                        \
    if (!attr_t[dev][S][O]) {                                                                                            \
      IIT_CUDA(cudaFuncSetAttribute(conv_win_kernel<RB, NTAPS, S, O>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WIN_SMEM_MAX)); \
      attr_t[dev][S][O] = true;                                                                                          \
    }                                                                                                                    \
    conv_win_kernel<RB, NTAPS, S, O><<<grid, WIN_CONV_THREADS, smem, st>>>(ma, mb, mo, (__nv_bfloat16*)out, p, bn_sum, bn_sq);       \
  }
      IIT_WIN_LAUNCH_T(4, 2)
      IIT_WIN_LAUNCH_T(3, 2)
      IIT_WIN_LAUNCH_T(4, 1)
      IIT_WIN_LAUNCH_T(3, 1)
#undef IIT_WIN_LAUNCH_T
      count_variant(RB == 128 ? KV_WIN_3X3 : KV_WIN_STEM);
      IIT_LAUNCH_CHECK(name);
      return 0;
    }
  }
  // as many tiles in flight as the shared-memory budget allows, at most 4
  int stages = 4;
  while (stages > 2 && win_smem_bytes(p, RB, stages) > WIN_SMEM_MAX) --stages;
  const size_t smem = win_smem_bytes(p, RB, stages);
  IIT_REQUIRE(smem <= WIN_SMEM_MAX, "%s: tile does not fit shared memory", name);
  static bool attr_set[kMaxDevices][5] = {};
#define IIT_WIN_LAUNCH(S)                                                                                              \
  case S:                                                                                                              \
    if (!attr_set[dev][S]) {                                                                                           \
      IIT_CUDA(cudaFuncSetAttribute(conv_win_kernel<RB, NTAPS, S, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WIN_SMEM_MAX)); \
      attr_set[dev][S] = true;                                                                                         \
    }                                                                                                                  \
    conv_win_kernel<RB, NTAPS, S, 0><<<grid, WIN_CONV_THREADS, smem, st>>>(ma, mb, ma, (__nv_bfloat16*)out, p, bn_sum, bn_sq);       \
    count_variant(RB == 128 ? KV_WIN_3X3 : KV_WIN_STEM);                                                               \
    break;
  switch (stages) {
    IIT_WIN_LAUNCH(2)
    IIT_WIN_LAUNCH(3)
    IIT_WIN_LAUNCH(4)
  }
#undef IIT_WIN_LAUNCH
  IIT_LAUNCH_CHECK(name);
  return 0;
}

// 3x3 / stride 1 / pad 1, 64 -> 64 channels. `flip` selects the dgrad tap table (taps mirrored, filter given as CRSK).
bool conv_win_3x3_supported(const iit_conv_t* d) {
  if (!(d->r == 3 && d->s == 3 && d->pad == 1 && d->stride == 1 && d->c == 64 && d->k == 64)) return false;
  WinParams p{};
  p.ntaps = 9;
  return win_plan(&p, 128, d->p, d->q, d->w + 2, 2, 1, 1, d->w + 3, -(d->w + 3));
}

int conv_win_3x3(const void* in, const void* wmat, void* out, const iit_conv_t* d, bool flip, cudaStream_t st, double* bn_sum,
                 double* bn_sq, const char* name, bool accumulate, const float* scale, const float* shift, const void* residual,
                 int relu) {
  WinParams p{};
  p.N = d->n; p.ntaps = 9;
  p.accumulate = accumulate ? 1 : 0;
  p.scale = scale; p.shift = shift; p.residual = (const __nv_bfloat16*)residual; p.relu = relu;
  const int PW = d->w + 2;
  IIT_REQUIRE(win_plan(&p, 128, d->p, d->q, PW, 2, 1, 1, PW + 1, -(PW + 1)), "%s: image does not fit the shifted-window kernel", name);
  p.x0 = -1; p.y0 = -1; p.y_step = p.band_h;
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) {
      const int t = r * 3 + s;
      // fprop reads x[h + r - 1, w + s - 1]; dgrad reads dy[h + 1 - r, w + 1 - s]
      const int dh = flip ? 1 - r : r - 1, dw = flip ? 1 - s : s - 1;
      p.doff[t] = (int16_t)(dh * PW + dw);
      p.wk[t] = (int8_t)t;
    }
  CUtensorMap ma, mb;
  if (win_encode_a(&ma, in, 64, d->w, d->h, d->n, PW, p.band_h + 2, 128, name)) return 1;
  if (win_encode_b(&mb, wmat, 9 * 64, 128, name)) return 1;
  return win_launch<128, 9>(ma, mb, out, p, st, bn_sum, bn_sq, name);
}

// ------------------------------------------------------------------------------------------------------------------
// Weight gradient of the same 3x3 / 64 -> 64 layers with shifted windows:
//     dW[k, r, s, c] = sum over padded positions j of  dy[j, k] * x[j + (r-1)*PW + (s-1), c]
// dy and x are loaded once per image as zero-padded tiles in the same raster (the halo of dy is zero, so halo positions
// contribute nothing). Pixels are the GEMM K dimension, both operands MN-major (a pixel's 64 channels are one 128-byte
// row). The 64-element M / N blocks of an MN-major operand are LBO bytes apart - and LBO may be ONE ROW (128 B) or one
// image row (PW*128 B), making the blocks overlapping shifted windows of the same tile (scratch/umma_mn_test.cu):
//     MMA 1 (M=128, N=192): A = [dy | dy shifted by PW rows], B = [x at +PW-1 | +PW | +PW+1]  -> taps r=2 (rows 0-63) and
//                                                                                           r=1 (rows 64-127), s=0,1,2
//     MMA 2 (M=128, N=192): A = [dy | (ignored)],            B = [x at -PW-1 | -PW | -PW+1]  -> taps r=0 (rows 0-63)
// Two MMAs per 16 pixels instead of nine loads + nine MMAs. The accumulators (2 x 192 TMEM columns) persist over all the
// images of a CTA; each CTA writes one partial dW, reduced in fixed order by tc_splitk_reduce (conv_tc.cu).
struct WgradWinParams {
  int N, PW, box_h;
  int ks1, ks2;           // 16-row K steps of MMA 1 / MMA 2
  int tile_rows;          // rows TMA delivers per tile
  int alloc_rows;         // rows reserved per tile (the extra ones are zeroed once and never written)
};

__device__ __forceinline__ uint64_t win_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

template <int STAGES>
__global__ void __launch_bounds__(WIN_THREADS) wgrad_win_kernel(const __grid_constant__ CUtensorMap map_dy,
                                                                const __grid_constant__ CUtensorMap map_x,
                                                                float* __restrict__ part, WgradWinParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t tile_alloc = (uint32_t)p.alloc_rows * 128;      // multiple of 1024
  const uint32_t stage_bytes = 2 * tile_alloc;                   // [dy tile][x tile]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)STAGES * stage_bytes);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // rows past the TMA box are read by the last K steps against zero gradient rows: they must be finite -> zero
  {
    const uint32_t tail = (uint32_t)(p.alloc_rows - p.tile_rows) * 128 / 16;   // uint4 per tile
    for (uint32_t i = threadIdx.x; i < tail * 2 * STAGES; i += blockDim.x) {
      const uint32_t tile = i / tail, off = i - tile * tail;
      reinterpret_cast<uint4*>(smem + (size_t)tile * tile_alloc + (size_t)p.tile_rows * 128)[off] = make_uint4(0, 0, 0, 0);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor-core reads
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const bool has_work = (int)blockIdx.x < p.N;

  if (warp == 0) {
    {   // warp-uniform loops, the elected lane issues (tc_common.cuh: elect_one)
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_dy)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int n = blockIdx.x; n < p.N; n += gridDim.x) {
        uint8_t* st = smem + (size_t)stage * stage_bytes;
        mbar_wait(&empty[stage], phase ^ 1);
        if (issuer) {
          mbar_expect_tx(&full[stage], 2u * p.tile_rows * 128);
          tma_load_4d(st, &map_dy, &full[stage], 0, -1, -1, n);
          tma_load_4d(st + tile_alloc, &map_x, &full[stage], 0, -1, -1, n);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (has_work) {
      const bool issuer = elect_one();
      // D=f32, A=B=bf16, both MN-major, N=192, M=128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(192 >> 3) << 17) |
                                 ((uint32_t)(128 >> 4) << 24);
      const uint32_t lbo_a = (uint32_t)p.PW * 128, lbo_b = 128;
      const uint64_t pa = win_desc_mn(0, lbo_a), pb = win_desc_mn(0, lbo_b);
      const uint32_t a_hi = (uint32_t)(pa >> 32), a_flags = (uint32_t)pa, b_hi = (uint32_t)(pb >> 32), b_flags = (uint32_t)pb;
      const uint32_t row = 128 / 16;                       // descriptor address units per tile row
      int stage = 0;
      uint32_t phase = 0, first = 1;
      for (int n = blockIdx.x; n < p.N; n += gridDim.x) {
        mbar_wait(&full[stage], phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t dy0 = smem_u32(smem + (size_t)stage * stage_bytes) >> 4, x0 = dy0 + (tile_alloc >> 4);
        // MMA 1: dy rows [0, 16*ks1) and [PW, PW+16*ks1); x rows from PW-1, PW, PW+1
        const uint32_t a1 = a_flags + dy0, b1 = b_flags + x0 + (uint32_t)(p.PW - 1) * row;
        // MMA 2: dy rows from PW+1; x rows from 0, 1, 2
        const uint32_t a2 = a_flags + dy0 + (uint32_t)(p.PW + 1) * row, b2 = b_flags + x0;
        if (issuer) {
          for (int kk = 0; kk < p.ks1; ++kk)
            umma_bf16_lohi(tmem_base, a1 + kk * 16 * row, a_hi, b1 + kk * 16 * row, b_hi, idesc, (uint32_t)(!first || kk != 0));
          for (int kk = 0; kk < p.ks2; ++kk)
            umma_bf16_lohi(tmem_base + 192, a2 + kk * 16 * row, a_hi, b2 + kk * 16 * row, b_hi, idesc, (uint32_t)(!first || kk != 0));
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        first = 0;
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (issuer) umma_commit(tmem_full);
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3;
    const int L = quarter * 32 + lane;          // TMEM lane
    const int k = L & 63, mb = L >> 6;
    float* base = part + ((int64_t)blockIdx.x * 64 + k) * 9 * 64;
    if (has_work) {
      mbar_wait(tmem_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
      if (which == 1 && mb == 1) break;        // rows 64-127 of the second accumulator are not a tap
      const int r = which == 0 ? (mb == 0 ? 2 : 1) : 0;
#pragma unroll 1
      for (int chunk = 0; chunk < 6; ++chunk) {
        uint32_t v[32];
        if (has_work) {
          tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(which * 192 + chunk * 32), v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        float* dst = base + (r * 3 + (chunk >> 1)) * 64 + (chunk & 1) * 32;
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<uint4*>(dst + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

static bool wgrad_win_plan(const iit_conv_t* d, WgradWinParams* p, int* grid) {
  if (!(d->r == 3 && d->s == 3 && d->pad == 1 && d->stride == 1 && d->c == 64 && d->k == 64)) return false;
  const int PW = d->w + 2, box_h = d->h + 2;
  if (PW > 256 || box_h > 256 || PW * box_h > 256) return false;      // whole padded image as one tile
  p->N = d->n; p->PW = PW; p->box_h = box_h;
  p->tile_rows = PW * box_h;
  p->ks1 = ((d->h + 1) * PW + 15) / 16;
  p->ks2 = (d->h * PW - 2 + 15) / 16;
  int need = PW + 16 * p->ks1;                                         // dy, MMA 1, second block
  if (PW + 1 + 16 * p->ks1 > need) need = PW + 1 + 16 * p->ks1;        // x, MMA 1
  if (2 * PW + 1 + 16 * p->ks2 > need) need = 2 * PW + 1 + 16 * p->ks2;  // dy, MMA 2, (ignored) second block
  if (need < p->tile_rows) need = p->tile_rows;
  p->alloc_rows = (need + 7) & ~7;
  *grid = d->n < kNumSMs ? d->n : kNumSMs;
  return true;
}

bool wgrad_win_supported(const iit_conv_t* d) {
  WgradWinParams p;
  int grid;
  return wgrad_win_plan(d, &p, &grid);
}

size_t wgrad_win_scratch_bytes(const iit_conv_t* d) {
  WgradWinParams p;
  int grid;
  if (!wgrad_win_plan(d, &p, &grid)) return 0;
  return (size_t)grid * 64 * 9 * 64 * sizeof(float);
}

// writes grid partial sums [split][k][r][s][c] into `part`; returns the number of splits through *splits
int wgrad_win(const void* dy, const void* x, float* part, const iit_conv_t* d, int* splits, cudaStream_t st, const char* name) {
  WgradWinParams p;
  int grid;
  IIT_REQUIRE(wgrad_win_plan(d, &p, &grid), "%s: shape does not fit the shifted-window weight-gradient kernel", name);
  constexpr int STAGES = 3;
  const size_t smem = (size_t)STAGES * 2 * p.alloc_rows * 128 + (2 * STAGES + 1) * 8 + 16 + 1024;
  IIT_REQUIRE(smem <= WIN_SMEM_MAX, "%s: tiles do not fit shared memory", name);
  CUtensorMap mdy, mx;
  if (win_encode_a(&mdy, dy, 64, d->q, d->p, d->n, p.PW, p.box_h, 128, name)) return 1;
  if (win_encode_a(&mx, x, 64, d->w, d->h, d->n, p.PW, p.box_h, 128, name)) return 1;
  static bool attr_set[kMaxDevices] = {};
  const int dev = current_device_slot();
  if (!attr_set[dev]) {
    IIT_CUDA(cudaFuncSetAttribute(wgrad_win_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WIN_SMEM_MAX));
    attr_set[dev] = true;
  }
  wgrad_win_kernel<STAGES><<<grid, WIN_THREADS, smem, st>>>(mdy, mx, part, p);
  count_variant(KV_WGRAD_WIN);
  IIT_LAUNCH_CHECK(name);
  *splits = grid;
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Weight gradient of the stem in its space-to-depth form:
//     dW2[k, R, S, c] = sum over output pixels (p, q) of  dy[p, q, k] * s2d[p + R, q + S, c]        (R, S < 4, c < 16)
// Same construction as above on the s2d raster (PW pixels per row; dy is loaded with PW - OW zero columns through TMA's
// out-of-bounds fill). An image is processed in bands of BH output rows; the dy band sits between two zeroed regions of
// shared memory so that both 64-row M blocks (dy, and dy shifted by one raster row) sweep every real row:
//     MMA 1 (M=128, N=64): A = [dy | dy shifted by PW rows], B = [s2d at +PW+S, S=0..3 (32-byte rows, LBO = 1 row)] -> R = 1, 0
//     MMA 2 (M=128, N=64): same A,                            B = [s2d at +3PW+S]                                  -> R = 3, 2
struct StemWgradParams {
  int N, OH, OW, PW, SH;   // images, output size, raster width (= s2d width), s2d height
  int BH, bands;           // output rows per band
  int ks;                  // 16-row K steps per band
  int dy_alloc, x_alloc;   // bytes reserved per dy / s2d tile (multiples of 1024)
  int dy_rows, x_rows;     // rows TMA delivers
};

constexpr int STEM_LEAD = 32;   // zeroed rows in front of each tile (TMA destination stays 1024-byte aligned)

template <int STAGES>
__global__ void __launch_bounds__(WIN_THREADS) stem_wgrad_win_kernel(const __grid_constant__ CUtensorMap map_dy,
                                                                     const __grid_constant__ CUtensorMap map_x,
                                                                     float* __restrict__ part, StemWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t stage_bytes = (uint32_t)(p.dy_alloc + p.x_alloc);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)STAGES * stage_bytes);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.N * p.bands;

  // everything TMA does not overwrite must be zero (dy) / finite (s2d): clear all stages once
  for (uint32_t i = threadIdx.x; i < (uint32_t)STAGES * stage_bytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const bool has_work = (int)blockIdx.x < total;

  if (warp == 0) {
    {   // warp-uniform loops, the elected lane issues (tc_common.cuh: elect_one)
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_dy)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int n = t / p.bands, band = t - n * p.bands;
        uint8_t* st = smem + (size_t)stage * stage_bytes;
        mbar_wait(&empty[stage], phase ^ 1);
        if (issuer) {
          mbar_expect_tx(&full[stage], (uint32_t)p.dy_rows * 128 + (uint32_t)p.x_rows * 32);
          tma_load_4d(st + STEM_LEAD * 128, &map_dy, &full[stage], 0, 0, band * p.BH, n);
          tma_load_4d(st + p.dy_alloc + STEM_LEAD * 32, &map_x, &full[stage], 0, 0, band * p.BH, n);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (has_work) {
      const bool issuer = elect_one();
      // D=f32, A=B=bf16, both MN-major, N=64, M=128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) |
                                 ((uint32_t)(128 >> 4) << 24);
      const uint64_t pa = win_desc_mn(0, (uint32_t)p.PW * 128);
      // s2d rows are 32 bytes: SWIZZLE_32B, 8-row groups 256 bytes apart, 16-channel N blocks one row (32 bytes) apart
      const uint64_t pb = (uint64_t)(32 >> 4) << 16 | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
      const uint32_t a_hi = (uint32_t)(pa >> 32), a_flags = (uint32_t)pa, b_hi = (uint32_t)(pb >> 32), b_flags = (uint32_t)pb;
      int stage = 0;
      uint32_t phase = 0, first = 1;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        mbar_wait(&full[stage], phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t dy0 = smem_u32(smem + (size_t)stage * stage_bytes) >> 4, x0 = dy0 + ((uint32_t)p.dy_alloc >> 4);
        // both tiles hold raster position j in tile row STEM_LEAD + j: for tap offset o the s2d row is the dy row + o.
        // M block 0 starts PW rows before the first real row (zeros), block 1 (LBO = PW rows) exactly on it.
        const uint32_t a0 = (uint32_t)(STEM_LEAD - p.PW);
        const uint32_t a = a_flags + dy0 + a0 * 8;
        const uint32_t b1 = b_flags + x0 + (a0 + p.PW) * 2, b2 = b_flags + x0 + (a0 + 3 * p.PW) * 2;
        if (issuer) {
          for (int kk = 0; kk < p.ks; ++kk) {
            const uint32_t acc = (uint32_t)(!first || kk != 0);
            umma_bf16_lohi(tmem_base, a + kk * 128, a_hi, b1 + kk * 32, b_hi, idesc, acc);
            umma_bf16_lohi(tmem_base + 64, a + kk * 128, a_hi, b2 + kk * 32, b_hi, idesc, acc);
          }
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        first = 0;
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (issuer) umma_commit(tmem_full);
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3;
    const int L = quarter * 32 + lane;
    const int k = L & 63, mb = L >> 6;
    float* base = part + ((int64_t)blockIdx.x * 64 + k) * 256;      // [k][R][S][c]
    if (has_work) {
      mbar_wait(tmem_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
#pragma unroll 1
    for (int chunk = 0; chunk < 4; ++chunk) {      // 32 columns = S pair {0,1} / {2,3} of accumulator chunk / 2
      uint32_t v[32];
      if (has_work) {
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32), v);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0u;
      }
      const int R = (chunk >> 1) == 0 ? (mb == 0 ? 1 : 0) : (mb == 0 ? 3 : 2);
      float* dst = base + R * 64 + (chunk & 1) * 32;
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<uint4*>(dst + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128));
  }
}

// sum of the per-CTA partials [split][64][4][4][16] -> dw [64,3,7,7] (given strides): inverse of stem_filter_s2d
__global__ void stem_wgrad_finish_kernel(const float* __restrict__ part, int splits, float* __restrict__ dw, int64_t sk, int64_t sc,
                                         int64_t sr, int64_t ss) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 256) return;
  const int ch = i & 15, S = (i >> 4) & 3, R = (i >> 6) & 3, k = i >> 8;
  if (ch >= 12) return;
  const int ab = ch / 3, c = ch - ab * 3, a = ab >> 1, b = ab & 1;
  const int r = 2 * R + a, s = 2 * S + b;
  if (r >= 7 || s >= 7) return;
  float acc = 0.f;
  for (int z = 0; z < splits; ++z) acc += part[(int64_t)z * 64 * 256 + i];   // fixed order: deterministic
  dw[k * sk + c * sc + r * sr + s * ss] = acc;
}

static bool stem_wgrad_plan(int n, int sh, int sw, StemWgradParams* p, int* grid, size_t* smem, int stages) {
  const int OH = sh - 3, OW = sw - 3, PW = sw;
  if (OH < 1 || OW < 1 || PW > STEM_LEAD) return false;      // one raster row of zeros must fit in front of the dy band
  const int a0 = STEM_LEAD - PW;                             // first tile row of M block 0 (block 1 starts at STEM_LEAD)
  for (int BH = OH; BH >= 1; --BH) {
    if (BH + 3 > 256) continue;
    const int ks = ((BH + 1) * PW + 15) / 16;
    const int dy_rows = BH * PW, x_rows = (BH + 3) * PW;
    int dy_alloc_rows = STEM_LEAD + 16 * ks;                 // block 1 reads tile rows [STEM_LEAD, STEM_LEAD + 16 ks)
    if (dy_alloc_rows < STEM_LEAD + dy_rows) dy_alloc_rows = STEM_LEAD + dy_rows;
    int x_alloc_rows = a0 + 3 * PW + 3 + 16 * ks;            // MMA 2, S = 3, last K step
    if (x_alloc_rows < STEM_LEAD + x_rows) x_alloc_rows = STEM_LEAD + x_rows;
    p->dy_alloc = (dy_alloc_rows * 128 + 1023) & ~1023;
    p->x_alloc = (x_alloc_rows * 32 + 1023) & ~1023;
    const size_t need = (size_t)stages * (p->dy_alloc + p->x_alloc) + (2 * stages + 1) * 8 + 16 + 1024;
    if (need > WIN_SMEM_MAX) continue;
    p->N = n; p->OH = OH; p->OW = OW; p->PW = PW; p->SH = sh; p->BH = BH; p->bands = (OH + BH - 1) / BH; p->ks = ks;
    p->dy_rows = dy_rows; p->x_rows = x_rows;
    const int64_t total = (int64_t)n * p->bands;
    *grid = total < kNumSMs ? (int)total : kNumSMs;
    *smem = need;
    return true;
  }
  return false;
}

}  // namespace iit

using namespace iit;

extern "C" size_t iit_stem_wgrad_s2d_scratch_bytes(int n, int sh, int sw) {
  StemWgradParams p;
  int grid; size_t smem;
  if (!stem_wgrad_plan(n, sh, sw, &p, &grid, &smem, 3)) return 0;
  return (size_t)grid * 64 * 256 * sizeof(float);
}

extern "C" int iit_stem_wgrad_s2d(const void* dy, const void* s2d, float* dw, const int64_t* dw_strides_kcrs, int n, int sh, int sw,
                                  void* scratch, size_t scratch_bytes, void* stream) {
  const char* name = "iit_stem_wgrad_s2d";
  IIT_REQUIRE(dy && s2d && dw && dw_strides_kcrs && scratch, "%s: null pointer", name);
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(s2d) | reinterpret_cast<uintptr_t>(scratch)) & 15) == 0,
              "%s: dy, s2d and scratch must be 16-byte aligned", name);
  // one raster row of zeros has to fit in front of the gradient tile
  IIT_REQUIRE(sw <= STEM_LEAD, "%s: s2d width %d exceeds %d (images wider than %d pixels take the im2col path)", name, sw, STEM_LEAD,
              2 * STEM_LEAD - 6);
  constexpr int STAGES = 3;
  StemWgradParams p;
  int grid; size_t smem;
  IIT_REQUIRE(stem_wgrad_plan(n, sh, sw, &p, &grid, &smem, STAGES), "%s: image %dx%d does not fit", name, sh, sw);
  const size_t need = (size_t)grid * 64 * 256 * sizeof(float);
  IIT_REQUIRE(scratch_bytes >= need, "%s: scratch too small: need %zu bytes, got %zu", name, need, scratch_bytes);
  CUtensorMap mdy, mx;
  // dy [n, OH, OW, 64] read on the PW-wide raster: the box is wider than the tensor, the excess columns are zero-filled
  if (win_encode_a(&mdy, dy, 64, p.OW, p.OH, n, p.PW, p.BH, 128, name)) return 1;
  if (win_encode_a(&mx, s2d, 16, sw, sh, n, p.PW, p.BH + 3, 32, name)) return 1;
  static bool attr_set[kMaxDevices] = {};
  const int dev = current_device_slot();
  if (!attr_set[dev]) {
    IIT_CUDA(cudaFuncSetAttribute(stem_wgrad_win_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WIN_SMEM_MAX));
    attr_set[dev] = true;
  }
  cudaStream_t st = as_stream(stream);
  stem_wgrad_win_kernel<STAGES><<<grid, WIN_THREADS, smem, st>>>(mdy, mx, (float*)scratch, p);
  count_variant(KV_WGRAD_WIN_STEM);
  IIT_LAUNCH_CHECK(name);
  stem_wgrad_finish_kernel<<<64, 256, 0, st>>>((const float*)scratch, grid, dw, dw_strides_kcrs[0], dw_strides_kcrs[1],
                                               dw_strides_kcrs[2], dw_strides_kcrs[3]);
  IIT_LAUNCH_CHECK("iit_stem_wgrad_s2d(finish)");
  return 0;
}

// Stem: y[n,p,q,k] = sum_{R,S<4} sum_{c<16} s2d[n, p+R, q+S, c] * w2[k, R, S, c], where s2d is the zero-padded image
// after a 2x2 space-to-depth ([N, (H+6)/2, (W+6)/2, 16] bf16, 12 real channels) and w2 the 7x7 filter regrouped
// accordingly ([64, 4, 4, 16] bf16): exactly the 7x7 / stride-2 / pad-3 convolution.
extern "C" int iit_conv_win_tma_store(int on) {
  const int before = g_win_tma_out;
  g_win_tma_out = on ? 1 : 0;
  return before;
}

extern "C" int iit_stem_conv_s2d(const void* s2d, const void* w2, void* y, int n, int sh, int sw, double* bn_sum, double* bn_sqsum,
                                 void* stream) {
  const char* name = "iit_stem_conv_s2d";
  IIT_REQUIRE(s2d && w2 && y && n >= 1 && sh >= 4 && sw >= 4, "%s: bad arguments", name);
  IIT_REQUIRE((bn_sum == nullptr) == (bn_sqsum == nullptr), "%s: bn_sum and bn_sqsum go together", name);
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(s2d) | reinterpret_cast<uintptr_t>(w2) | reinterpret_cast<uintptr_t>(y)) & 15) == 0,
              "%s: pointers must be 16-byte aligned", name);
  WinParams p{};
  p.N = n; p.ntaps = 16;
  // padded row length: 32 for the 31-wide image of the 56-pixel input (TMA zero-fills the extra column; the M-tile count
  // stays 7) so that an M tile is a 4 x 32 rectangle of the output and the epilogue can store by TMA
  const int OH = sh - 3, OW = sw - 3;
  int PW = sw;
  if (sw > 16 && sw < 32 && ((OH - 1) * 32 + OW + 127) / 128 <= ((OH - 1) * sw + OW + 127) / 128) PW = 32;
  IIT_REQUIRE(win_plan(&p, 32, OH, OW, PW, 3, 0, 0, 3 * PW + 3, 0), "%s: image %dx%d does not fit the shifted-window kernel", name, sh, sw);
  p.x0 = 0; p.y0 = 0; p.y_step = p.band_h;
  for (int R = 0; R < 4; ++R)
    for (int S = 0; S < 4; ++S) {
      p.doff[R * 4 + S] = (int16_t)(R * PW + S);
      p.wk[R * 4 + S] = (int8_t)(R * 4 + S);
    }
  CUtensorMap ma, mb;
  if (win_encode_a(&ma, s2d, 16, sw, sh, n, PW, p.band_h + 3, 32, name)) return 1;
  if (win_encode_b(&mb, w2, 16 * 16, 32, name)) return 1;
  return win_launch<32, 16>(ma, mb, y, p, as_stream(stream), bn_sum, bn_sqsum, name);
}
