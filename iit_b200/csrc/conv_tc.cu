// (b) ResNet-18 3x3 stride-1 convolutions (fprop and dgrad) on the 5th-generation tensor cores:
// TMA-fed, warp-specialised implicit GEMM with tcgen05.mma and a TMEM accumulator (bf16 in, fp32 acc).
//
//   D[128 pixels, BN filters] = sum over 9 taps, sum over C/64 chunks:  A_tap[128, 64] * W_tap[BN, 64]^T
//
// * A tile = a box of BW x BH x BNI output pixels (BW*BH*BNI = 128) of BNI images. For tap (r,s) the
//   operand is the SAME box of the NHWC input shifted by (r-pad, s-pad): one 4-D TMA load
//   (cp.async.bulk.tensor.4d, SWIZZLE_128B) whose out-of-bounds rows/columns - the padding halo - are
//   zero-filled by the TMA unit. No im2col buffer, no index arithmetic in the SM.
// * B tile = BN filters x 64 reduction elements of the KRSC (fprop) / CRSK (dgrad) filter matrix: one
//   2-D TMA load. Both land in the canonical K-major SWIZZLE_128B layout that the UMMA shared-memory
//   descriptors address directly.
// * Warp 0: TMA producer (one elected lane). Warp 1: TMEM allocator + MMA issuer (one elected lane,
//   tcgen05.mma.cta_group::1.kind::f16, 128 x BN x 16 per instruction, completion via tcgen05.commit ->
//   mbarrier). Warps 2-5: epilogue; each owns the 32 TMEM lanes it is allowed to read, pulls 32 fp32
//   columns per tcgen05.ld, converts to bf16 and stores 64 B runs of one output pixel's channels.
// * STAGES-deep full/empty mbarrier ring between producer and MMA; several CTAs per SM (<= 113 KB smem
//   and <= 128 TMEM columns each) so one tile's epilogue overlaps another tile's main loop.
//
// Every mbarrier wait is bounded: a protocol bug traps (CUDA error) instead of hanging the GPU.
#include "tc_common.cuh"

namespace iit {

constexpr int TC_BM = 128;      // pixels per tile (UMMA M)
constexpr int TC_BK = 64;       // bf16 elements per 128-byte swizzle row
constexpr int TC_UK = 16;       // UMMA K for 16-bit inputs
constexpr int TC_THREADS = 192;
constexpr int TC_CONV_THREADS = 64 + 256;   // fprop / dgrad: TMA warp, MMA warp, eight epilogue warps

struct TcParams {
  int N, P, Q;            // pixel grid enumerated by the tiles (a parity class of the output for strided dgrad)
  int OH, OW;             // full spatial size of the produced tensor
  int out_mul, out_oh, out_ow;  // produced pixel = (p*out_mul + out_oh, q*out_mul + out_ow)
  int in_mul;             // gathered row = p*in_mul + dh[tap] (the tensor map strides by in_mul too)
  int Cin, Cout;          // reduction channels per tap / produced channels
  int ntaps;              // 0 => the class receives no contribution: zeros are written
  int8_t dh[9], dw[9];    // per-tap offset into the gathered tensor
  int8_t wk[9];           // per-tap block index in the filter matrix (reduction offset = wk*Cin)
  int BW, BH, BNI;        // pixel box
  int tiles_w, tiles_h;   // boxes per image row / column
  int tiles_n;            // image blocks
  int accumulate;         // != 0: the produced values are ADDED to what `out` holds (st_or_add_bf16x8)
  // GEMM epilogues (the transformer's linear maps; single-CTA kernel only): applied to the fp32 accumulator before rounding
  int epi;                        // TC_EPI_*
  const float* bias;              // [Cout] fp32, added to every row (TC_EPI_BIAS / TC_EPI_GELU); may be null
  const float* scale;             // [Cout] fp32 or null: the accumulator is multiplied by it before `bias` is added (an
                                  // eval-mode BatchNorm folded into the convolution: scale = gamma / std, bias = beta - mean * scale)
  int relu;                       // != 0: max(0, .) after bias / residual
  const __nv_bfloat16* aux;       // laid out like `out`: the residual added (TC_EPI_BIAS, may be null) or the pre-activation
                                  // whose GELU derivative multiplies the result (TC_EPI_GELU_BWD)
  __nv_bfloat16* out2;            // TC_EPI_GELU: receives the pre-activation (may be null: not kept)
  int ksplit, kper;               // ksplit > 1: the reduction (one tap) is cut into ksplit ranges of kper 64-deep blocks, each
                                  // a tile of its own whose result is ADDED to `out` (zeroed by the host): few-tile, very deep
                                  // GEMMs (the unembedding's input gradient: 8 tiles x 786 blocks) then fill the machine
  int box[6];                     // TC_EPI_PATCH: output rows p in [box0, box1), columns q in [box2, box3), channels in [box4, box5)
                                  // take their value from `aux` (the cached source activation) instead of the convolution
};
enum { TC_EPI_NONE = 0, TC_EPI_BIAS = 1, TC_EPI_GELU = 2, TC_EPI_GELU_BWD = 3, TC_EPI_PATCH = 4 };

// gelu_new and its derivative with tanh.approx (2^-11 relative, below bf16's 2^-9 rounding): the same arithmetic as the
// element-wise kernels of transformer.cu
__device__ __forceinline__ float tc_tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float tc_gelu(float x) {
  return 0.5f * x * (1.f + tc_tanh_fast(0.7978845608028654f * (x + 0.044715f * x * x * x)));
}
__device__ __forceinline__ float tc_gelu_grad(float x) {
  const float k = 0.7978845608028654f;
  const float th = tc_tanh_fast(k * (x + 0.044715f * x * x * x));
  return 0.5f * (1.f + th) + 0.5f * x * (1.f - th * th) * k * (1.f + 3.f * 0.044715f * x * x);
}

// Persistent: gridDim.x CTAs walk the tile list (tile = pixel box x filter block; filter block fastest so that CTAs
// working at the same time share the pixel box through L2). The shared-memory ring runs across tile boundaries, and
// the accumulator is double-buffered in TMEM (2 x BN columns): while the epilogue warps drain tile i the MMA warp is
// already accumulating tile i+1.
// MT pixel boxes share one filter block per tile (MT = 2: 256 pixels x 128 filters). The kernel is bound by L2 -> shared
// memory traffic (ncu: ~11 TB/s through the crossbar at 36-42 % tensor-pipe activity with 128 x 128 tiles); a filter block
// that serves two pixel boxes cuts the bytes per FLOP by a quarter.
template <int BN, int STAGES, int MT, int OBUF>
__global__ void __launch_bounds__(TC_CONV_THREADS) conv3x3_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                                const __grid_constant__ CUtensorMap map_b,
                                                                const __grid_constant__ CUtensorMap map_o,
                                                                const __grid_constant__ CUtensorMap map_o2,
                                                                __nv_bfloat16* __restrict__ out, TcParams p,
                                                                double* __restrict__ bn_sum, double* __restrict__ bn_sq) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BN * TC_BK * 2;
  uint8_t* sA = smem;                         // [STAGES][MT] pixel boxes
  uint8_t* sB = smem + STAGES * MT * A_BYTES;
  uint8_t* s_out = sB + STAGES * B_BYTES;     // [OBUF][128 rows][128 bytes], SWIZZLE_128B: output tiles of the rounds (tc_round_store)
  uint64_t* full = reinterpret_cast<uint64_t*>(s_out + OBUF * TC_OUT_TILE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_stat = reinterpret_cast<float*>(s_out);           // [4 lane quarters][2][BN] (<= 8 KB) over the first output tile: tc_flush_stats

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kchunks = p.Cin / TC_BK;
  const int nk = p.ntaps * kchunks;
  const int nblk = p.Cout / BN;
  const int mtiles = p.tiles_w * p.tiles_h * p.tiles_n;            // pixel boxes
  const int real_tiles = ((mtiles + MT - 1) / MT) * nblk;
  const int ksplit = p.ksplit > 1 ? p.ksplit : 1;
  const int kper = p.ksplit > 1 ? p.kper : kchunks;                // k-blocks per tile and tap
  const int total_tiles = real_tiles * ksplit;                     // tile = split * real_tiles + (pixel box, filter block)

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM allocation is warp-collective; this warp also frees it
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * MT * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (nk > 0) {   // warp-uniform loop, one elected lane issues: the producer must stay far below the MMA time per k-block
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_o)) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int sp = tile / real_tiles, rt = tile - sp * real_tiles;
        const int kc_lo = sp * kper * TC_BK, kc_hi = min(p.Cin, (sp + 1) * kper * TC_BK);
        const int co0 = (rt % nblk) * BN;
        int n0[MT], p0[MT], q0[MT];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          int t = (rt / nblk) * MT + mt;
          if (t >= mtiles) t = mtiles - 1;            // odd tail: the box is loaded twice, its second copy never stored
          const int tw = t % p.tiles_w; t /= p.tiles_w;
          const int th = t % p.tiles_h; t /= p.tiles_h;
          n0[mt] = t * p.BNI; p0[mt] = th * p.BH * p.in_mul; q0[mt] = tw * p.BW * p.in_mul;
        }
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int dh = p.dh[tap], dw = p.dw[tap], wk0 = p.wk[tap] * p.Cin;
          for (int kc = kc_lo; kc < kc_hi; kc += TC_BK) {
            mbar_wait(&empty[stage], phase ^ 1);
            if (issuer) {
              mbar_expect_tx(&full[stage], MT * A_BYTES + B_BYTES);
#pragma unroll
              for (int mt = 0; mt < MT; ++mt)
                tma_load_4d(sA + (stage * MT + mt) * A_BYTES, &map_a, &full[stage], kc, q0[mt] + dw, p0[mt] + dh, n0[mt]);
              tma_load_2d(sB + stage * B_BYTES, &map_b, &full[stage], wk0 + kc, co0);
            }
            __syncwarp();
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (nk > 0) {   // the whole warp runs the loop; the elected lane issues (see elect_one)
      const bool issuer = elect_one();
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N at [17,23), M at [24,29)
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      // descriptor = constant high word + low word (1 << 16) | (address >> 4): operands are stepped with 32-bit adds
      const uint64_t proto = umma_desc_k_sw128(0);
      const uint32_t desc_hi = (uint32_t)(proto >> 32);
      const uint32_t a_lo0 = (uint32_t)proto + (smem_u32(sA) >> 4), b_lo0 = (uint32_t)proto + (smem_u32(sB) >> 4);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);   // the epilogue has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t tmem_d = tmem_base + acc * (MT * BN);
        const int sp = tile / real_tiles;
        const int nk_t = p.ntaps * (min(kchunks, (sp + 1) * kper) - sp * kper);     // k-blocks of this tile's reduction range
        for (int kb = 0; kb < nk_t; ++kb) {
          mbar_wait(&full[stage], phase);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_lo = a_lo0 + stage * (MT * A_BYTES >> 4), b_lo = b_lo0 + stage * (B_BYTES >> 4);
          if (issuer) {
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
              for (int k = 0; k < TC_BK / TC_UK; ++k)
                umma_bf16_lohi(tmem_d + mt * BN, a_lo + mt * (A_BYTES >> 4) + 2 * k, desc_hi, b_lo + 2 * k, desc_hi, idesc,
                               (uint32_t)((kb | k) != 0));
            umma_commit(&empty[stage]);  // frees this smem stage once the MMAs above have read it
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (issuer) umma_commit(&tmem_full[acc]);  // accumulator complete
        __syncwarp();
      }
    }
  } else {
    // epilogue, eight warps: warp w may read TMEM lanes [32*(w%4), 32*(w%4)+32), so two warps share each lane quarter and
    // take the two 32-column halves of every 64-channel round; the round's [128 x 64] tile goes out through one TMA store
    // (tc_round_store). With small reductions (1x1 and 64-channel layers, the d_model-64 GEMMs) the tile time IS the
    // epilogue time.
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;   // which of the two warps of this lane quarter
    const int m = quarter * 32 + lane;  // accumulator row == pixel of the box
    const int wl = m % p.BW;
    const int hl = (m / p.BW) % p.BH;
    const int nl = m / (p.BW * p.BH);
    const int e = threadIdx.x - 64;     // 0..255: index among the epilogue threads
    const bool leader = e == 0;         // issues the TMA stores
    // BatchNorm statistics: running sums in registers (tc_round_store), flushed when the filter block changes
    float rstat[BN / 64][4];
#pragma unroll
    for (int cc = 0; cc < BN / 64; ++cc) { rstat[cc][0] = 0.f; rstat[cc][1] = 0.f; rstat[cc][2] = 0.f; rstat[cc][3] = 0.f; }
    int stat_tiles = 0;                 // tiles summed since the last flush (fp32 running sums are kept short)
    const bool one_box = p.tiles_w == 1 && p.tiles_h == 1;     // the box covers whole images: no division per tile
    int run_co = -1;
    uint32_t it = 0, round = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int rt = ksplit > 1 ? tile % real_tiles : tile;
      const int co0 = (nblk > 1 ? rt % nblk : 0) * BN;
      if (bn_sum != nullptr && run_co >= 0 && (run_co != co0 || stat_tiles >= 16)) {
        tc_flush_stats<BN>(rstat, s_stat, quarter, half, lane, e, run_co, bn_sum, bn_sq, leader);
        stat_tiles = 0;
      }
      run_co = co0;
      ++stat_tiles;
      const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
     for (int mt = 0; mt < MT; ++mt) {
      int t = (nblk > 1 ? rt / nblk : rt) * MT + mt;
      const bool box_ok = t < mtiles;
      int tw = 0, th = 0;
      if (!one_box) { tw = t % p.tiles_w; t /= p.tiles_w; th = t % p.tiles_h; t /= p.tiles_h; }
      const int n_box = t * p.BNI, p_box = th * p.BH, q_box = tw * p.BW;
      const int n = n_box + nl, pp = p_box + hl, qq = q_box + wl;
      const bool ok = box_ok && n < p.N && pp < p.P && qq < p.Q;
#pragma unroll
      for (int cc = 0; cc < BN / 64; ++cc) {
        const int c = cc * 64 + half * 32;
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * (MT * BN) + mt * BN + c), v);
        if (cc == BN / 64 - 1 && mt == MT - 1) {
          // all of this warp's TMEM reads of the accumulator are complete: hand it back to the MMA warp
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty[acc])) : "memory");
        }
        if (p.epi != TC_EPI_NONE) {
          // this lane's own row (pixel n, pp, qq), channels co0 + c .. + 31, still fp32
          if (p.scale != nullptr) tc_scale_shift(v, p.scale + co0 + c, p.bias + co0 + c);
          else if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(p.bias + co0 + c + j);
              v[j] = __float_as_uint(__uint_as_float(v[j]) + b4.x); v[j + 1] = __float_as_uint(__uint_as_float(v[j + 1]) + b4.y);
              v[j + 2] = __float_as_uint(__uint_as_float(v[j + 2]) + b4.z); v[j + 3] = __float_as_uint(__uint_as_float(v[j + 3]) + b4.w);
            }
          }
          const int64_t own = (((int64_t)n * p.OH + (pp * p.out_mul + p.out_oh)) * p.OW + (qq * p.out_mul + p.out_ow)) * p.Cout + co0 + c;
          const bool in_pq = p.epi != TC_EPI_PATCH || (pp >= p.box[0] && pp < p.box[1] && qq >= p.box[2] && qq < p.box[3] &&
                                                       co0 + c < p.box[5] && co0 + c + 32 > p.box[4]);
          if (p.aux != nullptr && ok && in_pq) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              const uint4 a4 = *reinterpret_cast<const uint4*>(p.aux + own + j);
              const uint32_t w4[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
              for (int h2 = 0; h2 < 4; ++h2) {
                const float a0 = __uint_as_float(w4[h2] << 16), a1 = __uint_as_float(w4[h2] & 0xffff0000u);
                if (p.epi == TC_EPI_PATCH) {      // the interchange: dst[box] = src[box], inside the producing layer
                  const int ch = co0 + c + j + 2 * h2;
                  if (ch >= p.box[4] && ch < p.box[5]) v[j + 2 * h2] = __float_as_uint(a0);
                  if (ch + 1 >= p.box[4] && ch + 1 < p.box[5]) v[j + 2 * h2 + 1] = __float_as_uint(a1);
                } else if (p.epi == TC_EPI_GELU_BWD) {
                  v[j + 2 * h2] = __float_as_uint(__uint_as_float(v[j + 2 * h2]) * tc_gelu_grad(a0));
                  v[j + 2 * h2 + 1] = __float_as_uint(__uint_as_float(v[j + 2 * h2 + 1]) * tc_gelu_grad(a1));
                } else {
                  v[j + 2 * h2] = __float_as_uint(__uint_as_float(v[j + 2 * h2]) + a0);
                  v[j + 2 * h2 + 1] = __float_as_uint(__uint_as_float(v[j + 2 * h2 + 1]) + a1);
                }
              }
            }
          }
          if (p.relu) tc_relu(v);
          if (p.epi == TC_EPI_GELU) {
            if (p.out2 != nullptr) {     // the pre-activation, kept for the backward: a round of its own through the same tile
              float none[4] = {0.f, 0.f, 0.f, 0.f};
              tc_round_store<OBUF>(v, ok, s_out + (round++ % OBUF) * TC_OUT_TILE_BYTES, quarter, half, lane, leader, box_ok, &map_o2,
                                   co0 + cc * 64, q_box, p_box, n_box, false, none, false);
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(tc_gelu(__uint_as_float(v[j])));
          }
        }
        // the round's tile: staged, stored by TMA, and -- for the BatchNorm that follows -- its per-channel sums as stored
        tc_round_store<OBUF>(v, ok, s_out + (round++ % OBUF) * TC_OUT_TILE_BYTES, quarter, half, lane, leader, box_ok, &map_o,
                             co0 + cc * 64, q_box, p_box, n_box, p.accumulate != 0, rstat[cc], bn_sum != nullptr);
      }
     }  // mt
    }
    if (leader) tma_store_wait_all();    // every store of this CTA has been written before it may exit
    if (bn_sum != nullptr && run_co >= 0) tc_flush_stats<BN>(rstat, s_stat, quarter, half, lane, e, run_co, bn_sum, bn_sq, leader);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * MT * BN));
  }
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2) for layers with a multiple of 256 filters: the two CTAs of a cluster compute
// a tile of 256 pixels x 256 filters. Each CTA loads ITS pixel box (128 x 64 per k-block) and HALF of the filter block
// (128 x 64); the leader issues M=256, N=256 MMAs that read both halves. Per CTA and k-block that is 32 KB from L2 for
// 128 x 256 x 64 MACs - 128 FLOP per byte against 87 for the single-CTA 128 x 256 tile, which sat on the L2->SMEM limit -
// and 8 KB of shared-memory reads per 128 tensor cycles instead of 12.
// Barriers: `full` lives in the leader (both CTAs' TMA loads credit it); `empty` and `tmem_full` exist in both CTAs and
// are signalled by multicast commits; `tmem_empty` lives in the leader and is arrived on by the epilogue warps of both.
// ------------------------------------------------------------------------------------------------------------------
// BN = filters per tile (accumulator columns per CTA); each CTA loads BN / 2 filter rows. BN = 128 gives tiles of half the
// size at the traffic per FLOP of the single-CTA 128 x 256 tile: it is chosen when it fills the last wave better.
template <int BN, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_CONV_THREADS, 1)
conv3x3_tc2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_o, __nv_bfloat16* __restrict__ out, TcParams p,
                   double* __restrict__ bn_sum, double* __restrict__ bn_sq) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int BHALF = BN / 2;
  constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BHALF * TC_BK * 2;
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint8_t* s_out = sB + STAGES * B_BYTES;     // this CTA's two [128 rows][128 bytes] output tiles (tc_round_store)
  uint64_t* full = reinterpret_cast<uint64_t*>(s_out + 2 * TC_OUT_TILE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2] (the leader's are used)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_stat = reinterpret_cast<float*>(s_out);           // over the first output tile: tc_flush_stats

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int kchunks = p.Cin / TC_BK;
  const int nk = p.ntaps * kchunks;
  const int nblk = p.Cout / BN;
  const int mtiles = p.tiles_w * p.tiles_h * p.tiles_n;
  const int total_tiles = ((mtiles + 1) / 2) * nblk;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 16); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the same warp of both CTAs allocates (and frees) the pair's TMEM columns
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  cluster_sync_all();   // barriers of both CTAs initialised, TMEM allocated
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (nk > 0) {   // warp-uniform loops, the elected lane issues (tc_common.cuh: elect_one)
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t lead_full0 = mapa_u32(smem_u32(&full[0]), 0);
      for (int tile = pair; tile < total_tiles; tile += npairs) {
        const int co0 = (tile % nblk) * BN + (int)rank * BHALF;
        int t = (tile / nblk) * 2 + (int)rank;
        if (t >= mtiles) t = mtiles - 1;              // odd tail: the peer repeats the last box, its copy is never stored
        const int tw = t % p.tiles_w; t /= p.tiles_w;
        const int th = t % p.tiles_h; t /= p.tiles_h;
        const int n0 = t * p.BNI, p0 = th * p.BH * p.in_mul, q0 = tw * p.BW * p.in_mul;
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int dh = p.dh[tap], dw = p.dw[tap], wk0 = p.wk[tap] * p.Cin;
          for (int kc = 0; kc < p.Cin; kc += TC_BK) {
            mbar_wait(&empty[stage], phase ^ 1);      // multicast commit of the leader's MMAs: this CTA's stage is free
            if (issuer) {
              const uint32_t lead_full = lead_full0 + stage * 8;
              if (rank == 0) mbar_expect_tx(&full[stage], 2 * (A_BYTES + B_BYTES));
              tma2_load_4d(sA + stage * A_BYTES, &map_a, lead_full, kc, q0 + dw, p0 + dh, n0);
              tma2_load_2d(sB + stage * B_BYTES, &map_b, lead_full, wk0 + kc, co0);
            }
            __syncwarp();
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (nk > 0 && rank == 0) {
      const bool issuer = elect_one();
      // D=f32, A=B=bf16, K-major; N = BN at [17,23), M = 256 (both CTAs) at [24,29)
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0, it = 0;
      const uint64_t proto = umma_desc_k_sw128(0);
      const uint32_t desc_hi = (uint32_t)(proto >> 32);
      const uint32_t a_lo0 = (uint32_t)proto + (smem_u32(sA) >> 4), b_lo0 = (uint32_t)proto + (smem_u32(sB) >> 4);
      for (int tile = pair; tile < total_tiles; tile += npairs, ++it) {
        const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
        mbar_wait_cluster(&tmem_empty[acc], acc_phase ^ 1);   // epilogue warps of BOTH CTAs have drained it
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait_cluster(&full[stage], phase);              // both CTAs' loads of this stage have landed
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_lo = a_lo0 + stage * (A_BYTES >> 4), b_lo = b_lo0 + stage * (B_BYTES >> 4);
          if (issuer) {
#pragma unroll
            for (int k = 0; k < TC_BK / TC_UK; ++k)
              umma2_bf16_lohi(tmem_d, a_lo + 2 * k, desc_hi, b_lo + 2 * k, desc_hi, idesc, (uint32_t)((kb | k) != 0));
            umma2_commit_both(&empty[stage]);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (issuer) umma2_commit_both(&tmem_full[acc]);
        __syncwarp();
      }
    }
  } else {
    // eight epilogue warps per CTA, two per TMEM lane quarter on the two halves of each 64-channel round; each CTA stores
    // ITS 128 pixels of the pair's tile by TMA (tc_round_store)
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int m = quarter * 32 + lane;
    const int wl = m % p.BW;
    const int hl = (m / p.BW) % p.BH;
    const int nl = m / (p.BW * p.BH);
    const int e = threadIdx.x - 64;                        // 0..255
    const bool leader = e == 0;
    float rstat[BN / 64][4];                               // running BatchNorm statistics (tc_round_store / tc_flush_stats)
#pragma unroll
    for (int cc = 0; cc < BN / 64; ++cc) { rstat[cc][0] = 0.f; rstat[cc][1] = 0.f; rstat[cc][2] = 0.f; rstat[cc][3] = 0.f; }
    int stat_tiles = 0;
    const bool one_box = p.tiles_w == 1 && p.tiles_h == 1;
    int run_co = -1;
    uint32_t it = 0, round = 0;
    for (int tile = pair; tile < total_tiles; tile += npairs, ++it) {
      const int co0 = (tile % nblk) * BN;
      if (bn_sum != nullptr && run_co >= 0 && (run_co != co0 || stat_tiles >= 16)) {
        tc_flush_stats<BN>(rstat, s_stat, quarter, half, lane, e, run_co, bn_sum, bn_sq, leader);
        stat_tiles = 0;
      }
      run_co = co0;
      ++stat_tiles;
      const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
      const uint32_t lead_empty = mapa_u32(smem_u32(&tmem_empty[acc]), 0);
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      int t = (tile / nblk) * 2 + (int)rank;
      const bool box_ok = t < mtiles;
      int tw = 0, th = 0;
      if (!one_box) { tw = t % p.tiles_w; t /= p.tiles_w; th = t % p.tiles_h; t /= p.tiles_h; }
      const int n_box = t * p.BNI, p_box = th * p.BH, q_box = tw * p.BW;
      const bool ok = box_ok && n_box + nl < p.N && p_box + hl < p.P && q_box + wl < p.Q;
#pragma unroll
      for (int cc = 0; cc < BN / 64; ++cc) {
        const int c = cc * 64 + half * 32;
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + c), v);
        if (cc == BN / 64 - 1) {
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(lead_empty);
        }
        if (p.epi == TC_EPI_BIAS) {     // folded eval-mode BatchNorm: scale / shift (+ residual) (+ ReLU) on the fp32 accumulator
          if (p.scale != nullptr) tc_scale_shift(v, p.scale + co0 + c, p.bias + co0 + c);
          if (p.aux != nullptr && ok) {
            const int64_t own = (((int64_t)(n_box + nl) * p.OH + ((p_box + hl) * p.out_mul + p.out_oh)) * p.OW +
                                 ((q_box + wl) * p.out_mul + p.out_ow)) * p.Cout + co0 + c;
            tc_add_bf16_row(v, p.aux + own);
          }
          if (p.relu) tc_relu(v);
        }
        tc_round_store<2>(v, ok, s_out + (round++ & 1) * TC_OUT_TILE_BYTES, quarter, half, lane, leader, box_ok, &map_o,
                          co0 + cc * 64, q_box, p_box, n_box, p.accumulate != 0, rstat[cc], bn_sum != nullptr);
      }
    }
    if (leader) tma_store_wait_all();
    if (bn_sum != nullptr && run_co >= 0) tc_flush_stats<BN>(rstat, s_stat, quarter, half, lane, e, run_co, bn_sum, bn_sq, leader);
  }
  // neither CTA may leave (or free the pair's TMEM) while its peer can still signal its barriers or read its operands
  __syncwarp();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  cluster_sync_all();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN));
  }
}

// ---- host side -----------------------------------------------------------------------------------
static bool g_tile_pairs = true;
static bool g_bn256 = true;

static int pow2_ge(int v) { int r = 1; while (r < v) r <<= 1; return r; }

template <int BN, int MT>
static int launch_tc(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mo, const CUtensorMap& mo2, __nv_bfloat16* out,
                     const TcParams& p, int total_tiles, cudaStream_t st, double* bn_sum, double* bn_sq) {
  // Two output tiles (the TMA unit reads one while the warps fill the other) wherever shared memory has room for them next
  // to the operand ring: 256-filter tiles keep them by running a three-stage ring (their MMAs take 128 cycles per k-block
  // step, so three stages in flight cover the load latency as four do for the narrower tiles); 2 x 128 pixels x 128 filters
  // (four stages of 48 KB) has room for one.
  // 64-filter tiles stream long, narrow reductions (the transformer's [tokens x 3072] x [3072 x 64] maps: 24 KB per k-block and
  // 192 cycles of MMAs): seven stages keep ~170 KB of loads in flight per SM, what HBM latency x bandwidth asks for
  constexpr int STAGES = BN == 256 ? 3 : BN == 64 ? 7 : 4;
  constexpr int OBUF = 2;                                   // the statistics scratch lives in the first tile (tc_flush_stats)
  constexpr size_t smem = STAGES * (MT * TC_BM * TC_BK * 2 + BN * TC_BK * 2) + OBUF * TC_OUT_TILE_BYTES + (2 * STAGES + 4) * 8 + 16 +
                          1024;
  static_assert(smem <= 227 * 1024, "tile does not fit shared memory");
  auto kernel = conv3x3_tc_kernel<BN, STAGES, MT, OBUF>;
  static int ctas_per_sm_dev[kMaxDevices] = {};
  int& ctas_per_sm = ctas_per_sm_dev[current_device_slot()];
  if (!ctas_per_sm) {
    IIT_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    IIT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, TC_CONV_THREADS, smem));
    const int tmem_limit = 512 / (2 * MT * BN);   // every resident CTA holds 2*MT*BN TMEM columns for its whole life
    ctas_per_sm = occ < 1 ? 1 : (occ > tmem_limit ? tmem_limit : occ);
  }
  int grid = kNumSMs * ctas_per_sm;
  if (grid > total_tiles) grid = total_tiles;
  kernel<<<grid, TC_CONV_THREADS, smem, st>>>(ma, mb, mo, mo2, out, p, bn_sum, bn_sq);
  count_variant(BN == 64 ? KV_TC_64 : BN == 256 ? KV_TC_256 : MT == 2 ? KV_TC_128_MT2 : KV_TC_128);
  return 0;
}

// CTA-pair launch (256 pixels x BN filters per cluster of two)
static int g_pair_mode = 1;   // 0 off; 1 on (Cout % 256 == 0: 256-filter tiles; Cout == 128 with long reductions: 128); 2 without the 128-filter layers; 3 128-filter tiles everywhere
template <int BN>
static int launch_tc2(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mo, __nv_bfloat16* out, const TcParams& p,
                      int total_tiles, cudaStream_t st, double* bn_sum, double* bn_sq) {
  constexpr int STAGES = BN == 256 ? 6 : 8;
  constexpr size_t smem = STAGES * (TC_BM * TC_BK * 2 + (BN / 2) * TC_BK * 2) + 2 * TC_OUT_TILE_BYTES + (2 * STAGES + 4) * 8 + 16 +
                          1024;
  static_assert(smem <= 227 * 1024, "pair tile does not fit shared memory");
  static int max_clusters_dev[kMaxDevices] = {};
  int& max_clusters = max_clusters_dev[current_device_slot()];
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(TC_CONV_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (!max_clusters) {
    IIT_CUDA(cudaFuncSetAttribute(conv3x3_tc2_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfg.gridDim = dim3(kNumSMs);
    int n = 0;
    IIT_CUDA(cudaOccupancyMaxActiveClusters(&n, conv3x3_tc2_kernel<BN, STAGES>, &cfg));
    max_clusters = n < 1 ? 1 : (n > kNumSMs / 2 ? kNumSMs / 2 : n);
  }
  const int clusters = total_tiles < max_clusters ? total_tiles : max_clusters;
  cfg.gridDim = dim3(2 * clusters);
  IIT_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_tc2_kernel<BN, STAGES>, ma, mb, mo, out, p, bn_sum, bn_sq));
  count_variant(BN == 256 ? KV_TC2_256 : KV_TC2_128);
  return 0;
}

// Pixel box of a tile: BW x BH x BNI = `pixels` rows (powers of two; TMA enumerates w fastest, then h, then the image).
// The box that wastes the fewest rows on positions outside the P x Q grid wins; among equals the one with the largest
// spatial extent (longer contiguous runs). A 7 x 7 or 11 x 11 grid is covered exactly by 1 x 1 x 128 (the same pixel of
// 128 images) where 8 x 8 x 2 / 16 x 8 x 1 used 77 % / 47 % of the rows.
static void plan_box(int P, int Q, int pixels, int* BW, int* BH, int* BNI) {
  double best = -1.0;
  for (int bw = 1; bw <= pixels && bw <= 256; bw <<= 1)
    for (int bh = 1; bw * bh <= pixels && bh <= 256; bh <<= 1) {
      const int bn = pixels / (bw * bh);
      if (bn > 256) continue;
      const double covered = (double)(((Q + bw - 1) / bw) * bw) * (((P + bh - 1) / bh) * bh);
      const double score = (double)P * Q / covered + 1e-6 * (bw * bh) + 1e-9 * bw;
      if (score > best) { best = score; *BW = bw; *BH = bh; *BNI = bn; }
    }
}

static int encode_nhwc_strided(CUtensorMap* m, const void* ptr, int C, int W, int H, int N, int bw, int bh, int bn, int es_hw,
                               const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  // with an element stride e the box is given in traversed elements: e*n to load n
  cuuint32_t box[4] = {64, (cuuint32_t)(bw * es_hw), (cuuint32_t)(bh * es_hw), (cuuint32_t)bn};
  cuuint32_t es[4] = {1, (cuuint32_t)es_hw, (cuuint32_t)es_hw, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

// The produced tensor as the epilogue's TMA stores see it: the pixel grid (p, q) of this launch -- the whole [N, OH, OW, Cout]
// tensor, or one parity class of it for a strided dgrad (base shifted to the class's first pixel, pixel strides multiplied)
// -- in boxes of 64 channels x BW x BH x BNI pixels, SWIZZLE_128B (tc_round_store).
static int encode_out(CUtensorMap* m, void* out, const TcParams& tp, const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  uint8_t* base = reinterpret_cast<uint8_t*>(out) + ((size_t)tp.out_oh * tp.OW + tp.out_ow) * tp.Cout * 2;
  cuuint64_t dims[4] = {(cuuint64_t)tp.Cout, (cuuint64_t)tp.Q, (cuuint64_t)tp.P, (cuuint64_t)tp.N};
  cuuint64_t strides[3] = {(cuuint64_t)tp.out_mul * tp.Cout * 2, (cuuint64_t)tp.out_mul * tp.OW * tp.Cout * 2,
                           (cuuint64_t)tp.OH * tp.OW * tp.Cout * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)tp.BW, (cuuint32_t)tp.BH, (cuuint32_t)tp.BNI};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: output tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

static int encode_filter(CUtensorMap* m, const void* wmat, int Kg, int rows, int BN, const char* name) {
  EncodeTiledFn enc = get_encode();
  IIT_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is unavailable in this driver", name);
  cuuint64_t dims[2] = {(cuuint64_t)Kg, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)Kg * 2};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)BN};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(wmat), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  IIT_REQUIRE(r == CUDA_SUCCESS, "%s: filter tensor map rejected (CUresult %d)", name, (int)r);
  return 0;
}

// One launch: the produced pixels (p*out_mul+out_oh, q*out_mul+out_ow), p<P, q<Q, from the taps listed in `tp`.
static int conv_tc_launch(const void* gathered, int gC, int gW, int gH, const void* wmat, int Kg, void* out, TcParams tp,
                          cudaStream_t st, const char* name, double* bn_sum = nullptr, double* bn_sq = nullptr) {
  plan_box(tp.P, tp.Q, TC_BM, &tp.BW, &tp.BH, &tp.BNI);
  tp.tiles_w = (tp.Q + tp.BW - 1) / tp.BW;
  tp.tiles_h = (tp.P + tp.BH - 1) / tp.BH;
  tp.tiles_n = (tp.N + tp.BNI - 1) / tp.BNI;
  // 256 filters per tile where the layer has them: with N = 256 an MMA reads 12 KB of shared memory per 128 tensor cycles
  // (96 B/clk); with N = 128 it is 8 KB per 64 cycles, exactly the 128 B/clk shared memory delivers
  const int BN = (tp.Cout % 256 == 0 && g_bn256) ? 256 : (tp.Cout % 128 == 0) ? 128 : 64;
  IIT_REQUIRE(tp.ntaps > 0 && tp.Cin >= TC_BK, "%s: empty reduction", name);
  CUtensorMap ma, mb, mo, mo2;
  if (encode_nhwc_strided(&ma, gathered, gC, gW, gH, tp.N, tp.BW, tp.BH, tp.BNI, tp.in_mul, name)) return 1;
  if (encode_out(&mo, out, tp, name)) return 1;
  mo2 = mo;
  if (tp.out2 != nullptr && encode_out(&mo2, tp.out2, tp, name)) return 1;
  const int64_t mtiles = (int64_t)tp.tiles_w * tp.tiles_h * tp.tiles_n;
  // CTA pairs: whenever every SM gets a tile; with fewer boxes (layer4 at batch 4096: 128) only for long reductions, where
  // the halved filter traffic outweighs the pair's start-up (measured: +3 % on 4608-deep, -8 % on 256-deep reductions)
  // ... and not for reductions of one or two k-blocks (the 1x1 downsample 128 -> 256: 24.6 us on pairs, 20.5 us on single
  // CTAs, which keep two output tiles in flight)
  const bool pair_epi = tp.epi == TC_EPI_NONE || (tp.epi == TC_EPI_BIAS && tp.scale != nullptr && tp.out2 == nullptr);
  // 128-filter layers with a long reduction (layer2's 3x3s) run on pairs too: 256 pixels x 128 filters per pair, each CTA
  // loads half the filter block (layer2 fprop 61.0 -> 58.9 us, dgrad 59.9 -> 57.9 us against two pixel boxes per single CTA;
  // mode 2 keeps them on single CTAs)
  const bool pair_cout = tp.Cout % 256 == 0 || (g_pair_mode != 2 && tp.Cout == 128 && tp.ntaps * tp.Cin >= 1024);
  if (g_pair_mode && pair_epi && pair_cout && tp.ntaps > 0 && tp.ntaps * tp.Cin >= 512 &&
      (mtiles >= kNumSMs || (mtiles >= kNumSMs / 2 && tp.ntaps * tp.Cin >= 2304))) {
    // 256-filter tiles unless mode 3 asks for 128 (measured: half-width tiles fill the last wave better, 7 waves of 74
    // clusters instead of 4 on layer3, but pay the pixel box twice and end up slower: 76.6 vs 70.5 us)
    const int64_t pairs = (mtiles + 1) / 2;
    const int64_t t256 = pairs * (tp.Cout / 256), t128 = pairs * (tp.Cout / 128);
    const int BN2 = (g_pair_mode == 3 || tp.Cout % 256 != 0) ? 128 : 256;
    if (encode_filter(&mb, wmat, Kg, tp.Cout, BN2 / 2, name)) return 1;
    const int64_t total2 = BN2 == 256 ? t256 : t128;
    IIT_REQUIRE(total2 < (1ll << 31), "%s: too many tiles", name);
    if (BN2 == 256 ? launch_tc2<256>(ma, mb, mo, (__nv_bfloat16*)out, tp, (int)total2, st, bn_sum, bn_sq)
                   : launch_tc2<128>(ma, mb, mo, (__nv_bfloat16*)out, tp, (int)total2, st, bn_sum, bn_sq)) return 1;
    IIT_LAUNCH_CHECK(name);
    return 0;
  }
  if (encode_filter(&mb, wmat, Kg, tp.Cout, BN, name)) return 1;
  // few tiles and a very deep single-tap reduction: cut the reduction so that every SM gets a tile; the partial results are
  // added onto the zeroed output by the accumulate epilogue (bf16 reductions: at most a few dozen addends per element)
  const int kchunks_all = tp.Cin / TC_BK;
  const int64_t base_tiles = mtiles * (tp.Cout / BN);
  if (tp.ntaps == 1 && tp.epi == TC_EPI_NONE && !tp.accumulate && bn_sum == nullptr && base_tiles * 2 <= kNumSMs && kchunks_all >= 64 &&
      tp.out_mul == 1) {
    int want = (int)(kNumSMs / base_tiles);
    if (want > kchunks_all / 8) want = kchunks_all / 8;
    if (want > 1) {
      tp.kper = (kchunks_all + want - 1) / want;
      tp.ksplit = (kchunks_all + tp.kper - 1) / tp.kper;
      tp.accumulate = 1;
      IIT_CUDA(cudaMemsetAsync(out, 0, (size_t)tp.N * tp.OH * tp.OW * tp.Cout * sizeof(__nv_bfloat16), st));
    }
  }
  // two pixel boxes per filter block whenever that still leaves a tile for every SM
  // ... and the reduction is long enough to pay for the doubled epilogue per tile (measured: 1152-deep gains, 576-deep loses)
  const int MT = (BN == 128 && g_tile_pairs && ((mtiles + 1) / 2) * (tp.Cout / BN) >= kNumSMs && tp.ntaps * tp.Cin >= 1024) ? 2 : 1;
  const int64_t total = ((mtiles + MT - 1) / MT) * (tp.Cout / BN) * (tp.ksplit > 1 ? tp.ksplit : 1);
  IIT_REQUIRE(total < (1ll << 31), "%s: too many tiles", name);
  int rc = BN == 256 ? launch_tc<256, 1>(ma, mb, mo, mo2, (__nv_bfloat16*)out, tp, (int)total, st, bn_sum, bn_sq)
         : BN == 128 ? (MT == 2 ? launch_tc<128, 2>(ma, mb, mo, mo2, (__nv_bfloat16*)out, tp, (int)total, st, bn_sum, bn_sq)
                                : launch_tc<128, 1>(ma, mb, mo, mo2, (__nv_bfloat16*)out, tp, (int)total, st, bn_sum, bn_sq))
                     : launch_tc<64, 1>(ma, mb, mo, mo2, (__nv_bfloat16*)out, tp, (int)total, st, bn_sum, bn_sq);
  if (rc) return rc;
  IIT_LAUNCH_CHECK(name);
  return 0;
}

static int check_tc_ptrs(const void* a, const void* b, const void* c, const char* name) {
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15) == 0,
              "%s: pointers must be 16-byte aligned", name);
  return 0;
}

// y[n,p,q,k] = sum_{r,s,c} x[n, p*stride - pad + r, q*stride - pad + s, c] * w[k,r,s,c]
static int conv_tc_fprop(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, cudaStream_t st, double* bn_sum,
                         double* bn_sq, const TcParams* epi = nullptr, const char* name = "iit_conv2d_fprop_tc") {
  if (check_tc_ptrs(x, w_krsc, y, name)) return 1;
  TcParams tp{};
  if (epi != nullptr) {
    tp.epi = epi->epi; tp.bias = epi->bias; tp.aux = epi->aux; tp.out2 = epi->out2; tp.scale = epi->scale; tp.relu = epi->relu;
    for (int i = 0; i < 6; ++i) tp.box[i] = epi->box[i];
  }
  tp.N = d->n; tp.P = d->p; tp.Q = d->q; tp.OH = d->p; tp.OW = d->q;
  tp.out_mul = 1; tp.out_oh = 0; tp.out_ow = 0; tp.in_mul = d->stride;
  tp.Cin = d->c; tp.Cout = d->k;
  tp.ntaps = d->r * d->s;
  for (int r = 0; r < d->r; ++r)
    for (int s = 0; s < d->s; ++s) {
      int t = r * d->s + s;
      tp.dh[t] = (int8_t)(r - d->pad); tp.dw[t] = (int8_t)(s - d->pad); tp.wk[t] = (int8_t)t;
    }
  return conv_tc_launch(x, d->c, d->w, d->h, w_krsc, d->r * d->s * d->c, y, tp, st, name, bn_sum, bn_sq);
}

// dx[n,h,w,c] = sum_{r,s,k} dy[n,(h+pad-r)/stride,(w+pad-s)/stride,k] * w[k,r,s,c]  over exact divisions.
// For stride 2 the output splits into four parity classes, each a stride-1 gather of dy with its own taps.
// `accumulate`: dx += the result (one reduction per element; classes no tap reaches then need no launch and no memset).
static int conv_tc_dgrad(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, cudaStream_t st,
                         bool accumulate = false, const TcParams* epi = nullptr, double* col_sum = nullptr,
                         double* col_sq = nullptr, const char* name_override = nullptr) {
  const char* name = name_override ? name_override : accumulate ? "iit_conv2d_dgrad_tc_accumulate" : "iit_conv2d_dgrad_tc";
  if (check_tc_ptrs(dy, w_crsk, dx, name)) return 1;
  const int st_ = d->stride;
  // Parity classes of dx that no filter tap reaches (three of the four for a 1x1 stride-2 convolution) are zeros: one memset
  // of dx replaces their launches (each would push zeros through the tile epilogue into 128-byte rows 2 pixels apart).
  bool zero_class = false;
  for (int a = 0; a < st_; ++a)
    for (int b = 0; b < st_; ++b) {
      bool any = false;
      for (int r = 0; r < d->r; ++r)
        for (int s = 0; s < d->s; ++s) any = any || ((a + d->pad - r) % st_ == 0 && (b + d->pad - s) % st_ == 0);
      zero_class = zero_class || !any;
    }
  if (zero_class && !accumulate)
    IIT_CUDA(cudaMemsetAsync(dx, 0, (size_t)d->n * d->h * d->w * d->c * sizeof(__nv_bfloat16), st));
  for (int a = 0; a < st_; ++a)
    for (int b = 0; b < st_; ++b) {
      TcParams tp{};
      tp.N = d->n; tp.OH = d->h; tp.OW = d->w;
      tp.P = (d->h - a + st_ - 1) / st_; tp.Q = (d->w - b + st_ - 1) / st_;
      if (tp.P <= 0 || tp.Q <= 0) continue;
      tp.out_mul = st_; tp.out_oh = a; tp.out_ow = b; tp.in_mul = 1;
      tp.Cin = d->k; tp.Cout = d->c;
      tp.accumulate = accumulate ? 1 : 0;
      if (epi != nullptr) { tp.epi = epi->epi; tp.bias = epi->bias; tp.aux = epi->aux; tp.out2 = epi->out2; }
      int nt = 0;
      for (int r = 0; r < d->r; ++r) {
        if ((a + d->pad - r) % st_ != 0) continue;
        for (int s = 0; s < d->s; ++s) {
          if ((b + d->pad - s) % st_ != 0) continue;
          // floor division is exact here; negative numerators stay exact multiples
          tp.dh[nt] = (int8_t)((a + d->pad - r) / st_); tp.dw[nt] = (int8_t)((b + d->pad - s) / st_);
          tp.wk[nt] = (int8_t)(r * d->s + s);
          ++nt;
        }
      }
      tp.ntaps = nt;
      if (nt == 0) continue;     // zeroed above
      int rc = conv_tc_launch(dy, d->k, d->q, d->p, w_crsk, d->r * d->s * d->k, dx, tp, st, name, col_sum, col_sq);
      if (rc) return rc;
    }
  return 0;
}

// ------------------------------------------------------------------------------------------------
// wgrad for 3x3 stride-1 convolutions on the tensor cores.
//   dW[k, r, s, c] = sum over pixels  dy[pix, k] * x[pix shifted by (r-pad, s-pad), c]
// Per filter row r one CTA accumulates the three taps s=0..2 of a [128 filters x CB channels] block in
// TMEM (3*CB fp32 columns) while streaming boxes of 64 pixels (the GEMM K dimension):
//   A = dy box  [64 pixels x 128 filters]   (two 64-channel TMA boxes; the second is out of bounds and
//                                             therefore zero-filled when the layer has only 64 filters)
//   B = x box shifted by the tap [64 pixels x CB channels], three of them per stage
// Both operands are "MN-major" (channels contiguous, pixels strided): exactly how NHWC rows land in
// shared memory, so no transpose is needed; the UMMA descriptors say so (a_major = b_major = MN).
// Pixels are split across gridDim.x CTAs; partial sums go to fp32 scratch and are reduced in fixed order.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

struct TcWgradParams {
  int N, P, Q, Cin, Cout, pad, stride, R, S;
  int BW, BH, BNI, tiles_w, tiles_h, tiles_n;
  int chunks_per_split, total_chunks;
  int ncblk;
};

template <int CB, int S, int STAGES>
__global__ void __launch_bounds__(TC_THREADS) wgrad3x3_tc_kernel(const __grid_constant__ CUtensorMap map_dy,
                                                                 const __grid_constant__ CUtensorMap map_x,
                                                                 float* __restrict__ part, TcWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t BOX = 64 * 64 * 2;                 // one TMA box: 64 pixels x 64 channels bf16
  constexpr uint32_t A_BYTES = 2 * BOX;                 // 128 filters
  constexpr uint32_t B_TAP = (CB / 64) * BOX;           // CB channels
  constexpr uint32_t STAGE_BYTES = A_BYTES + S * B_TAP;
  constexpr uint32_t TMEM_COLS = (S * CB <= 64) ? 64 : (S * CB <= 128) ? 128 : (S * CB <= 256) ? 256 : 512;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x, r = blockIdx.y;
  const int kblk = blockIdx.z / p.ncblk, cblk = blockIdx.z - kblk * p.ncblk;
  const int chunk_lo = split * p.chunks_per_split;
  int chunk_hi = chunk_lo + p.chunks_per_split;
  if (chunk_hi > p.total_chunks) chunk_hi = p.total_chunks;
  const int nchunks = chunk_hi > chunk_lo ? chunk_hi - chunk_lo : 0;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    {   // warp-uniform loop, the elected lane issues (tc_common.cuh: elect_one); the box walks the pixel grid by counters
      const bool issuer = elect_one();
      if (issuer) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_dy)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
      }
      int stage = 0;
      uint32_t phase = 0;
      int t = chunk_lo;
      int tw = t % p.tiles_w; t /= p.tiles_w;
      int th = t % p.tiles_h;
      int tn = t / p.tiles_h;
      const int k0 = kblk * 128, c0 = cblk * CB, roff = r - p.pad;
      for (int ch = chunk_lo; ch < chunk_hi; ++ch) {
        const int n0 = tn * p.BNI, p0 = th * p.BH, q0 = tw * p.BW;
        uint8_t* st = smem + stage * STAGE_BYTES;
        mbar_wait(&empty[stage], phase ^ 1);
        if (issuer) {
          mbar_expect_tx(&full[stage], STAGE_BYTES);
          tma_load_4d(st, &map_dy, &full[stage], k0, q0, p0, n0);
          tma_load_4d(st + BOX, &map_dy, &full[stage], k0 + 64, q0, p0, n0);
#pragma unroll
          for (int s = 0; s < S; ++s)
#pragma unroll
            for (int j = 0; j < CB / 64; ++j)
              tma_load_4d(st + A_BYTES + s * B_TAP + j * BOX, &map_x, &full[stage], c0 + j * 64, q0 * p.stride + s - p.pad,
                          p0 * p.stride + roff, n0);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if (++tw == p.tiles_w) { tw = 0; if (++th == p.tiles_h) { th = 0; ++tn; } }
      }
    }
  } else if (warp == 1) {
    if (nchunks > 0) {
      const bool issuer = elect_one();
      // D=f32, A=B=bf16, A and B MN-major (bits 15, 16), N = CB, M = 128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(CB >> 3) << 17) |
                                 ((uint32_t)(128 >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      // descriptor = constant high word + low word (LBO << 16) | (address >> 4): stepped with 32-bit adds (tc_common.cuh)
      const uint64_t proto = umma_desc_mn_sw128(0, BOX);
      const uint32_t d_hi = (uint32_t)(proto >> 32), lo0 = (uint32_t)proto + (smem_u32(smem) >> 4);
      for (int i = 0; i < nchunks; ++i) {
        mbar_wait(&full[stage], phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_lo = lo0 + stage * (STAGE_BYTES >> 4);
        const uint32_t b_lo = a_lo + (A_BYTES >> 4);
        if (issuer) {
#pragma unroll
          for (int s = 0; s < S; ++s) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {   // 64 pixels = 4 x UMMA_K(16); 16 pixel rows = 2048 bytes
              umma_bf16_lohi(tmem_base + s * CB, a_lo + k * (2048 >> 4), d_hi, b_lo + s * (B_TAP >> 4) + k * (2048 >> 4), d_hi,
                             idesc, (uint32_t)((i | k) != 0));
            }
          }
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (issuer) umma_commit(tmem_full);
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int k = kblk * 128 + m;
    const int64_t rsc = (int64_t)p.R * S * p.Cin;
    float* dst = part + ((int64_t)split * p.Cout + k) * rsc + (int64_t)(r * S) * p.Cin + cblk * CB;
    if (nchunks > 0) {
      mbar_wait(tmem_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int c = 0; c < CB; c += 32) {
        uint32_t v[32];
        if (nchunks > 0) {
          tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(s * CB + c), v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (k < p.Cout) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(dst + (int64_t)s * p.Cin + c + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
  }
}

// out[i] (+)= sum over splits of part[z][i], in a fixed order (deterministic). A block is 64 float4 columns x 4 groups of
// splits: the small filters (layer1: 9216 float4 columns, 148 splits) otherwise leave 36 blocks walking 148 partials each.
// Four partials in flight per thread; the four group sums are added in order by the first group.
__global__ void __launch_bounds__(256) tc_splitk_reduce_kernel(const float* __restrict__ part, float* __restrict__ out,
                                                               int64_t n, int splits, int accumulate) {
  __shared__ float4 s_part[3][64];
  const int64_t n4 = n >> 2;
  const int lane_col = threadIdx.x & 63, zg = threadIdx.x >> 6;
  const int per = (splits + 3) >> 2;
  const int z0 = zg * per, z1 = min(splits, z0 + per);
  for (int64_t base = (int64_t)blockIdx.x * 64; base < n4; base += (int64_t)gridDim.x * 64) {   // block-uniform trip count
    const int64_t i = base + lane_col;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4) {
      int z = z0;
      for (; z + 4 <= z1; z += 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = reinterpret_cast<const float4*>(part + (int64_t)(z + u) * n)[i];
#pragma unroll
        for (int u = 0; u < 4; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
      }
      for (; z < z1; ++z) {
        float4 v = reinterpret_cast<const float4*>(part + (int64_t)z * n)[i];
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
    }
    if (zg > 0) s_part[zg - 1][lane_col] = s;
    __syncthreads();
    if (zg == 0 && i < n4) {
      float4 t = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      t.x += s.x; t.y += s.y; t.z += s.z; t.w += s.w;
#pragma unroll
      for (int g = 0; g < 3; ++g) {
        const float4 v = s_part[g][lane_col];
        t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
      }
      reinterpret_cast<float4*>(out)[i] = t;
    }
    __syncthreads();
  }
}

// few splits, many weights (layers 3-4): one thread per float4 column, four partials in flight
__global__ void __launch_bounds__(256) tc_splitk_reduce_flat_kernel(const float* __restrict__ part, float* __restrict__ out,
                                                                    int64_t n, int splits, int accumulate) {
  const int64_t n4 = n >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 s = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    int z = 0;
    for (; z + 4 <= splits; z += 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = reinterpret_cast<const float4*>(part + (int64_t)(z + u) * n)[i];
#pragma unroll
      for (int u = 0; u < 4; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
    for (; z < splits; ++z) {
      float4 v = reinterpret_cast<const float4*>(part + (int64_t)z * n)[i];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = s;
  }
}

// The reduction written straight into parameter gradients whose layout differs from the [k][r][s][c] weight-gradient matrix
// (the transformer's per-head [H, d_model, d_head] and [in][out] matrices): destination element idx of map m, decomposed by
// the map's shape, reads the matrix at offset + a*stride0 + b*stride1 + c*stride2 from every partial and ADDS the sum to
// dst[idx] (the optimizer's gradient buffer): no intermediate matrix, no strided addition by autograd.
__global__ void __launch_bounds__(256) tc_splitk_reduce_map_kernel(const float* __restrict__ part, int64_t n, int splits,
                                                                   const iit_wgrad_map_t* __restrict__ maps) {
  const iit_wgrad_map_t mp = maps[blockIdx.y];
  const int64_t numel = (int64_t)mp.size[0] * mp.size[1] * mp.size[2];
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < numel; idx += (int64_t)gridDim.x * 256) {
    const int c = (int)(idx % mp.size[2]);
    const int64_t t = idx / mp.size[2];
    const int b = (int)(t % mp.size[1]), a = (int)(t / mp.size[1]);
    const int64_t src = mp.offset + a * mp.stride[0] + b * mp.stride[1] + c * mp.stride[2];
    // four partials in flight per thread (a serial walk over the splits made this kernel 10 % of the IOI step: 13 us per launch
    // for a few MB), summed in the fixed order 0, 1, 2, ...
    float s = 0.f;
    int z = 0;
    for (; z + 4 <= splits; z += 4) {
      const float a = part[(int64_t)z * n + src], b = part[(int64_t)(z + 1) * n + src];
      const float c = part[(int64_t)(z + 2) * n + src], d = part[(int64_t)(z + 3) * n + src];
      s += a; s += b; s += c; s += d;
    }
    for (; z < splits; ++z) s += part[(int64_t)z * n + src];
    mp.dst[idx] += s;
  }
}

static void launch_splitk_reduce(const float* part, float* out, int64_t n, int splits, int accumulate, cudaStream_t st) {
  if (splits >= 16) tc_splitk_reduce_kernel<<<grid_for(n / 4, 64), 256, 0, st>>>(part, out, n, splits, accumulate);
  else tc_splitk_reduce_flat_kernel<<<grid_for(n / 4, 256), 256, 0, st>>>(part, out, n, splits, accumulate);
}

static void wgrad_tc_plan(const iit_conv_t* d, TcWgradParams* p, int* splits, int* CB) {
  p->N = d->n; p->P = d->p; p->Q = d->q; p->Cin = d->c; p->Cout = d->k; p->pad = d->pad;
  p->stride = d->stride; p->R = d->r; p->S = d->s;
  plan_box(d->p, d->q, 64, &p->BW, &p->BH, &p->BNI);
  p->tiles_w = (d->q + p->BW - 1) / p->BW;
  p->tiles_h = (d->p + p->BH - 1) / p->BH;
  p->tiles_n = (d->n + p->BNI - 1) / p->BNI;
  p->total_chunks = p->tiles_w * p->tiles_h * p->tiles_n;
  *CB = (d->c % 128 == 0) ? 128 : 64;
  p->ncblk = d->c / *CB;
  int groups = d->r * ((d->k + 127) / 128) * p->ncblk;
  // one CTA per SM (192 KB of stages, 512 TMEM columns): the grid must not exceed the SM count, or the few CTAs left over
  // run as a second wave and double the kernel time (50 x 3 = 150 CTAs did exactly that on layer2)
  int want = kNumSMs / groups;
  int maxs = (p->total_chunks + 7) / 8;   // at least 8 chunks (512 pixels) per CTA
  if (want > maxs) want = maxs;
  if (want < 1) want = 1;
  p->chunks_per_split = (p->total_chunks + want - 1) / want;
  *splits = (p->total_chunks + p->chunks_per_split - 1) / p->chunks_per_split;
}

template <int CB, int S>
static int launch_wgrad_tc(const CUtensorMap& mdy, const CUtensorMap& mx, float* part, const TcWgradParams& p, dim3 grid,
                           cudaStream_t st) {
  constexpr int STAGES = (S == 3 && CB == 128) ? 3 : 4;
  constexpr size_t smem = STAGES * (2 + S * (CB / 64)) * 8192 + (2 * STAGES + 1) * 8 + 16 + 1024;
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device_slot()];
  if (!attr_set) {
    IIT_CUDA(cudaFuncSetAttribute(wgrad3x3_tc_kernel<CB, S, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  wgrad3x3_tc_kernel<CB, S, STAGES><<<grid, TC_THREADS, smem, st>>>(mdy, mx, part, p);
  count_variant(KV_WGRAD_TC);
  return 0;
}

// conv_win.cu: shifted-window kernel for 3x3 / stride-1 / 64 -> 64 convolutions on small images
bool conv_win_3x3_supported(const iit_conv_t* d);
int conv_win_3x3(const void* in, const void* wmat, void* out, const iit_conv_t* d, bool flip, cudaStream_t st, double* bn_sum,
                 double* bn_sq, const char* name, bool accumulate = false, const float* scale = nullptr,
                 const float* shift = nullptr, const void* residual = nullptr, int relu = 0);
bool wgrad_win_supported(const iit_conv_t* d);
size_t wgrad_win_scratch_bytes(const iit_conv_t* d);
int wgrad_win(const void* dy, const void* x, float* part, const iit_conv_t* d, int* splits, cudaStream_t st, const char* name);
static bool g_use_win = true;

}  // namespace iit

using namespace iit;

extern "C" int iit_conv_pair_enable(int on) {
  const int was = g_pair_mode;
  g_pair_mode = on;
  return was;
}

extern "C" int iit_conv_win_enable(int on) {
  const int was = g_use_win ? 1 : 0;
  g_use_win = on != 0;
  return was;
}

extern "C" int iit_conv2d_tc_supported(const iit_conv_t* d, int dtype) {
  if (!d || dtype != IIT_BF16) return 0;
  const bool k3 = d->r == 3 && d->s == 3 && d->pad == 1;
  const bool k1 = d->r == 1 && d->s == 1 && d->pad == 0;
  return (k3 || k1) && (d->stride == 1 || d->stride == 2) && d->c % 64 == 0 && d->k % 64 == 0;
}

extern "C" int iit_conv2d_fprop_tc(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, double* bn_sum,
                                   double* bn_sqsum, void* stream) {
  IIT_REQUIRE(x && w_krsc && y && d, "iit_conv2d_fprop_tc: null pointer");
  IIT_REQUIRE((bn_sum == nullptr) == (bn_sqsum == nullptr), "iit_conv2d_fprop_tc: bn_sum and bn_sqsum go together");
  IIT_REQUIRE(iit_conv2d_tc_supported(d, IIT_BF16), "iit_conv2d_fprop_tc: only bf16 3x3/pad-1 or 1x1/pad-0 convolutions, stride 1 "
              "or 2, with channel counts that are multiples of 64");
  if (g_use_win && conv_win_3x3_supported(d)) {
    if (check_tc_ptrs(x, w_krsc, y, "iit_conv2d_fprop_tc")) return 1;
    return conv_win_3x3(x, w_krsc, y, d, false, as_stream(stream), bn_sum, bn_sqsum, "iit_conv2d_fprop_tc");
  }
  return conv_tc_fprop(x, w_krsc, y, d, as_stream(stream), bn_sum, bn_sqsum);
}

// y = act(conv(x, w) * scale[k] + shift[k] (+ residual)): an eval-mode BatchNorm (and the block's residual addition and
// ReLU) folded into the convolution's epilogue, on the fp32 accumulator
extern "C" int iit_conv2d_fprop_tc_affine(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, const float* scale,
                                          const float* shift, const void* residual, int relu, void* stream) {
  const char* name = "iit_conv2d_fprop_tc_affine";
  IIT_REQUIRE(x && w_krsc && y && d && scale && shift, "%s: null pointer", name);
  IIT_REQUIRE(iit_conv2d_tc_supported(d, IIT_BF16), "%s: only bf16 3x3/pad-1 or 1x1/pad-0 convolutions, stride 1 or 2, with channel "
              "counts that are multiples of 64", name);
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(scale) | reinterpret_cast<uintptr_t>(shift) | reinterpret_cast<uintptr_t>(residual)) & 15) == 0,
              "%s: scale, shift and residual must be 16-byte aligned", name);
  if (check_tc_ptrs(x, w_krsc, y, name)) return 1;
  if (g_use_win && conv_win_3x3_supported(d))
    return conv_win_3x3(x, w_krsc, y, d, false, as_stream(stream), nullptr, nullptr, name, false, scale, shift, residual, relu);
  TcParams e{};
  e.epi = TC_EPI_BIAS; e.scale = scale; e.bias = shift; e.aux = (const __nv_bfloat16*)residual; e.relu = relu;
  return conv_tc_fprop(x, w_krsc, y, d, as_stream(stream), nullptr, nullptr, &e, name);
}

static int dgrad_tc_entry(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, void* stream, bool accumulate,
                          const char* name) {
  IIT_REQUIRE(dy && w_crsk && dx && d, "%s: null pointer", name);
  IIT_REQUIRE(iit_conv2d_tc_supported(d, IIT_BF16), "%s: only bf16 3x3/pad-1 or 1x1/pad-0 convolutions, stride 1 "
              "or 2, with channel counts that are multiples of 64", name);
  if (g_use_win && conv_win_3x3_supported(d)) {
    if (check_tc_ptrs(dy, w_crsk, dx, name)) return 1;
    return conv_win_3x3(dy, w_crsk, dx, d, true, as_stream(stream), nullptr, nullptr, name, accumulate);
  }
  return conv_tc_dgrad(dy, w_crsk, dx, d, as_stream(stream), accumulate);
}

extern "C" int iit_conv2d_dgrad_tc(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, void* stream) {
  return dgrad_tc_entry(dy, w_crsk, dx, d, stream, false, "iit_conv2d_dgrad_tc");
}

extern "C" int iit_conv2d_dgrad_tc_accumulate(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, void* stream) {
  return dgrad_tc_entry(dy, w_crsk, dx, d, stream, true, "iit_conv2d_dgrad_tc_accumulate");
}

// ---- the interchange patch inside the producing convolution (north_star (a)) ---------------------------------------------
extern "C" int iit_conv2d_patch_fusable(const iit_conv_t* d) {
  // the tiled kernel carries the epilogue; the shifted-window kernel (64 -> 64 on small images, the stem) does not
  return d && iit_conv2d_tc_supported(d, IIT_BF16) && !(g_use_win && conv_win_3x3_supported(d)) ? 1 : 0;
}

extern "C" int iit_conv2d_fprop_tc_patch(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, double* bn_sum,
                                         double* bn_sqsum, const void* src, const int32_t* box6, void* stream) {
  const char* name = "iit_conv2d_fprop_tc_patch";
  IIT_REQUIRE(x && w_krsc && y && d && src && box6, "%s: null pointer", name);
  IIT_REQUIRE((bn_sum == nullptr) == (bn_sqsum == nullptr), "%s: bn_sum and bn_sqsum come together", name);
  IIT_REQUIRE(iit_conv2d_patch_fusable(d), "%s: this convolution runs on a kernel without the patch epilogue", name);
  IIT_REQUIRE(box6[0] >= 0 && box6[1] <= d->p && box6[2] >= 0 && box6[3] <= d->q && box6[4] >= 0 && box6[5] <= d->k &&
              box6[0] <= box6[1] && box6[2] <= box6[3] && box6[4] <= box6[5], "%s: box outside the output tensor", name);
  IIT_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0, "%s: pointers must be 16-byte aligned", name);
  TcParams e{};
  e.epi = TC_EPI_PATCH;
  e.aux = (const __nv_bfloat16*)src;
  for (int i = 0; i < 6; ++i) e.box[i] = box6[i];
  return conv_tc_fprop(x, w_krsc, y, d, as_stream(stream), bn_sum, bn_sqsum, &e, name);
}

// ---- the transformer's linear maps with their element-wise neighbours in the GEMM epilogue ------------------------------
static void gemm_desc(iit_conv_t* d, int64_t rows, int n_in, int n_out) {
  d->n = (int)rows; d->h = 1; d->w = 1; d->c = n_in; d->k = n_out; d->r = 1; d->s = 1; d->stride = 1; d->pad = 0; d->p = 1; d->q = 1;
}

extern "C" int iit_gemm_bias_act_tc(const void* x, const void* w_oi, const float* bias, const void* resid, void* y, void* pre_out,
                                    int64_t rows, int n_in, int n_out, int act, void* stream) {
  const char* name = "iit_gemm_bias_act_tc";
  IIT_REQUIRE(x && w_oi && y && rows > 0 && rows < (1ll << 31), "%s: bad arguments", name);
  IIT_REQUIRE(n_in % 64 == 0 && n_out % 64 == 0, "%s: widths must be multiples of 64 (got %d -> %d)", name, n_in, n_out);
  IIT_REQUIRE(act == 0 || act == 1, "%s: act must be 0 (identity) or 1 (gelu_new)", name);
  IIT_REQUIRE(!(act == 1 && resid != nullptr), "%s: a residual cannot be combined with the activation", name);
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(bias) | reinterpret_cast<uintptr_t>(resid) | reinterpret_cast<uintptr_t>(pre_out)) & 15) == 0,
              "%s: pointers must be 16-byte aligned", name);
  iit_conv_t d;
  gemm_desc(&d, rows, n_in, n_out);
  TcParams e{};
  e.epi = act == 1 ? TC_EPI_GELU : TC_EPI_BIAS;
  e.bias = bias; e.aux = (const __nv_bfloat16*)resid; e.out2 = (__nv_bfloat16*)pre_out;
  return conv_tc_fprop(x, w_oi, y, &d, as_stream(stream), nullptr, nullptr, &e, name);
}

extern "C" int iit_gemm_dgrad_gelu_tc(const void* dy, const void* w_io, const void* pre, void* dpre, double* col_sum,
                                      double* col_sq, int64_t rows, int n_in, int n_out, void* stream) {
  const char* name = "iit_gemm_dgrad_gelu_tc";
  IIT_REQUIRE(dy && w_io && pre && dpre && rows > 0 && rows < (1ll << 31), "%s: bad arguments", name);
  IIT_REQUIRE(n_in % 64 == 0 && n_out % 64 == 0, "%s: widths must be multiples of 64 (got %d -> %d)", name, n_in, n_out);
  IIT_REQUIRE((col_sum == nullptr) == (col_sq == nullptr), "%s: col_sum and col_sq come together", name);
  IIT_REQUIRE((reinterpret_cast<uintptr_t>(pre) & 15) == 0, "%s: pointers must be 16-byte aligned", name);
  iit_conv_t d;
  gemm_desc(&d, rows, n_in, n_out);
  TcParams e{};
  e.epi = TC_EPI_GELU_BWD;
  e.aux = (const __nv_bfloat16*)pre;
  return conv_tc_dgrad(dy, w_io, dpre, &d, as_stream(stream), false, &e, col_sum, col_sq, name);
}

extern "C" size_t iit_conv2d_wgrad_tc_scratch_bytes(const iit_conv_t* d) {
  if (!d || !iit_conv2d_tc_supported(d, IIT_BF16)) return 0;
  TcWgradParams p;
  int splits, CB;
  wgrad_tc_plan(d, &p, &splits, &CB);
  size_t need = (size_t)splits * d->k * d->r * d->s * d->c * sizeof(float);
  const size_t win = wgrad_win_scratch_bytes(d);     // enough for either kernel, whichever the switch selects
  return need > win ? need : win;
}

static int wgrad_tc_impl(const void* dy, const void* x, float* dw, const iit_conv_t* d, void* scratch, size_t scratch_bytes,
                         void* stream, int accumulate, const iit_wgrad_map_t* maps = nullptr, int n_maps = 0,
                         int64_t max_map_numel = 0) {
  if (maps != nullptr) dw = reinterpret_cast<float*>(scratch);     // unused placeholder that passes the pointer checks
  IIT_REQUIRE(dy && x && dw && d, "iit_conv2d_wgrad_tc: null pointer");
  IIT_REQUIRE(iit_conv2d_tc_supported(d, IIT_BF16), "iit_conv2d_wgrad_tc: only bf16 3x3/pad-1 or 1x1/pad-0 convolutions, stride 1 "
              "or 2, with channel counts that are multiples of 64");
  IIT_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 15) == 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0,
              "iit_conv2d_wgrad_tc: scratch and dw must be 16-byte aligned");
  if (g_use_win && wgrad_win_supported(d)) {
    const size_t need_win = wgrad_win_scratch_bytes(d);
    IIT_REQUIRE(scratch && scratch_bytes >= need_win, "iit_conv2d_wgrad_tc: scratch too small: need %zu bytes, got %zu", need_win,
                scratch_bytes);
    IIT_REQUIRE(((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x)) & 15) == 0,
                "iit_conv2d_wgrad_tc: dy and x must be 16-byte aligned");
    int nsplit = 0;
    int rc = wgrad_win(dy, x, (float*)scratch, d, &nsplit, as_stream(stream), "iit_conv2d_wgrad_tc");
    if (rc) return rc;
    const int64_t n = (int64_t)d->k * d->r * d->s * d->c;
    if (maps != nullptr)
      tc_splitk_reduce_map_kernel<<<dim3((unsigned)grid_for(max_map_numel, 256, 16), (unsigned)n_maps), 256, 0, as_stream(stream)>>>(
          (const float*)scratch, n, nsplit, maps);
    else
      launch_splitk_reduce((const float*)scratch, dw, n, nsplit, accumulate, as_stream(stream));
    IIT_LAUNCH_CHECK("iit_conv2d_wgrad_tc(reduce)");
    return 0;
  }
  TcWgradParams p;
  int splits, CB;
  wgrad_tc_plan(d, &p, &splits, &CB);
  size_t need = (size_t)splits * d->k * d->r * d->s * d->c * sizeof(float);
  IIT_REQUIRE(scratch && scratch_bytes >= need, "iit_conv2d_wgrad_tc: scratch too small: need %zu bytes, got %zu", need, scratch_bytes);
  CUtensorMap mdy, mx;
  if (encode_nhwc_strided(&mdy, dy, d->k, d->q, d->p, d->n, p.BW, p.BH, p.BNI, 1, "iit_conv2d_wgrad_tc(dy)")) return 1;
  if (encode_nhwc_strided(&mx, x, d->c, d->w, d->h, d->n, p.BW, p.BH, p.BNI, d->stride, "iit_conv2d_wgrad_tc(x)")) return 1;
  cudaStream_t st = as_stream(stream);
  dim3 grid((unsigned)splits, (unsigned)d->r, (unsigned)(((d->k + 127) / 128) * p.ncblk));
  int rc;
  if (d->s == 3) rc = CB == 128 ? launch_wgrad_tc<128, 3>(mdy, mx, (float*)scratch, p, grid, st)
                                : launch_wgrad_tc<64, 3>(mdy, mx, (float*)scratch, p, grid, st);
  else rc = CB == 128 ? launch_wgrad_tc<128, 1>(mdy, mx, (float*)scratch, p, grid, st)
                      : launch_wgrad_tc<64, 1>(mdy, mx, (float*)scratch, p, grid, st);
  if (rc) return rc;
  IIT_LAUNCH_CHECK("iit_conv2d_wgrad_tc");
  int64_t n = (int64_t)d->k * d->r * d->s * d->c;
  if (maps != nullptr)
    tc_splitk_reduce_map_kernel<<<dim3((unsigned)grid_for(max_map_numel, 256, 16), (unsigned)n_maps), 256, 0, st>>>(
        (const float*)scratch, n, splits, maps);
  else
    launch_splitk_reduce((const float*)scratch, dw, n, splits, accumulate, st);
  IIT_LAUNCH_CHECK("iit_conv2d_wgrad_tc(reduce)");
  return 0;
}

extern "C" int iit_conv2d_wgrad_tc_scatter(const void* dy, const void* x, const iit_conv_t* d, void* scratch, size_t scratch_bytes,
                                           const iit_wgrad_map_t* maps_dev, int n_maps, int64_t max_map_numel, void* stream) {
  IIT_REQUIRE(maps_dev && n_maps >= 1 && n_maps <= 16 && max_map_numel >= 1, "iit_conv2d_wgrad_tc_scatter: 1..16 maps (device array) expected");
  return wgrad_tc_impl(dy, x, nullptr, d, scratch, scratch_bytes, stream, 1, maps_dev, n_maps, max_map_numel);
}

extern "C" int iit_conv2d_wgrad_tc(const void* dy, const void* x, float* dw, const iit_conv_t* d, void* scratch,
                                   size_t scratch_bytes, void* stream) {
  return wgrad_tc_impl(dy, x, dw, d, scratch, scratch_bytes, stream, 0);
}

extern "C" int iit_conv2d_wgrad_tc_accumulate(const void* dy, const void* x, float* dw, const iit_conv_t* d, void* scratch,
                                              size_t scratch_bytes, void* stream) {
  return wgrad_tc_impl(dy, x, dw, d, scratch, scratch_bytes, stream, 1);
}
