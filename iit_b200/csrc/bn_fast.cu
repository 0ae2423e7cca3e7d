// (b) BatchNorm kernels, 16-byte edition, and the fused stem tail (BN + ReLU + MaxPool 3x3/2).
//
// Same arithmetic as the kernels in layers.cu, restructured for HBM throughput:
//   * every thread moves 16 bytes per access (8 bf16 / 4 fp32 channels of one pixel);
//   * the launch geometry makes a thread's channel group loop-invariant, so the per-channel constants
//     (mean, 1/std, gamma, beta, the two backward sums) sit in registers as fp32 - no per-element table reads
//     and no fp64->fp32 conversions in the loop (fp64 issue rate on B200 is 1/64 of fp32);
//   * the stem's BN -> ReLU -> MaxPool chain is one kernel forward (reads the 28x28 conv output once, writes
//     the 14x14 pooled tensor and the window index) and two kernels backward (reduce, apply) that recompute
//     the normalised activation from the conv output instead of reading stored intermediates.
// layers.cu's entry points call the *_fast launchers when `bn_fast_ok` holds and keep their generic
// kernels otherwise.
#include <type_traits>

#include "common.cuh"

namespace iit {

template <typename T> struct Pack;
template <> struct Pack<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void unpack(const uint4& t, float (&v)[4]) {
    v[0] = __uint_as_float(t.x); v[1] = __uint_as_float(t.y); v[2] = __uint_as_float(t.z); v[3] = __uint_as_float(t.w);
  }
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) { unpack(*reinterpret_cast<const uint4*>(p), v); }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Pack<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) { unpack(*reinterpret_cast<const uint4*>(p), v); }
  static __device__ __forceinline__ void unpack(const uint4& t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {   // bf16 -> fp32 is a 16-bit shift
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <typename T> __device__ __forceinline__ float round_to(float v) { return to_f<T>(from_f<T>(v)); }

template <int NP>
__device__ __forceinline__ void load_consts(const float* __restrict__ a, int c0, float (&v)[NP]) {
#pragma unroll
  for (int j = 0; j < NP; j += 4) {
    float4 t = *reinterpret_cast<const float4*>(a + c0 + j);
    v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
  }
}

// ---- forward: y = [relu]((x - mean) * invstd * gamma + beta [+ res]) ---------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) bn_act_fwd_fast_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                                                              int64_t npacks, int C, const float* __restrict__ mean,
                                                              const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, int relu) {
  constexpr int NP = Pack<T>::N;
  const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, stride = (int64_t)gridDim.x * 256;
  const int c0 = (int)((tid * NP) % C);   // invariant: stride * NP is a multiple of C
  float mu[NP], is[NP], g[NP], b[NP];
  load_consts<NP>(mean, c0, mu); load_consts<NP>(invstd, c0, is); load_consts<NP>(gamma, c0, g); load_consts<NP>(beta, c0, b);
  for (int64_t i = tid; i < npacks; i += stride) {
    float v[NP], o[NP];
    Pack<T>::load(x + i * NP, v);
#pragma unroll
    for (int j = 0; j < NP; ++j) o[j] = (v[j] - mu[j]) * is[j] * g[j] + b[j];
    if (res) {
      float r[NP];
      Pack<T>::load(res + i * NP, r);
#pragma unroll
      for (int j = 0; j < NP; ++j) o[j] += r[j];
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < NP; ++j) o[j] = fmaxf(o[j], 0.f);
    }
    Pack<T>::store(y + i * NP, o);
  }
}

// ---- per-channel reductions: block = CGB channel groups x RL row lanes ------------------------------------------
template <int NRED, int NP>
__device__ __forceinline__ void reduce_finish(float (&acc)[NRED][NP], int cg, int rl, int CGB, int RL, int c0, double* const* outs,
                                              float* smem) {
  const int W = CGB * NP;   // smem layout [NRED][RL][W]
#pragma unroll
  for (int r = 0; r < NRED; ++r)
#pragma unroll
    for (int j = 0; j < NP; ++j) smem[(r * RL + rl) * W + cg * NP + j] = acc[r][j];
  __syncthreads();
  for (int i = threadIdx.x; i < NRED * W; i += blockDim.x) {
    const int r = i / W, ch = i - r * W;
    float s = 0.f;
    for (int l = 0; l < RL; ++l) s += smem[(r * RL + l) * W + ch];
    atomicAdd(&outs[r][c0 + ch], (double)s);
  }
}

struct Red2 { double* p[2]; };

template <typename T>
__global__ void __launch_bounds__(256) bn_stats_fast_kernel(const T* __restrict__ x, int64_t rows, int C, int CGB, Red2 o) {
  extern __shared__ float smem[];
  constexpr int NP = Pack<T>::N;
  const int RL = 256 / CGB;
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int cbase = blockIdx.y * CGB * NP, c = cbase + cg * NP;
  float acc[2][NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  const int64_t step = (int64_t)gridDim.x * RL;
  int64_t r = (int64_t)blockIdx.x * RL + rl;
  for (; r + 3 * step < rows; r += 4 * step) {   // four independent 16-byte loads in flight per thread
    float v[4][NP];
#pragma unroll
    for (int u = 0; u < 4; ++u) Pack<T>::load(x + (r + u * step) * C + c, v[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int j = 0; j < NP; ++j) { acc[0][j] += v[u][j]; acc[1][j] += v[u][j] * v[u][j]; }
  }
  for (; r < rows; r += step) {
    float v[NP];
    Pack<T>::load(x + r * C + c, v);
#pragma unroll
    for (int j = 0; j < NP; ++j) { acc[0][j] += v[j]; acc[1][j] += v[j] * v[j]; }
  }
  double* outs[2] = {o.p[0], o.p[1]};
  reduce_finish<2, NP>(acc, cg, rl, CGB, RL, cbase, outs, smem);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_reduce_fast_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                 const T* __restrict__ x, int64_t rows, int C, int CGB,
                                                                 const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                 int relu, Red2 o) {
  extern __shared__ float smem[];
  constexpr int NP = Pack<T>::N;
  const int RL = 256 / CGB;
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int cbase = blockIdx.y * CGB * NP, c = cbase + cg * NP;
  float mu[NP], is[NP], acc[2][NP];
  load_consts<NP>(mean, c, mu); load_consts<NP>(invstd, c, is);
#pragma unroll
  for (int j = 0; j < NP; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  const int64_t step = (int64_t)gridDim.x * RL;
  int64_t r = (int64_t)blockIdx.x * RL + rl;
  for (; r + step < rows; r += 2 * step) {      // two rows (six 16-byte loads) in flight per thread
    uint4 rg[2], rx[2], ry[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      rg[u] = *reinterpret_cast<const uint4*>(dy + (r + u * step) * C + c);
      rx[u] = *reinterpret_cast<const uint4*>(x + (r + u * step) * C + c);
      if (relu) ry[u] = *reinterpret_cast<const uint4*>(y + (r + u * step) * C + c);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float g[NP], xv[NP];
      Pack<T>::unpack(rg[u], g);
      Pack<T>::unpack(rx[u], xv);
      if (relu) {
        float yv[NP];
        Pack<T>::unpack(ry[u], yv);
#pragma unroll
        for (int j = 0; j < NP; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < NP; ++j) { acc[0][j] += g[j] * (xv[j] - mu[j]) * is[j]; acc[1][j] += g[j]; }
    }
  }
  for (; r < rows; r += step) {
    float g[NP], xv[NP];
    Pack<T>::load(dy + r * C + c, g);
    Pack<T>::load(x + r * C + c, xv);
    if (relu) {
      float yv[NP];
      Pack<T>::load(y + r * C + c, yv);
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < NP; ++j) { acc[0][j] += g[j] * (xv[j] - mu[j]) * is[j]; acc[1][j] += g[j]; }
  }
  double* outs[2] = {o.p[0], o.p[1]};
  reduce_finish<2, NP>(acc, cg, rl, CGB, RL, cbase, outs, smem);
}

// dx = gamma * invstd * (g - dbeta/rows - xhat * dgamma/rows),  g = relu'(y) * dy;  dres = g
template <typename T>
__global__ void __launch_bounds__(256, 4) bn_bwd_apply_fast_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                const T* __restrict__ x, T* __restrict__ dx, T* __restrict__ dres,
                                                                int64_t npacks, int C, float inv_rows, const float* __restrict__ mean,
                                                                const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                                const double* __restrict__ dgamma, const double* __restrict__ dbeta,
                                                                int relu) {
  constexpr int NP = Pack<T>::N;
  const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, stride = (int64_t)gridDim.x * 256;
  const int c0 = (int)((tid * NP) % C);
  float mu[NP], is[NP], k[NP], dbn[NP], dgn[NP];
  load_consts<NP>(mean, c0, mu); load_consts<NP>(invstd, c0, is); load_consts<NP>(gamma, c0, k);
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    k[j] *= is[j];
    dbn[j] = (float)dbeta[c0 + j] * inv_rows;
    dgn[j] = (float)dgamma[c0 + j] * inv_rows;
  }
  for (int64_t i = tid; i < npacks; i += stride) {
    float g[NP], xv[NP], o[NP];
    Pack<T>::load(dy + i * NP, g);
    Pack<T>::load(x + i * NP, xv);
    if (relu) {
      float yv[NP];
      Pack<T>::load(y + i * NP, yv);
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
    if (dres) Pack<T>::store(dres + i * NP, g);
#pragma unroll
    for (int j = 0; j < NP; ++j) o[j] = k[j] * (g[j] - dbn[j] - (xv[j] - mu[j]) * is[j] * dgn[j]);
    Pack<T>::store(dx + i * NP, o);
  }
}

// ---- fused stem tail: BN -> ReLU -> MaxPool2d(3, 2, 1) -----------------------------------------------------------
// forward: one thread = one pooled pixel x one channel pack
template <typename T>
__global__ void __launch_bounds__(256) bn_relu_pool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx,
                                                               int N, int H, int W, int C, int P, int Q,
                                                               const float* __restrict__ mean, const float* __restrict__ invstd,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta) {
  constexpr int NP = Pack<T>::N;
  const uint32_t CPK = C / NP, PXB = 256 / CPK;          // channel packs per pixel, pixels per block pass
  const uint32_t cp = threadIdx.x % CPK, px = threadIdx.x / CPK;
  float mu[NP], is[NP], g[NP], b[NP];
  load_consts<NP>(mean, cp * NP, mu); load_consts<NP>(invstd, cp * NP, is); load_consts<NP>(gamma, cp * NP, g);
  load_consts<NP>(beta, cp * NP, b);
  float sgn[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) sgn[j] = (is[j] * g[j] < 0.f) ? -1.f : 1.f;
  uint32_t sgn2[NP / 2];     // the same as packed bf16 pairs (0x3f80 = 1, 0xbf80 = -1)
#pragma unroll
  for (int j = 0; j < NP / 2; ++j)
    sgn2[j] = (sgn[2 * j] < 0.f ? 0xbf80u : 0x3f80u) | ((sgn[2 * j + 1] < 0.f ? 0xbf80u : 0x3f80u) << 16);
  const uint32_t total = (uint32_t)N * P * Q;            // 32-bit index arithmetic: 64-bit divisions cost ~100 instructions
  // (n, p, q) follows pix incrementally: a division costs more than the rest of the index arithmetic together
  const uint32_t stride = gridDim.x * PXB, first = blockIdx.x * PXB + px;
  const int sq = (int)(stride % Q), sp = (int)((stride / Q) % P), sn = (int)(stride / ((uint32_t)Q * P));
  int q = (int)(first % Q), p = (int)((first / Q) % P), n = (int)(first / ((uint32_t)Q * P));
  for (uint32_t pix = first; pix < total; pix += stride, q += sq, p += sp, n += sn) {
    if (q >= Q) { q -= Q; ++p; }
    if (p >= P) { p -= P; ++n; }
    // all nine window loads are issued before any is used (clamped addresses, validity applied afterwards):
    // one memory round trip per pooled pixel instead of nine dependent ones
    uint4 raw[9];
    bool valid[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = p * 2 - 1 + r;
      const int ihc = min(max(ih, 0), H - 1);
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int iw = q * 2 - 1 + s;
        const int iwc = min(max(iw, 0), W - 1);
        valid[r * 3 + s] = ih >= 0 && ih < H && iw >= 0 && iw < W;
        raw[r * 3 + s] = *reinterpret_cast<const uint4*>(x + (((int64_t)n * H + ihc) * W + iwc) * C + cp * NP);
      }
    }
    // BN is a monotonic map per channel (increasing when gamma*invstd >= 0, decreasing otherwise), ReLU and the
    // rounding to T are monotonic too: the window winner is the largest (smallest) RAW value, found with one
    // compare per element; the affine map, ReLU and rounding are applied to the nine winners only. The pooled value
    // is bit-identical to pooling the rounded activations; among activations that tie after ReLU / rounding the
    // winner is the one with the largest pre-rounding value (the fp32 reference's choice).
    float best[NP];
    uint32_t ipk[NP / 4];      // the winners' window slots, one byte per channel
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      // two channels per instruction on packed bf16: key = raw * (+-1), running maximum with HMNMX2, then the first
      // tap equal to the maximum (scanning backwards so that the earliest one is kept)
      const uint32_t NEG_INF2 = 0xff80ff80u;
      uint32_t key[9][4], best2[4] = {NEG_INF2, NEG_INF2, NEG_INF2, NEG_INF2};
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const uint32_t w4[4] = {raw[t].x, raw[t].y, raw[t].z, raw[t].w};
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          __nv_bfloat162 k2 = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&w4[w]), *reinterpret_cast<const __nv_bfloat162*>(&sgn2[w]));
          uint32_t kb = *reinterpret_cast<uint32_t*>(&k2);
          kb = valid[t] ? kb : NEG_INF2;
          key[t][w] = kb;
          __nv_bfloat162 m2 = __hmax2_nan(*reinterpret_cast<__nv_bfloat162*>(&best2[w]), *reinterpret_cast<__nv_bfloat162*>(&kb));
          best2[w] = *reinterpret_cast<uint32_t*>(&m2);
        }
      }
      uint32_t idx2[4] = {0u, 0u, 0u, 0u};
#pragma unroll
      for (int t = 8; t >= 0; --t) {
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint32_t m = __heq2_mask(*reinterpret_cast<__nv_bfloat162*>(&key[t][w]), *reinterpret_cast<__nv_bfloat162*>(&best2[w]));
          idx2[w] = (idx2[w] & ~m) | ((uint32_t)(t | (t << 16)) & m);
        }
      }
      ipk[0] = __byte_perm(idx2[0], idx2[1], 0x6420);
      ipk[1] = __byte_perm(idx2[2], idx2[3], 0x6420);
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        __nv_bfloat162 v2 = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&best2[w]), *reinterpret_cast<const __nv_bfloat162*>(&sgn2[w]));
        const uint32_t vb = *reinterpret_cast<uint32_t*>(&v2);
        best[2 * w] = __uint_as_float(vb << 16);
        best[2 * w + 1] = __uint_as_float(vb & 0xffff0000u);
      }
    } else {
      float bkey[NP];
      int bi[NP];
#pragma unroll
      for (int j = 0; j < NP; ++j) { bkey[j] = -INFINITY; bi[j] = 0; }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        float v[NP];
        Pack<T>::unpack(raw[t], v);
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          const float key = v[j] * sgn[j];
          const bool take = valid[t] & !(key <= bkey[j]);   // unordered compare: a NaN is taken and then kept
          bkey[j] = take ? key : bkey[j];
          bi[j] = take ? t : bi[j];
        }
      }
#pragma unroll
      for (int j = 0; j < NP; ++j) best[j] = bkey[j] * sgn[j];
#pragma unroll
      for (int j = 0; j < NP; j += 4) ipk[j / 4] = (uint32_t)bi[j] | ((uint32_t)bi[j + 1] << 8) | ((uint32_t)bi[j + 2] << 16) | ((uint32_t)bi[j + 3] << 24);
    }
#pragma unroll
    for (int j = 0; j < NP; ++j) best[j] = fmaxf((best[j] - mu[j]) * is[j] * g[j] + b[j], 0.f);
    Pack<T>::store(y + (int64_t)pix * C + cp * NP, best);
    uint32_t* ip = reinterpret_cast<uint32_t*>(idx + (int64_t)pix * C + cp * NP);
#pragma unroll
    for (int j = 0; j < NP / 4; ++j) ip[j] = ipk[j];
  }
}

// gradient reaching the ReLU output at input pixel (h, w): sum of the pooled gradients whose window picked it.
// The (up to) 2 x 2 candidate windows are loaded together from clamped addresses; invalid ones are masked afterwards.
template <typename T>
__device__ __forceinline__ void pool_bwd_gather(const T* __restrict__ dy, const int8_t* __restrict__ idx, int n, int h, int w, int c,
                                                int C, int P, int Q, float (&g)[Pack<T>::N]) {
  constexpr int NP = Pack<T>::N;
  const int pc[2] = {h >> 1, (h + 1) >> 1}, qc[2] = {w >> 1, (w + 1) >> 1};
  uint4 raw[4];
  uint32_t kraw[4][NP / 4];
  int slot[4];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int t = a * 2 + b;
      const bool ok = pc[a] < P && qc[b] < Q && (a == 0 || pc[1] != pc[0]) && (b == 0 || qc[1] != qc[0]);
      const int p = min(pc[a], P - 1), q = min(qc[b], Q - 1);
      slot[t] = ok ? (h - (pc[a] * 2 - 1)) * 3 + (w - (qc[b] * 2 - 1)) : -1;   // -1 never matches a stored slot
      const int64_t op = (((int64_t)n * P + p) * Q + q) * C + c;
      raw[t] = *reinterpret_cast<const uint4*>(dy + op);
#pragma unroll
      for (int j = 0; j < NP / 4; ++j) kraw[t][j] = *reinterpret_cast<const uint32_t*>(idx + op + 4 * j);
    }
  if constexpr (std::is_same<T, __nv_bfloat16>::value) {
    // four channels per compare (__vcmpeq4 on the slot bytes), two per masked packed add; an element is the winner of
    // more than one window only occasionally, so the bf16 additions are mostly exact
    uint32_t g2[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint32_t s4 = (uint32_t)(slot[t] & 0xff) * 0x01010101u;
      const uint32_t w4[4] = {raw[t].x, raw[t].y, raw[t].z, raw[t].w};
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        const uint32_t m4 = __vcmpeq4(kraw[t][hlf], s4);
        const uint32_t mlo = __byte_perm(m4, 0u, 0x1100), mhi = __byte_perm(m4, 0u, 0x3322);
        uint32_t a = w4[2 * hlf] & mlo, b = w4[2 * hlf + 1] & mhi;
        __nv_bfloat162 sa = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&g2[2 * hlf]), *reinterpret_cast<__nv_bfloat162*>(&a));
        __nv_bfloat162 sb = __hadd2(*reinterpret_cast<__nv_bfloat162*>(&g2[2 * hlf + 1]), *reinterpret_cast<__nv_bfloat162*>(&b));
        g2[2 * hlf] = *reinterpret_cast<uint32_t*>(&sa);
        g2[2 * hlf + 1] = *reinterpret_cast<uint32_t*>(&sb);
      }
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) { g[2 * w] = __uint_as_float(g2[w] << 16); g[2 * w + 1] = __uint_as_float(g2[w] & 0xffff0000u); }
  } else {
#pragma unroll
    for (int j = 0; j < NP; ++j) g[j] = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      float d[NP];
      Pack<T>::unpack(raw[t], d);
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const int k = (int)(int8_t)((kraw[t][j >> 2] >> (8 * (j & 3))) & 0xff);
        if (k == slot[t]) g[j] += d[j];
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256, 4) bn_relu_pool_bwd_reduce_kernel(const T* __restrict__ dy, const int8_t* __restrict__ idx,
                                                                      const T* __restrict__ x, int N, int H, int W, int C, int P,
                                                                      int Q, int CGB, const float* __restrict__ mean,
                                                                      const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                                      const float* __restrict__ beta, Red2 o) {
  extern __shared__ float smem[];
  constexpr int NP = Pack<T>::N;
  const int RL = 256 / CGB;                               // CGB == C / NP here: one block covers every channel
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int cbase = 0, c = cg * NP;
  float mu[NP], is[NP], gm[NP], bt[NP], acc[2][NP];
  load_consts<NP>(mean, c, mu); load_consts<NP>(invstd, c, is); load_consts<NP>(gamma, c, gm); load_consts<NP>(beta, c, bt);
#pragma unroll
  for (int j = 0; j < NP; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  const uint32_t rows = (uint32_t)N * H * W;
  const uint32_t stride = gridDim.x * RL, first = blockIdx.x * RL + rl;
  const int sw = (int)(stride % W), sh = (int)((stride / W) % H), sn = (int)(stride / ((uint32_t)W * H));
  int w = (int)(first % W), h = (int)((first / W) % H), n = (int)(first / ((uint32_t)W * H));
  for (uint32_t r = first; r < rows; r += stride, w += sw, h += sh, n += sn) {
    if (w >= W) { w -= W; ++h; }
    if (h >= H) { h -= H; ++n; }
    float g[NP], xv[NP];
    Pack<T>::load(x + (int64_t)r * C + c, xv);
    pool_bwd_gather<T>(dy, idx, n, h, w, c, C, P, Q, g);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const float xh = (xv[j] - mu[j]) * is[j];
      const float gj = (xh * gm[j] + bt[j]) > 0.f ? g[j] : 0.f;
      acc[0][j] += gj * xh; acc[1][j] += gj;
    }
  }
  double* outs[2] = {o.p[0], o.p[1]};
  reduce_finish<2, NP>(acc, cg, rl, CGB, RL, cbase, outs, smem);
}

template <typename T>
__global__ void __launch_bounds__(256, 4) bn_relu_pool_bwd_apply_kernel(const T* __restrict__ dy, const int8_t* __restrict__ idx,
                                                                     const T* __restrict__ x, T* __restrict__ dx, int N, int H, int W,
                                                                     int C, int P, int Q, float inv_rows,
                                                                     const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                     const double* __restrict__ dgamma, const double* __restrict__ dbeta) {
  constexpr int NP = Pack<T>::N;
  const uint32_t CPK = C / NP, PXB = 256 / CPK;
  const int c = (int)(threadIdx.x % CPK) * NP;
  const uint32_t px = threadIdx.x / CPK;
  float mu[NP], is[NP], gm[NP], bt[NP], dbn[NP], dgn[NP];
  load_consts<NP>(mean, c, mu); load_consts<NP>(invstd, c, is); load_consts<NP>(gamma, c, gm); load_consts<NP>(beta, c, bt);
#pragma unroll
  for (int j = 0; j < NP; ++j) { dbn[j] = (float)dbeta[c + j] * inv_rows; dgn[j] = (float)dgamma[c + j] * inv_rows; }
  const uint32_t rows = (uint32_t)N * H * W;
  const uint32_t stride = gridDim.x * PXB, first = blockIdx.x * PXB + px;
  const int sw = (int)(stride % W), sh = (int)((stride / W) % H), sn = (int)(stride / ((uint32_t)W * H));
  int w = (int)(first % W), h = (int)((first / W) % H), n = (int)(first / ((uint32_t)W * H));
  for (uint32_t r = first; r < rows; r += stride, w += sw, h += sh, n += sn) {
    if (w >= W) { w -= W; ++h; }
    if (h >= H) { h -= H; ++n; }
    float g[NP], xv[NP], o[NP];
    Pack<T>::load(x + (int64_t)r * C + c, xv);
    pool_bwd_gather<T>(dy, idx, n, h, w, c, C, P, Q, g);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const float xh = (xv[j] - mu[j]) * is[j];
      const float gj = (xh * gm[j] + bt[j]) > 0.f ? g[j] : 0.f;
      o[j] = gm[j] * is[j] * (gj - dbn[j] - xh * dgn[j]);
    }
    Pack<T>::store(dx + (int64_t)r * C + c, o);
  }
}

// ---- launch geometry ------------------------------------------------------------------------------------------
static inline int pack_n(int dtype) { return dtype == IIT_BF16 ? 8 : 4; }

bool bn_fast_ok(int dtype, int C, const void* a, const void* b, const void* c, const void* d, const void* e) {
  if (dtype != IIT_BF16 && dtype != IIT_F32) return false;
  const int np = pack_n(dtype);
  if (C % np != 0 || (256 * np) % C != 0) return false;        // a thread's channels must be loop-invariant
  const uintptr_t m = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
                      reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(e);
  return (m & 15) == 0;
}

static inline void red_geometry(int64_t rows, int C, int np, int* CGB, dim3* grid, size_t* smem) {
  const int CG = C / np;
  *CGB = CG < 64 ? CG : 64;
  const int RL = 256 / *CGB;
  const int chunks = CG / *CGB;
  int64_t bx = (rows + (int64_t)RL * 8 - 1) / ((int64_t)RL * 8);
  int64_t cap = (int64_t)kNumSMs * 6 / chunks;
  if (cap < 1) cap = 1;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  *grid = dim3((unsigned)bx, (unsigned)chunks);
  *smem = (size_t)2 * RL * *CGB * np * sizeof(float);
}

int bn_stats_fast(const void* x, int dtype, int64_t rows, int C, double* sum, double* sq, cudaStream_t st) {
  int CGB; dim3 grid; size_t smem;
  red_geometry(rows, C, pack_n(dtype), &CGB, &grid, &smem);
  Red2 o{{sum, sq}};
  IIT_DISPATCH_DTYPE(dtype, T, { bn_stats_fast_kernel<T><<<grid, 256, smem, st>>>((const T*)x, rows, C, CGB, o); })
  IIT_LAUNCH_CHECK("iit_bn_stats");
  return 0;
}

int bn_act_fwd_fast(const void* x, const void* res, void* y, int dtype, int64_t rows, int C, const float* mean, const float* invstd,
                    const float* gamma, const float* beta, int relu, cudaStream_t st) {
  const int64_t npacks = rows * C / pack_n(dtype);
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_act_fwd_fast_kernel<T><<<grid_for(npacks, 256 * 2), 256, 0, st>>>((const T*)x, (const T*)res, (T*)y, npacks, C, mean, invstd,
                                                                        gamma, beta, relu);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_fwd");
  return 0;
}

int bn_bwd_reduce_fast(const void* dy, const void* y, const void* x, int dtype, int64_t rows, int C, const float* mean,
                       const float* invstd, int relu, double* dgamma, double* dbeta, cudaStream_t st) {
  int CGB; dim3 grid; size_t smem;
  red_geometry(rows, C, pack_n(dtype), &CGB, &grid, &smem);
  Red2 o{{dgamma, dbeta}};
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_reduce_fast_kernel<T><<<grid, 256, smem, st>>>((const T*)dy, (const T*)y, (const T*)x, rows, C, CGB, mean, invstd, relu, o);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_bwd_reduce");
  return 0;
}

int bn_bwd_apply_fast(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows, int C,
                      const float* mean, const float* invstd, const float* gamma, const double* dgamma, const double* dbeta,
                      int relu, cudaStream_t st) {
  const int64_t npacks = rows * C / pack_n(dtype);
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_apply_fast_kernel<T><<<grid_for(npacks, 256 * 2), 256, 0, st>>>((const T*)dy, (const T*)y, (const T*)x, (T*)dx, (T*)dres,
                                                                          npacks, C, 1.f / (float)rows, mean, invstd, gamma, dgamma,
                                                                          dbeta, relu);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_bwd_apply");
  return 0;
}

}  // namespace iit

using namespace iit;

static int pool_args_ok(const char* name, int dtype, int n, int h, int w, int c) {
  IIT_REQUIRE(dtype == IIT_BF16 || dtype == IIT_F32, "%s: unsupported dtype %d", name, dtype);
  const int np = dtype == IIT_BF16 ? 8 : 4;
  IIT_REQUIRE(n >= 1 && h >= 1 && w >= 1 && c % np == 0 && 256 % (c / np) == 0,
              "%s: channels must be a multiple of %d with 256 %% (c / %d) == 0; got n=%d h=%d w=%d c=%d", name, np, np, n, h, w, c);
  IIT_REQUIRE((int64_t)n * h * w < (1ll << 31), "%s: more than 2^31 pixels", name);
  return 0;
}

extern "C" int iit_bn_relu_maxpool_fwd(const void* x, void* y, int8_t* idx, int dtype, int n, int h, int w, int c, const float* mean,
                                       const float* invstd, const float* gamma, const float* beta, void* stream) {
  IIT_REQUIRE(x && y && idx && mean && invstd && gamma && beta, "iit_bn_relu_maxpool_fwd: null pointer");
  if (pool_args_ok("iit_bn_relu_maxpool_fwd", dtype, n, h, w, c)) return 1;
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0 && (reinterpret_cast<uintptr_t>(idx) & 3) == 0,
              "iit_bn_relu_maxpool_fwd: x and y must be 16-byte aligned, idx 4-byte aligned");
  const int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  const int64_t total = (int64_t)n * P * Q * (c / (dtype == IIT_BF16 ? 8 : 4));
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_relu_pool_fwd_kernel<T><<<grid_for(total, 256, 16), 256, 0, as_stream(stream)>>>((const T*)x, (T*)y, idx, n, h, w, c, P, Q, mean,
                                                                                   invstd, gamma, beta);
  })
  IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_fwd");
  return 0;
}

extern "C" int iit_bn_relu_maxpool_bwd(const void* dy, const int8_t* idx, const void* x, void* dx, int dtype, int n, int h, int w,
                                       int c, const float* mean, const float* invstd, const float* gamma, const float* beta,
                                       double* dgamma, double* dbeta, void* stream) {
  IIT_REQUIRE(dy && idx && x && dx && mean && invstd && gamma && beta && dgamma && dbeta, "iit_bn_relu_maxpool_bwd: null pointer");
  if (pool_args_ok("iit_bn_relu_maxpool_bwd", dtype, n, h, w, c)) return 1;
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0 &&
              (reinterpret_cast<uintptr_t>(idx) & 3) == 0, "iit_bn_relu_maxpool_bwd: x, dy, dx must be 16-byte aligned, idx 4-byte aligned");
  const int np = dtype == IIT_BF16 ? 8 : 4;
  const int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  const int64_t rows = (int64_t)n * h * w;
  cudaStream_t st = as_stream(stream);
  const int CGB = c / np;                      // pool_args_ok: divides 256, so one block spans all channels
  const size_t smem = (size_t)2 * 256 * np * sizeof(float);
  const dim3 grid((unsigned)grid_for(rows * CGB, 256, 6));
  Red2 o{{dgamma, dbeta}};
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_relu_pool_bwd_reduce_kernel<T><<<grid, 256, smem, st>>>((const T*)dy, idx, (const T*)x, n, h, w, c, P, Q, CGB, mean, invstd,
                                                              gamma, beta, o);
  })
  IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_bwd(reduce)");
  const int64_t total = rows * (c / np);
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_relu_pool_bwd_apply_kernel<T><<<grid_for(total, 256, 16), 256, 0, st>>>((const T*)dy, idx, (const T*)x, (T*)dx, n, h, w, c, P, Q,
                                                                          1.f / (float)rows, mean, invstd, gamma, beta, dgamma, dbeta);
  })
  IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_bwd(apply)");
  return 0;
}
