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
                                                              const float* __restrict__ beta, int relu,
                                                              uint8_t* __restrict__ mask) {
  constexpr int NP = Pack<T>::N;
  constexpr int U = 2;                    // packs (and their residuals) in flight per thread: 23.7 -> 23.4 ms per step against one
  const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, stride = (int64_t)gridDim.x * 256;
  const int c0 = (int)((tid * NP) % C);   // invariant: stride * NP is a multiple of C
  float mu[NP], is[NP], g[NP], b[NP];
  load_consts<NP>(mean, c0, mu); load_consts<NP>(invstd, c0, is); load_consts<NP>(gamma, c0, g); load_consts<NP>(beta, c0, b);
  auto finish = [&](int64_t i, const uint4& rx, const uint4& rr) {
    float v[NP], o[NP];
    Pack<T>::unpack(rx, v);
#pragma unroll
    for (int j = 0; j < NP; ++j) o[j] = (v[j] - mu[j]) * is[j] * g[j] + b[j];
    if (res) {
      float r[NP];
      Pack<T>::unpack(rr, r);
#pragma unroll
      for (int j = 0; j < NP; ++j) o[j] += r[j];
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < NP; ++j) o[j] = fmaxf(o[j], 0.f);
      if (mask != nullptr) {   // one byte per pack: bit j = the STORED value of element j is > 0 (what the backward tests)
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) bits |= (round_to<T>(o[j]) > 0.f ? 1u : 0u) << j;
        mask[i] = (uint8_t)bits;
      }
    }
    Pack<T>::store(y + i * NP, o);
  };
  int64_t i = tid;
  for (; i + (U - 1) * stride < npacks; i += U * stride) {
    uint4 rx[U], rr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      rx[u] = *reinterpret_cast<const uint4*>(x + (i + u * stride) * NP);
      rr[u] = make_uint4(0, 0, 0, 0);
      if (res) rr[u] = *reinterpret_cast<const uint4*>(res + (i + u * stride) * NP);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) finish(i + u * stride, rx[u], rr[u]);
  }
  for (; i < npacks; i += stride) {
    const uint4 rx = *reinterpret_cast<const uint4*>(x + i * NP);
    uint4 rr = make_uint4(0, 0, 0, 0);
    if (res) rr = *reinterpret_cast<const uint4*>(res + i * NP);
    finish(i, rx, rr);
  }
}

// ReLU mask of an activation tensor (for forwards that did not take the kernel above)
template <typename T>
__global__ void __launch_bounds__(256) relu_mask_kernel(const T* __restrict__ y, int64_t npacks, uint8_t* __restrict__ mask) {
  constexpr int NP = Pack<T>::N;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < npacks; i += (int64_t)gridDim.x * 256) {
    float v[NP];
    Pack<T>::load(y + i * NP, v);
    uint32_t bits = 0;
#pragma unroll
    for (int j = 0; j < NP; ++j) bits |= (v[j] > 0.f ? 1u : 0u) << j;
    mask[i] = (uint8_t)bits;
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
constexpr int RED_THREADS = 1024;     // bn_bwd_reduce_fast_kernel: one block per SM

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

// dgamma[c] = sum g * xhat,  dbeta[c] = sum g,  g = relu'(y) * dy,  xhat = (x - mean) * invstd.
// The loop accumulates sum(g * (x - mean)) and sum(g) -- the mean stays in the loop (the uncentred form sum(g x) - mean * sum(g)
// cancels badly in fp32), invstd is applied once per block and channel before the atomics. The kernel fits 64 registers: four blocks of 256 threads per SM with RU rows (2 x 16 bytes + the mask byte each) in
// flight per thread, launched as ONE wave (red_geometry). The previous form (xhat in the loop, 121 registers, two blocks
// per SM, 1.7 waves on the 2 x 2 layers) reached 42 % of HBM bandwidth.
// MODE: 0 = no ReLU, 1 = ReLU through the bit mask the forward kept, 2 = ReLU through the stored activation y.
template <typename T, int RU, int MODE>
__global__ void __launch_bounds__(RED_THREADS, 1) bn_bwd_reduce_fast_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                    const T* __restrict__ x, int64_t rows, int C, int CGB,
                                                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                    Red2 o, const uint8_t* __restrict__ mask) {
  extern __shared__ float smem[];
  constexpr int NP = Pack<T>::N;
  const int RL = RED_THREADS / CGB;
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int cbase = blockIdx.y * CGB * NP, c = cbase + cg * NP;
  const int CP = C / NP, cpk = c / NP;           // packs per row, this thread's pack within a row (mask index)
  float acc[2][NP], mu[NP];                      // [0] = sum g * (x - mean), [1] = sum g
  load_consts<NP>(mean, c, mu);
#pragma unroll
  for (int j = 0; j < NP; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  auto add = [&](const uint4& rg, const uint4& rx, const uint4& ry, uint32_t mb) {
    float g[NP], xv[NP];
    Pack<T>::unpack(rg, g);
    Pack<T>::unpack(rx, xv);
    if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = (mb >> j) & 1u ? g[j] : 0.f;
    } else if (MODE == 2) {
      float yv[NP];
      Pack<T>::unpack(ry, yv);
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < NP; ++j) { acc[0][j] = fmaf(g[j], xv[j] - mu[j], acc[0][j]); acc[1][j] += g[j]; }
  };
  const int64_t step = (int64_t)gridDim.x * RL;
  int64_t r = (int64_t)blockIdx.x * RL + rl;
  for (; r + (RU - 1) * step < rows; r += RU * step) {      // RU rows in flight per thread
    uint4 rg[RU], rx[RU], ry[RU];
    uint32_t mb[RU];
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      rg[u] = *reinterpret_cast<const uint4*>(dy + (r + u * step) * C + c);
      rx[u] = *reinterpret_cast<const uint4*>(x + (r + u * step) * C + c);
      ry[u] = make_uint4(0, 0, 0, 0); mb[u] = 0;
      if (MODE == 1) mb[u] = mask[(r + u * step) * CP + cpk];
      else if (MODE == 2) ry[u] = *reinterpret_cast<const uint4*>(y + (r + u * step) * C + c);
    }
#pragma unroll
    for (int u = 0; u < RU; ++u) add(rg[u], rx[u], ry[u], mb[u]);
  }
  for (; r < rows; r += step) {
    const uint4 rg = *reinterpret_cast<const uint4*>(dy + r * C + c), rx = *reinterpret_cast<const uint4*>(x + r * C + c);
    uint4 ry = make_uint4(0, 0, 0, 0);
    uint32_t mb = 0;
    if (MODE == 1) mb = mask[r * CP + cpk];
    else if (MODE == 2) ry = *reinterpret_cast<const uint4*>(y + r * C + c);
    add(rg, rx, ry, mb);
  }
  // block totals through shared memory ([2][RL][W]); `parts` threads per (sum, channel) add RL / parts rows each and meet in
  // a shuffle; the block's share goes to the fp64 totals with one conversion each (fp64 arithmetic runs at 1/64 rate)
  const int W = CGB * NP;
#pragma unroll
  for (int q = 0; q < 2; ++q)
#pragma unroll
    for (int j = 0; j < NP; ++j) smem[(q * RL + rl) * W + cg * NP + j] = acc[q][j];
  __syncthreads();
  const int parts = RED_THREADS / (2 * W) > 32 ? 32 : RED_THREADS / (2 * W);     // 1 ... 8 for W = 512 ... 64; a power of two
  const int outi = threadIdx.x / parts, part = threadIdx.x % parts;     // outi = q * W + ch
  float v = 0.f;
  if (outi < 2 * W) {
    const int q = outi / W, ch = outi - q * W;
    for (int l = part; l < RL; l += parts) v += smem[(q * RL + l) * W + ch];
  }
  for (int o2 = parts >> 1; o2 > 0; o2 >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o2);
  __syncthreads();
  if (part == 0 && outi < 2 * W) smem[outi] = v;           // [0..W) = sum g (x - mean), [W..2W) = sum g
  __syncthreads();
  for (int ch = threadIdx.x; ch < W; ch += blockDim.x) {
    atomicAdd(&o.p[0][cbase + ch], (double)(invstd[cbase + ch] * smem[ch]));
    atomicAdd(&o.p[1][cbase + ch], (double)smem[W + ch]);
  }
}

// dx = gamma * invstd * (g - dbeta/rows - xhat * dgamma/rows),  g = relu'(y) * dy;  dres = g
// Two packs per thread and iteration are loaded before either is used (48 KB of loads in flight per SM at 3 blocks; with one
// pack and 4 blocks the kernel was latency-bound: 24.4 -> 24.1 ms per step).
// `zb`: the box of x (rows p, columns q of the hw x w image, channels) that an interchange patch overwrote in the forward
// pass: its gradient is zero (the patched values are constants; CopySlices backward), written here instead of by a
// separate zeroing kernel on dx.
struct ZeroBox { int on, hw, w, p0, p1, q0, q1, c0, c1; };

template <typename T>
__global__ void __launch_bounds__(256, 3)
bn_bwd_apply_fast_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, T* __restrict__ dx,
                         T* __restrict__ dres, int64_t npacks, int C, float inv_rows, const float* __restrict__ mean,
                         const float* __restrict__ invstd, const float* __restrict__ gamma, const double* __restrict__ dgamma,
                         const double* __restrict__ dbeta, int relu, const uint8_t* __restrict__ mask, ZeroBox zb) {
  constexpr int NP = Pack<T>::N;
  constexpr int U = 2;
  const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, stride = (int64_t)gridDim.x * 256;
  const int c0 = (int)((tid * NP) % C);
  // o = k * (g - dbn - (x - mu) * c1),  k = gamma * invstd,  dbn = dbeta / rows,  c1 = invstd * dgamma / rows. The fp64 sums
  // are converted ONCE per block into shared memory: every thread converting its own 16 values (fp64 conversions issue at
  // 1/64 rate) was a fixed ~9 us per launch, more than the 2 x 2 layers' whole pass.
  extern __shared__ float s_c[];           // [3][C]
  for (int ch = threadIdx.x; ch < C; ch += 256) {
    const float is = invstd[ch];
    s_c[ch] = gamma[ch] * is;
    s_c[C + ch] = (float)dbeta[ch] * inv_rows;
    s_c[2 * C + ch] = is * ((float)dgamma[ch] * inv_rows);
  }
  __syncthreads();
  float mu[NP], k[NP], dbn[NP], c1[NP];
  load_consts<NP>(mean, c0, mu); load_consts<NP>(s_c, c0, k); load_consts<NP>(s_c + C, c0, dbn); load_consts<NP>(s_c + 2 * C, c0, c1);
  auto finish = [&](int64_t i, const uint4& rg, const uint4& rx, const uint4& ry, uint32_t mb) {
    float g[NP], xv[NP], o[NP];
    Pack<T>::unpack(rg, g);
    Pack<T>::unpack(rx, xv);
    if (relu && mask != nullptr) {
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = (mb >> j) & 1u ? g[j] : 0.f;
    } else if (relu) {
      float yv[NP];
      Pack<T>::unpack(ry, yv);
#pragma unroll
      for (int j = 0; j < NP; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
    if (dres) Pack<T>::store(dres + i * NP, g);
#pragma unroll
    for (int j = 0; j < NP; ++j) o[j] = k[j] * (g[j] - dbn[j] - (xv[j] - mu[j]) * c1[j]);
    if (zb.on && c0 >= zb.c0 && c0 < zb.c1) {       // channel ranges are multiples of the pack width (host-checked)
      const int pix = (int)(((i * NP) / C) % zb.hw), pp = pix / zb.w, qq = pix - pp * zb.w;
      if (pp >= zb.p0 && pp < zb.p1 && qq >= zb.q0 && qq < zb.q1) {
#pragma unroll
        for (int j = 0; j < NP; ++j) o[j] = 0.f;
      }
    }
    Pack<T>::store(dx + i * NP, o);
  };
  int64_t i = tid;
  for (; i + (U - 1) * stride < npacks; i += U * stride) {
    uint4 rg[U], rx[U], ry[U];
    uint32_t mb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      rg[u] = *reinterpret_cast<const uint4*>(dy + (i + u * stride) * NP);
      rx[u] = *reinterpret_cast<const uint4*>(x + (i + u * stride) * NP);
      ry[u] = make_uint4(0, 0, 0, 0); mb[u] = 0;
      if (relu && mask != nullptr) mb[u] = mask[i + u * stride];
      else if (relu) ry[u] = *reinterpret_cast<const uint4*>(y + (i + u * stride) * NP);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) finish(i + u * stride, rg[u], rx[u], ry[u], mb[u]);
  }
  for (; i < npacks; i += stride) {
    const uint4 rg = *reinterpret_cast<const uint4*>(dy + i * NP), rx = *reinterpret_cast<const uint4*>(x + i * NP);
    uint4 ry = make_uint4(0, 0, 0, 0);
    uint32_t mb = 0;
    if (relu && mask != nullptr) mb = mask[i];
    else if (relu) ry = *reinterpret_cast<const uint4*>(y + i * NP);
    finish(i, rg, rx, ry, mb);
  }
}

// ---- fused stem tail: BN -> ReLU -> MaxPool2d(3, 2, 1) -----------------------------------------------------------
// forward: one thread = one pooled pixel x one channel pack
template <typename T>
__global__ void __launch_bounds__(256) bn_relu_pool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx,
                                                               T* __restrict__ xwin, int N, int H, int W, int C, int P, int Q,
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
    // the winners' raw values (exact copies of elements of x): the backward's per-channel sums run over these pooled
    // tensors instead of gathering through the full-resolution ones
    if (xwin != nullptr) Pack<T>::store(xwin + (int64_t)pix * C + cp * NP, best);
#pragma unroll
    for (int j = 0; j < NP; ++j) best[j] = fmaxf((best[j] - mu[j]) * is[j] * g[j] + b[j], 0.f);
    Pack<T>::store(y + (int64_t)pix * C + cp * NP, best);
    uint32_t* ip = reinterpret_cast<uint32_t*>(idx + (int64_t)pix * C + cp * NP);
#pragma unroll
    for (int j = 0; j < NP / 4; ++j) ip[j] = ipk[j];
  }
}

// Backward. The per-channel sums (dgamma, dbeta) need, for every pooled pixel, the gradient, the ReLU mask and the
// normalised value of the window's winner: the forward saved the winners' raw values (xwin), so the sums are the generic
// BN reduction over the POOLED tensors (bn_bwd_reduce_fast_kernel on dy, y, xwin: a quarter of the pixels, no gather).
//
// dx: one thread = a 2 x 2 block of input pixels x one channel pack. The block (rows 2a, 2a+1; columns 2b, 2b+1) is
// covered by the four windows (a, b), (a, b+1), (a+1, b), (a+1, b+1) only, so four gradient packs and four slot words
// serve four outputs (the per-pixel gather needed four of each per output). Slot of pixel (2a+i, 2b+j) inside window
// (a+u, b+v): row 1 + i - 2u, column 1 + j - 2v when both lie in 0..2.
template <typename T>
__global__ void __launch_bounds__(256, 2) bn_relu_pool_bwd_apply_kernel(const T* __restrict__ dy, const int8_t* __restrict__ idx,
                                                                     const T* __restrict__ x, T* __restrict__ dx, int N, int H, int W,
                                                                     int C, int P, int Q, float inv_rows,
                                                                     const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                     const double* __restrict__ dgamma, const double* __restrict__ dbeta) {
  constexpr int NP = Pack<T>::N;
  const uint32_t CPK = C / NP, PXB = 256 / CPK;
  const int c = (int)(threadIdx.x % CPK) * NP;
  const uint32_t px = threadIdx.x / CPK;
  // four constants per channel (two blocks of 2 x 2 pixels in flight leave no registers for six):
  //   activation = x * A + B0 (its sign is the ReLU mask);  dx = A * g + K0 + K1 * x
  float A[NP], B0[NP], K0[NP], K1[NP];
  extern __shared__ float s_d[];            // [2][C]: dbeta / rows, dgamma / rows, converted from fp64 once per block
  for (int ch = threadIdx.x; ch < C; ch += 256) { s_d[ch] = (float)dbeta[ch] * inv_rows; s_d[C + ch] = (float)dgamma[ch] * inv_rows; }
  __syncthreads();
  {
    float mu[NP], is[NP], gm[NP], bt[NP];
    load_consts<NP>(mean, c, mu); load_consts<NP>(invstd, c, is); load_consts<NP>(gamma, c, gm); load_consts<NP>(beta, c, bt);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const float dbn = s_d[c + j], dgn = s_d[C + c + j];
      A[j] = gm[j] * is[j];
      B0[j] = bt[j] - mu[j] * A[j];
      K1[j] = -A[j] * dgn * is[j];
      K0[j] = -A[j] * dbn - K1[j] * mu[j];
    }
  }
  const int HB = (H + 1) >> 1, WB = (W + 1) >> 1;
  const uint32_t blocks = (uint32_t)N * HB * WB;
  const uint32_t stride = gridDim.x * PXB, first = blockIdx.x * PXB + px;
  const int sb = (int)(stride % WB), sa = (int)((stride / WB) % HB), sn = (int)(stride / ((uint32_t)WB * HB));
  int bb = (int)(first % WB), a = (int)((first / WB) % HB), n = (int)(first / ((uint32_t)WB * HB));
  for (uint32_t r = first; r < blocks; r += stride, bb += sb, a += sa, n += sn) {
    if (bb >= WB) { bb -= WB; ++a; }
    if (a >= HB) { a -= HB; ++n; }
    // the four windows: clamped loads, invalid ones get slot words that match nothing
    uint4 g4[4];
    uint32_t k4[4][NP / 4];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        const int t = u * 2 + v;
        const bool ok = a + u < P && bb + v < Q;
        const int64_t op = (((int64_t)n * P + min(a + u, P - 1)) * Q + min(bb + v, Q - 1)) * C + c;
        g4[t] = *reinterpret_cast<const uint4*>(dy + op);
        if constexpr (NP == 8) {
          const uint2 kk = *reinterpret_cast<const uint2*>(idx + op);
          k4[t][0] = ok ? kk.x : 0xffffffffu; k4[t][1] = ok ? kk.y : 0xffffffffu;
        } else {
          const uint32_t kk = *reinterpret_cast<const uint32_t*>(idx + op);
          k4[t][0] = ok ? kk : 0xffffffffu;
        }
      }
    uint4 xr[4];
    bool pok[4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int h = 2 * a + i, w = 2 * bb + j;
        pok[i * 2 + j] = h < H && w < W;
        xr[i * 2 + j] = *reinterpret_cast<const uint4*>(x + (((int64_t)n * H + min(h, H - 1)) * W + min(w, W - 1)) * C + c);
      }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (!pok[i * 2 + j]) continue;
        float g[NP], xv[NP], o[NP];
#pragma unroll
        for (int e = 0; e < NP; ++e) g[e] = 0.f;
#pragma unroll
        for (int u = 0; u <= i; ++u)
#pragma unroll
          for (int v = 0; v <= j; ++v) {
            const int t = u * 2 + v;
            const uint32_t slot = (uint32_t)((1 + i - 2 * u) * 3 + (1 + j - 2 * v));
            float d[NP];
            Pack<T>::unpack(g4[t], d);
            if constexpr (NP == 8) {
              const uint32_t s4 = slot * 0x01010101u;
              const uint32_t m0 = __vcmpeq4(k4[t][0], s4), m1 = __vcmpeq4(k4[t][1], s4);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if ((m0 >> (8 * e)) & 1u) g[e] += d[e];
                if ((m1 >> (8 * e)) & 1u) g[4 + e] += d[4 + e];
              }
            } else {
#pragma unroll
              for (int e = 0; e < NP; ++e)
                if (((k4[t][e >> 2] >> (8 * (e & 3))) & 0xffu) == slot) g[e] += d[e];
            }
          }
        Pack<T>::unpack(xr[i * 2 + j], xv);
#pragma unroll
        for (int e = 0; e < NP; ++e) {
          const float ge = fmaf(xv[e], A[e], B0[e]) > 0.f ? g[e] : 0.f;
          o[e] = fmaf(A[e], ge, fmaf(K1[e], xv[e], K0[e]));
        }
        Pack<T>::store(dx + (((int64_t)n * H + 2 * a + i) * W + 2 * bb + j) * C + c, o);
      }
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

// one wave of 1024-thread blocks (bn_bwd_reduce_fast_kernel): block = CGB channel groups x RL row lanes
static inline void red_geometry_1024(int64_t rows, int C, int np, int* CGB, dim3* grid, size_t* smem) {
  const int CG = C / np;
  *CGB = CG < 64 ? CG : 64;
  const int RL = RED_THREADS / *CGB;
  const int chunks = CG / *CGB;
  int64_t bx = (rows + (int64_t)RL * 4 - 1) / ((int64_t)RL * 4);
  int64_t cap = kNumSMs / chunks;
  if (cap < 1) cap = 1;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  *grid = dim3((unsigned)bx, (unsigned)chunks);
  *smem = (size_t)2 * RL * *CGB * np * sizeof(float);      // 64 KB (bf16) / 32 KB (fp32)
}

template <typename K>
static int red_opt_in(K kernel, size_t smem) {
  IIT_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return 0;
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
                    const float* gamma, const float* beta, int relu, uint8_t* mask, cudaStream_t st) {
  const int64_t npacks = rows * C / pack_n(dtype);
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_act_fwd_fast_kernel<T><<<grid_for(npacks, 256 * 2), 256, 0, st>>>((const T*)x, (const T*)res, (T*)y, npacks, C, mean, invstd,
                                                                        gamma, beta, relu, mask);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_fwd");
  return 0;
}

int relu_mask_fast(const void* y, int dtype, int64_t rows, int C, uint8_t* mask, cudaStream_t st) {
  const int64_t npacks = rows * C / pack_n(dtype);
  IIT_DISPATCH_DTYPE(dtype, T, { relu_mask_kernel<T><<<grid_for(npacks, 256 * 2), 256, 0, st>>>((const T*)y, npacks, mask); })
  IIT_LAUNCH_CHECK("iit_bn_act_fwd(mask)");
  return 0;
}

int bn_bwd_reduce_fast(const void* dy, const void* y, const void* x, int dtype, int64_t rows, int C, const float* mean,
                       const float* invstd, int relu, double* dgamma, double* dbeta, const uint8_t* mask, cudaStream_t st) {
  int CGB; dim3 grid; size_t smem;
  red_geometry_1024(rows, C, pack_n(dtype), &CGB, &grid, &smem);
  Red2 o{{dgamma, dbeta}};
  IIT_DISPATCH_DTYPE(dtype, T, {   // four rows in flight per thread (three when the stored activation is read as well)
    auto k0 = bn_bwd_reduce_fast_kernel<T, 4, 0>; auto k1 = bn_bwd_reduce_fast_kernel<T, 4, 1>; auto k2 = bn_bwd_reduce_fast_kernel<T, 3, 2>;
    auto k = !relu ? k0 : (mask != nullptr ? k1 : k2);
    if (red_opt_in(k, smem)) return 3;
    k<<<grid, RED_THREADS, smem, st>>>((const T*)dy, (const T*)y, (const T*)x, rows, C, CGB, mean, invstd, o, mask);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_bwd_reduce");
  return 0;
}

int bn_bwd_apply_fast(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows, int C,
                      const float* mean, const float* invstd, const float* gamma, const double* dgamma, const double* dbeta,
                      int relu, const uint8_t* mask, cudaStream_t st, const int* zero_box_hw) {
  const int64_t npacks = rows * C / pack_n(dtype);
  ZeroBox zb{};
  if (zero_box_hw != nullptr) {     // h, w, p0, p1, q0, q1, c0, c1
    zb.on = 1; zb.hw = zero_box_hw[0] * zero_box_hw[1]; zb.w = zero_box_hw[1];
    zb.p0 = zero_box_hw[2]; zb.p1 = zero_box_hw[3]; zb.q0 = zero_box_hw[4]; zb.q1 = zero_box_hw[5];
    zb.c0 = zero_box_hw[6]; zb.c1 = zero_box_hw[7];
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_apply_fast_kernel<T><<<grid_for(npacks, 256 * 2, 6), 256, (size_t)3 * C * sizeof(float), st>>>((const T*)dy, (const T*)y, (const T*)x, (T*)dx, (T*)dres,
                                                                          npacks, C, 1.f / (float)rows, mean, invstd, gamma, dgamma,
                                                                          dbeta, relu, mask, zb);
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

extern "C" int iit_bn_relu_maxpool_fwd(const void* x, void* y, int8_t* idx, void* xwin, int dtype, int n, int h, int w, int c,
                                       const float* mean, const float* invstd, const float* gamma, const float* beta, void* stream) {
  IIT_REQUIRE(x && y && idx && mean && invstd && gamma && beta, "iit_bn_relu_maxpool_fwd: null pointer");
  if (pool_args_ok("iit_bn_relu_maxpool_fwd", dtype, n, h, w, c)) return 1;
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(xwin)) & 15) == 0 &&
              (reinterpret_cast<uintptr_t>(idx) & 7) == 0,
              "iit_bn_relu_maxpool_fwd: x, y and xwin must be 16-byte aligned, idx 8-byte aligned");
  const int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  const int64_t total = (int64_t)n * P * Q * (c / (dtype == IIT_BF16 ? 8 : 4));
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_relu_pool_fwd_kernel<T><<<grid_for(total, 256, 16), 256, 0, as_stream(stream)>>>((const T*)x, (T*)y, idx, (T*)xwin, n, h, w, c, P, Q,
                                                                                   mean, invstd, gamma, beta);
  })
  IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_fwd");
  return 0;
}

extern "C" int iit_bn_relu_maxpool_bwd(const void* dy, const int8_t* idx, const void* x, const void* y, const void* xwin, void* dx,
                                       int dtype, int n, int h, int w, int c, const float* mean, const float* invstd,
                                       const float* gamma, const float* beta, double* dgamma, double* dbeta, void* stream) {
  IIT_REQUIRE(dy && idx && x && y && xwin && dx && mean && invstd && gamma && beta && dgamma && dbeta,
              "iit_bn_relu_maxpool_bwd: null pointer");
  if (pool_args_ok("iit_bn_relu_maxpool_bwd", dtype, n, h, w, c)) return 1;
  IIT_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx) |
                reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(xwin)) & 15) == 0 &&
              (reinterpret_cast<uintptr_t>(idx) & 7) == 0,
              "iit_bn_relu_maxpool_bwd: x, y, xwin, dy, dx must be 16-byte aligned, idx 8-byte aligned");
  const int np = dtype == IIT_BF16 ? 8 : 4;
  const int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  const int64_t rows = (int64_t)n * h * w, prows = (int64_t)n * P * Q;
  cudaStream_t st = as_stream(stream);
  // per-channel sums over the pooled tensors: mask = y > 0, normalised value from the winner's raw value
  {
    int CGB; dim3 grid; size_t smem;
    red_geometry_1024(prows, c, np, &CGB, &grid, &smem);
    Red2 o{{dgamma, dbeta}};
    IIT_DISPATCH_DTYPE(dtype, T, {
      if (red_opt_in(bn_bwd_reduce_fast_kernel<T, 3, 2>, smem)) return 3;
      bn_bwd_reduce_fast_kernel<T, 3, 2><<<grid, RED_THREADS, smem, st>>>((const T*)dy, (const T*)y, (const T*)xwin, prows, c, CGB, mean,
                                                                       invstd, o, nullptr);
    })
    IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_bwd(reduce)");
  }
  const int64_t total = (int64_t)n * ((h + 1) / 2) * ((w + 1) / 2) * (c / np);
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_relu_pool_bwd_apply_kernel<T><<<grid_for(total, 256, 12), 256, (size_t)2 * c * sizeof(float), st>>>((const T*)dy, idx, (const T*)x, (T*)dx, n, h, w, c, P, Q,
                                                                          1.f / (float)rows, mean, invstd, gamma, beta, dgamma, dbeta);
  })
  IIT_LAUNCH_CHECK("iit_bn_relu_maxpool_bwd(apply)");
  return 0;
}
