// (b) the HBM-bound layers around the convolutions: BatchNorm2d (training statistics, fused
// scale/shift + residual + ReLU, two-pass backward), ReLU, MaxPool 3x3/2, global average pool,
// the 512->10 linear head; plus the flat-buffer Adam step and gradient norm.
//
// Activations are NHWC: a [rows, C] matrix with C innermost. Threads own 4 consecutive channels
// (16 B fp32 / 8 B bf16 vector accesses) and stride over rows, so every warp access is a run of
// full cache lines. Per-channel reductions finish in fp64 global atomics: fp32 partial sums of at
// most a few hundred terms per thread, fp64 across blocks (E[x^2]-mean^2 needs the headroom).
#include "common.cuh"

namespace iit {

// ------------------------------------------------------------------------------------------------
// per-channel reductions over rows. NRED values reduced per element, via functor f(row-values) .
// block = 256 threads = CG channel groups (C/4, capped at 64) x RL row lanes
// ------------------------------------------------------------------------------------------------
template <int NRED>
__device__ __forceinline__ void channel_reduce_finish(float (&acc)[NRED][4], int cg, int rl, int CGB, int RL, int c0, int C,
                                                      double* const* outs, float* smem) {
  // smem layout: [NRED][RL][CGB*4]
  const int W = CGB * 4;
#pragma unroll
  for (int r = 0; r < NRED; ++r)
#pragma unroll
    for (int j = 0; j < 4; ++j) smem[(r * RL + rl) * W + cg * 4 + j] = acc[r][j];
  __syncthreads();
  for (int i = threadIdx.x; i < NRED * W; i += blockDim.x) {
    int r = i / W, ch = i - r * W;
    if (c0 + ch < C) {
      float s = 0.f;
      for (int l = 0; l < RL; ++l) s += smem[(r * RL + l) * W + ch];
      atomicAdd(&outs[r][c0 + ch], (double)s);
    }
  }
}

struct RedPtrs2 { double* p[2]; };

template <typename T>
__global__ void __launch_bounds__(256) bn_stats_kernel(const T* __restrict__ x, int64_t rows, int C, RedPtrs2 o) {
  extern __shared__ float smem[];
  const int CG = C / 4;
  const int CGB = CG < 64 ? CG : 64;      // channel groups handled by one block
  const int RL = 256 / CGB;
  const int nchunk = (CG + CGB - 1) / CGB;
  const int chunk = blockIdx.y;            // which slab of channels
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int c = (chunk * CGB + cg) * 4;
  (void)nchunk;
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  if (c < C && rl < RL) {
    for (int64_t r = (int64_t)blockIdx.x * RL + rl; r < rows; r += (int64_t)gridDim.x * RL) {
      float v[4];
      Vec4<T>::load(x + r * C + c, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) { acc[0][j] += v[j]; acc[1][j] += v[j] * v[j]; }
    }
  }
  double* outs[2] = {o.p[0], o.p[1]};
  if (rl < RL) channel_reduce_finish<2>(acc, cg, rl, CGB, RL, chunk * CGB * 4, C, outs, smem);
  else __syncthreads();
}

__global__ void bn_finalize_kernel(double* __restrict__ sum, double* __restrict__ sq, int64_t rows, int C,
                                   float eps, float momentum, float* __restrict__ mean, float* __restrict__ invstd,
                                   float* __restrict__ rmean, float* __restrict__ rvar, long long* __restrict__ tracked,
                                   int clear) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && tracked) *tracked += 1;            // num_batches_tracked, folded into this launch
  if (c >= C) return;
  double m = sum[c] / (double)rows;
  double var = sq[c] / (double)rows - m * m;
  if (clear) { sum[c] = 0.0; sq[c] = 0.0; }        // consume-and-clear: the accumulators are zero for their next use
  if (var < 0) var = 0;
  mean[c] = (float)m;
  invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
  if (rmean) {
    double unb = rows > 1 ? var * (double)rows / (double)(rows - 1) : var;
    rmean[c] = (float)((1.0 - momentum) * rmean[c] + momentum * m);
    rvar[c] = (float)((1.0 - momentum) * rvar[c] + momentum * unb);
  }
}

__global__ void bn_eval_affine_kernel(const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ rmean,
                                      const float* __restrict__ rvar, int C, float eps, float* __restrict__ scale,
                                      float* __restrict__ shift) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float is = (float)(1.0 / sqrt((double)rvar[c] + (double)eps));      // as bn_eval_invstd_kernel
  const float sc = gamma[c] * is;
  scale[c] = sc;
  shift[c] = beta[c] - rmean[c] * sc;
}

__global__ void bn_eval_invstd_kernel(const float* __restrict__ rvar, int C, float eps, float* __restrict__ invstd) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) invstd[c] = 1.0f / sqrtf(rvar[c] + eps);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_act_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y,
                                                         int64_t n4, int C, const float* __restrict__ mean,
                                                         const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, int relu) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t e = i * 4;
    int c = (int)(e % C);
    float v[4], o[4];
    Vec4<T>::load(x + e, v);
    float4 mu = *reinterpret_cast<const float4*>(mean + c), is = *reinterpret_cast<const float4*>(invstd + c);
    float4 g = *reinterpret_cast<const float4*>(gamma + c), b = *reinterpret_cast<const float4*>(beta + c);
    // same operation order as (x - mean) * invstd * gamma + beta
    o[0] = (v[0] - mu.x) * is.x * g.x + b.x;
    o[1] = (v[1] - mu.y) * is.y * g.y + b.y;
    o[2] = (v[2] - mu.z) * is.z * g.z + b.z;
    o[3] = (v[3] - mu.w) * is.w * g.w + b.w;
    if (res) {
      float rv[4];
      Vec4<T>::load(res + e, rv);
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] += rv[j];
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = fmaxf(o[j], 0.f);
    }
    Vec4<T>::store(y + e, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                            const T* __restrict__ x, int64_t rows, int C,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd,
                                                            int relu, RedPtrs2 o) {
  extern __shared__ float smem[];
  const int CG = C / 4;
  const int CGB = CG < 64 ? CG : 64;
  const int RL = 256 / CGB;
  const int chunk = blockIdx.y;
  const int cg = threadIdx.x % CGB, rl = threadIdx.x / CGB;
  const int c = (chunk * CGB + cg) * 4;
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  if (c < C && rl < RL) {
    float4 mu = *reinterpret_cast<const float4*>(mean + c), is = *reinterpret_cast<const float4*>(invstd + c);
    float m4[4] = {mu.x, mu.y, mu.z, mu.w}, i4[4] = {is.x, is.y, is.z, is.w};
    for (int64_t r = (int64_t)blockIdx.x * RL + rl; r < rows; r += (int64_t)gridDim.x * RL) {
      float g[4], xv[4];
      Vec4<T>::load(dy + r * C + c, g);
      Vec4<T>::load(x + r * C + c, xv);
      if (relu) {
        float yv[4];
        Vec4<T>::load(y + r * C + c, yv);
#pragma unroll
        for (int j = 0; j < 4; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) { acc[0][j] += g[j] * (xv[j] - m4[j]) * i4[j]; acc[1][j] += g[j]; }
    }
  }
  double* outs[2] = {o.p[0], o.p[1]};
  if (rl < RL) channel_reduce_finish<2>(acc, cg, rl, CGB, RL, chunk * CGB * 4, C, outs, smem);
  else __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                           const T* __restrict__ x, T* __restrict__ dx, T* __restrict__ dres,
                                                           int64_t n4, int C, float inv_rows, const float* __restrict__ mean,
                                                           const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                           const double* __restrict__ dgamma, const double* __restrict__ dbeta,
                                                           int relu) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t e = i * 4;
    int c = (int)(e % C);
    float g[4], xv[4], o[4];
    Vec4<T>::load(dy + e, g);
    Vec4<T>::load(x + e, xv);
    if (relu) {
      float yv[4];
      Vec4<T>::load(y + e, yv);
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    }
    if (dres) Vec4<T>::store(dres + e, g);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float is = invstd[c + j];
      float xh = (xv[j] - mean[c + j]) * is;
      float dg = (float)dgamma[c + j], db = (float)dbeta[c + j];
      o[j] = gamma[c + j] * is * (g[j] - db * inv_rows - xh * dg * inv_rows);
    }
    Vec4<T>::store(dx + e, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) relu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float v[4];
    Vec4<T>::load(x + i * 4, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = fmaxf(v[j], 0.f);
    Vec4<T>::store(y + i * 4, v);
  }
}
// y = x / scale (IEEE division in fp32, rounded to T: what torch's `tensor / scalar` gives)
template <typename T>
__global__ void __launch_bounds__(256) div_scalar_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t n4, float scale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float v[4];
    Vec4<T>::load(x + i * 4, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = v[j] / scale;
    Vec4<T>::store(y + i * 4, v);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) relu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, T* __restrict__ dx,
                                                       int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float g[4], yv[4];
    Vec4<T>::load(dy + i * 4, g);
    Vec4<T>::load(y + i * 4, yv);
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
    Vec4<T>::store(dx + i * 4, g);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) add_relu_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y,
                                                       int64_t n4, int relu) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float u[4], v[4];
    Vec4<T>::load(a + i * 4, u);
    Vec4<T>::load(b + i * 4, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) { u[j] += v[j]; if (relu) u[j] = fmaxf(u[j], 0.f); }
    Vec4<T>::store(y + i * 4, u);
  }
}

// ------------------------------------------------------------------------------------------------
// MaxPool2d(3, 2, 1), NHWC. One thread = 4 channels of one output pixel.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int8_t* __restrict__ idx,
                                                          int N, int H, int W, int C, int P, int Q) {
  const int C4 = C / 4;
  const int64_t total = (int64_t)N * P * Q * C4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c4 = (int)(i % C4);
    int64_t pix = i / C4;
    int q = (int)(pix % Q);
    int64_t t2 = pix / Q;
    int p = (int)(t2 % P);
    int n = (int)(t2 / P);
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    int bi[4] = {0, 0, 0, 0};
    bool any = false;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      int ih = p * 2 - 1 + r;
      if (ih < 0 || ih >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        int iw = q * 2 - 1 + s;
        if (iw < 0 || iw >= W) continue;
        float v[4];
        Vec4<T>::load(x + (((int64_t)n * H + ih) * W + iw) * C + c4 * 4, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // first maximum in window scan order, NaN propagates (torch max_pool2d)
          if (!any || v[j] > best[j] || v[j] != v[j]) { best[j] = v[j]; bi[j] = r * 3 + s; }
        }
        any = true;
      }
    }
    Vec4<T>::store(y + pix * C + c4 * 4, best);
    char4 o = make_char4((char)bi[0], (char)bi[1], (char)bi[2], (char)bi[3]);
    *reinterpret_cast<char4*>(idx + pix * C + c4 * 4) = o;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dy, const int8_t* __restrict__ idx,
                                                          T* __restrict__ dx, int N, int H, int W, int C, int P, int Q) {
  const int C4 = C / 4;
  const int64_t total = (int64_t)N * H * W * C4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c4 = (int)(i % C4);
    int64_t pix = i / C4;
    int w = (int)(pix % W);
    int64_t t2 = pix / W;
    int h = (int)(t2 % H);
    int n = (int)(t2 / H);
    float g[4] = {0.f, 0.f, 0.f, 0.f};
    // output windows containing (h, w): p*2-1 <= h <= p*2+1
    int p_lo = h >> 1, p_hi = (h + 1) >> 1;   // (h-1)/2 rounded up .. (h+1)/2
    int q_lo = w >> 1, q_hi = (w + 1) >> 1;
    for (int p = p_lo; p <= p_hi; ++p) {
      if (p >= P) continue;
      int r = h - (p * 2 - 1);
      if (r < 0 || r > 2) continue;
      for (int q = q_lo; q <= q_hi; ++q) {
        if (q >= Q) continue;
        int s = w - (q * 2 - 1);
        if (s < 0 || s > 2) continue;
        int64_t op = (((int64_t)n * P + p) * Q + q) * C + c4 * 4;
        char4 k = *reinterpret_cast<const char4*>(idx + op);
        float d[4];
        Vec4<T>::load(dy + op, d);
        int slot = r * 3 + s;
        if (k.x == slot) g[0] += d[0];
        if (k.y == slot) g[1] += d[1];
        if (k.z == slot) g[2] += d[2];
        if (k.w == slot) g[3] += d[3];
      }
    }
    Vec4<T>::store(dx + pix * C + c4 * 4, g);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const T* __restrict__ x, float* __restrict__ y, int N, int HW, int C) {
  const int64_t total = (int64_t)N * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t n = i / C;
    float s = 0.f;
    for (int j = 0; j < HW; ++j) s += to_f<T>(x[(n * HW + j) * C + c]);
    y[i] = s / (float)HW;
  }
}
template <typename T>
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  const int64_t total = (int64_t)N * HW * C;
  const float inv = 1.f / (float)HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t n = i / ((int64_t)HW * C);
    dx[i] = from_f<T>(dy[n * C + c] * inv);
  }
}

// ------------------------------------------------------------------------------------------------
// Linear head (small: 512 -> 10). One warp per output element for fwd / dx; dw by column threads.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ b, float* __restrict__ y, int N, int I, int O) {
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  int64_t total = (int64_t)N * O;
  int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t e = warp; e < total; e += nw) {
    int64_t n = e / O;
    int o = (int)(e - n * O);
    float s = 0.f;
    for (int i = lane; i < I; i += 32) s = fmaf(x[n * I + i], w[(int64_t)o * I + i], s);
    s = warp_sum(s);
    if (lane == 0) y[e] = s + (b ? b[o] : 0.f);
  }
}
__global__ void __launch_bounds__(256) linear_dx_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                        float* __restrict__ dx, int N, int I, int O) {
  int64_t total = (int64_t)N * I;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = e / I;
    int i = (int)(e - n * I);
    float s = 0.f;
    for (int o = 0; o < O; ++o) s = fmaf(dy[n * O + o], w[(int64_t)o * I + i], s);
    dx[e] = s;
  }
}
// dw[o,i] = sum_n dy[n,o] x[n,i], db[o] = sum_n dy[n,o]. The head is tiny (O = 10) but N is the batch: rows are split
// over gridDim.y blocks; a block covers 64 columns x 4 row lanes, keeps all O (<= 16) outputs of its column in registers
// so that x is read once, and adds its partial sums to the zeroed dw / db with fp32 atomics.
constexpr int LIN_MAX_O = 16;
__global__ void __launch_bounds__(256) linear_dw_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                        float* __restrict__ dw, float* __restrict__ db, int N, int I, int O,
                                                        int rows_per_block) {
  __shared__ float s_part[4][LIN_MAX_O][64];
  const int col = threadIdx.x & 63, rl = threadIdx.x >> 6;
  const int i = blockIdx.x * 64 + col;
  const int n0 = blockIdx.y * rows_per_block;
  const int n1 = min(N, n0 + rows_per_block);
  float acc[LIN_MAX_O];
#pragma unroll
  for (int o = 0; o < LIN_MAX_O; ++o) acc[o] = 0.f;
  if (i < I) {
    for (int n = n0 + rl; n < n1; n += 4) {
      const float xv = x[(int64_t)n * I + i];
#pragma unroll
      for (int o = 0; o < LIN_MAX_O; ++o)
        if (o < O) acc[o] = fmaf(dy[(int64_t)n * O + o], xv, acc[o]);
    }
  }
#pragma unroll
  for (int o = 0; o < LIN_MAX_O; ++o) s_part[rl][o][col] = acc[o];
  __syncthreads();
  for (int e = threadIdx.x; e < O * 64; e += 256) {
    const int o = e >> 6, c = e & 63;
    const int ii = blockIdx.x * 64 + c;
    if (ii < I) atomicAdd(&dw[(int64_t)o * I + ii], s_part[0][o][c] + s_part[1][o][c] + s_part[2][o][c] + s_part[3][o][c]);
  }
  if (db && blockIdx.x == 0 && threadIdx.x < O) {
    float sb = 0.f;
    for (int n = n0; n < n1; ++n) sb += dy[(int64_t)n * O + threadIdx.x];
    atomicAdd(&db[threadIdx.x], sb);
  }
}

// ------------------------------------------------------------------------------------------------
// optimiser pieces on a flat fp32 buffer
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sqnorm_kernel(const float* __restrict__ g, int64_t n, double* __restrict__ out) {
  __shared__ float s[8];
  float acc = 0.f;
  int64_t n4 = n >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(g)[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) { float v = g[n4 * 4 + threadIdx.x]; acc += v * v; }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int i = 0; i < 8; ++i) t += (double)s[i];
    atomicAdd(out, t);
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t n, float step_size, float b1, float b2, float eps,
                                                   float bc2_sqrt, const double* __restrict__ sqnorm, float max_norm) {
  float clip = 1.f;
  if (sqnorm && max_norm > 0.f) {
    float total = (float)sqrt(*sqnorm);
    float coef = max_norm / (total + 1e-6f);
    clip = coef < 1.f ? coef : 1.f;     // torch.nn.utils.clip_grad_norm_: clamp(max_norm/(norm+1e-6), max=1)
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * clip;
    float mi = m[i] + (gi - m[i]) * (1.f - b1);          // exp_avg.lerp_(grad, 1-beta1)
    float vi = v[i] * b2 + gi * gi * (1.f - b2);          // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2)
    m[i] = mi; v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Device-resident step counter and learning rate: nothing about the update is baked into the launch
// arguments, so the launch can sit inside a captured CUDA graph and still follow the schedule.
__global__ void adam_tick_kernel(long long* step) { *step += 1; }

__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                       float* __restrict__ v, int64_t n, const float* __restrict__ lr_dev, double b1d,
                                                       double b2d, float eps, const long long* __restrict__ step_dev,
                                                       const double* __restrict__ sqnorm, float max_norm) {
  // bias corrections in double, as torch.optim.Adam computes them in Python floats
  const double step = (double)*step_dev;
  const double bc1 = 1.0 - pow(b1d, step), bc2 = 1.0 - pow(b2d, step);
  const float step_size = (float)((double)*lr_dev / bc1), bc2_sqrt = (float)sqrt(bc2);
  const float b1 = (float)b1d, b2 = (float)b2d;
  float clip = 1.f;
  if (sqnorm && max_norm > 0.f) {
    float total = (float)sqrt(*sqnorm);
    float coef = max_norm / (total + 1e-6f);
    clip = coef < 1.f ? coef : 1.f;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * clip;
    float mi = m[i] + (gi - m[i]) * (1.f - b1);
    float vi = v[i] * b2 + gi * gi * (1.f - b2);
    m[i] = mi; v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// bn_fast.cu: 16-byte-per-thread kernels with register-resident per-channel constants
bool bn_fast_ok(int dtype, int C, const void* a, const void* b, const void* c, const void* d, const void* e);
int bn_stats_fast(const void* x, int dtype, int64_t rows, int C, double* sum, double* sq, cudaStream_t st);
int bn_act_fwd_fast(const void* x, const void* res, void* y, int dtype, int64_t rows, int C, const float* mean, const float* invstd,
                    const float* gamma, const float* beta, int relu, uint8_t* mask, cudaStream_t st);
int relu_mask_fast(const void* y, int dtype, int64_t rows, int C, uint8_t* mask, cudaStream_t st);
int bn_bwd_reduce_fast(const void* dy, const void* y, const void* x, int dtype, int64_t rows, int C, const float* mean,
                       const float* invstd, int relu, double* dgamma, double* dbeta, const uint8_t* mask, cudaStream_t st);
int bn_bwd_apply_fast(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows, int C,
                      const float* mean, const float* invstd, const float* gamma, const double* dgamma, const double* dbeta,
                      int relu, const uint8_t* mask, cudaStream_t st, const int* zero_box_hw = nullptr);

// fp64 per-channel sums of a BatchNorm backward -> fp32 parameter gradients, written or accumulated in place (the
// optimizer's flat gradient buffer), and the layer's persistent accumulator cleared for its next backward: one launch
// instead of a fill, two casts and two accumulation kernels per layer
__global__ void bn_param_grads_kernel(double* __restrict__ sums, int c, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                      int accumulate, int clear) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c) return;
  const float dg = (float)sums[i], db = (float)sums[c + i];
  dgamma[i] = accumulate ? dgamma[i] + dg : dg;
  dbeta[i] = accumulate ? dbeta[i] + db : db;
  if (clear) { sums[i] = 0.0; sums[c + i] = 0.0; }
}

}  // namespace iit

using namespace iit;

static inline size_t red_smem(int C) {
  int CG = C / 4, CGB = CG < 64 ? CG : 64, RL = 256 / CGB;
  return (size_t)2 * RL * CGB * 4 * sizeof(float);
}
static inline dim3 red_grid(int64_t rows, int C) {
  int CG = C / 4, CGB = CG < 64 ? CG : 64, RL = 256 / CGB;
  int chunks = (CG + CGB - 1) / CGB;
  int64_t bx = (rows + RL * 8 - 1) / (RL * 8);
  int64_t cap = (int64_t)kNumSMs * 4 / chunks;
  if (cap < 1) cap = 1;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  return dim3((unsigned)bx, (unsigned)chunks);
}

extern "C" int iit_bn_stats(const void* x, int dtype, int64_t rows, int c, double* sum, double* sqsum, void* stream) {
  IIT_REQUIRE(x && sum && sqsum && rows >= 1, "iit_bn_stats: bad arguments");
  if (bn_fast_ok(dtype, c, x, nullptr, nullptr, nullptr, nullptr)) return bn_stats_fast(x, dtype, rows, c, sum, sqsum, as_stream(stream));
  IIT_REQUIRE(c % 4 == 0 && (c / 4 <= 64 ? 256 % (c / 4) == 0 : (c / 4) % 64 == 0), "iit_bn_stats: unsupported channel count %d", c);
  RedPtrs2 o{{sum, sqsum}};
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_stats_kernel<T><<<red_grid(rows, c), 256, red_smem(c), as_stream(stream)>>>((const T*)x, rows, c, o);
  })
  IIT_LAUNCH_CHECK("iit_bn_stats");
  return 0;
}

extern "C" int iit_bn_finalize(double* sum, double* sqsum, int64_t rows, int c, float eps, float momentum,
                               float* mean, float* invstd, float* running_mean, float* running_var,
                               int64_t* num_batches_tracked, int clear_sums, void* stream) {
  IIT_REQUIRE(sum && sqsum && mean && invstd && rows >= 1, "iit_bn_finalize: bad arguments");
  IIT_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "iit_bn_finalize: running stats go together");
  bn_finalize_kernel<<<(c + 127) / 128, 128, 0, as_stream(stream)>>>(sum, sqsum, rows, c, eps, momentum, mean, invstd,
                                                                    running_mean, running_var,
                                                                    reinterpret_cast<long long*>(num_batches_tracked), clear_sums);
  IIT_LAUNCH_CHECK("iit_bn_finalize");
  return 0;
}

extern "C" int iit_bn_eval_affine(const float* gamma, const float* beta, const float* running_mean, const float* running_var, int c,
                                  float eps, float* scale, float* shift, void* stream) {
  IIT_REQUIRE(gamma && beta && running_mean && running_var && scale && shift && c >= 1, "iit_bn_eval_affine: bad arguments");
  bn_eval_affine_kernel<<<(c + 127) / 128, 128, 0, as_stream(stream)>>>(gamma, beta, running_mean, running_var, c, eps, scale, shift);
  IIT_LAUNCH_CHECK("iit_bn_eval_affine");
  return 0;
}

extern "C" int iit_bn_param_grads(double* sums, int c, float* dgamma, float* dbeta, int accumulate, int clear_sums, void* stream) {
  IIT_REQUIRE(sums && dgamma && dbeta && c >= 1, "iit_bn_param_grads: bad arguments");
  bn_param_grads_kernel<<<(c + 127) / 128, 128, 0, as_stream(stream)>>>(sums, c, dgamma, dbeta, accumulate, clear_sums);
  IIT_LAUNCH_CHECK("iit_bn_param_grads");
  return 0;
}

extern "C" int iit_bn_eval_invstd(const float* running_var, int c, float eps, float* invstd, void* stream) {
  IIT_REQUIRE(running_var && invstd, "iit_bn_eval_invstd: null pointer");
  bn_eval_invstd_kernel<<<(c + 127) / 128, 128, 0, as_stream(stream)>>>(running_var, c, eps, invstd);
  IIT_LAUNCH_CHECK("iit_bn_eval_invstd");
  return 0;
}

extern "C" int iit_bn_act_fwd(const void* x, const void* residual, void* y, int dtype, int64_t rows, int c, const float* mean,
                              const float* invstd, const float* gamma, const float* beta, int relu, uint8_t* relu_mask,
                              void* stream) {
  IIT_REQUIRE(x && y && mean && invstd && gamma && beta, "iit_bn_act_fwd: null pointer");
  if (!relu) relu_mask = nullptr;
  if (bn_fast_ok(dtype, c, x, residual, y, mean, invstd) && bn_fast_ok(dtype, c, gamma, beta, nullptr, nullptr, nullptr))
    return bn_act_fwd_fast(x, residual, y, dtype, rows, c, mean, invstd, gamma, beta, relu, relu_mask, as_stream(stream));
  IIT_REQUIRE(c % 4 == 0, "iit_bn_act_fwd: channels must be a multiple of 4");
  int64_t n4 = rows * c / 4;
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_act_fwd_kernel<T><<<grid_for(n4, 256 * 2), 256, 0, as_stream(stream)>>>((const T*)x, (const T*)residual, (T*)y, n4, c,
                                                                               mean, invstd, gamma, beta, relu);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_fwd");
  // the generic kernel does not write the mask: derive it from y where the 16-byte backward kernels could ask for it
  if (relu_mask && bn_fast_ok(dtype, c, y, nullptr, nullptr, nullptr, nullptr))
    return relu_mask_fast(y, dtype, rows, c, relu_mask, as_stream(stream));
  return 0;
}

extern "C" int iit_bn_act_bwd_reduce(const void* dy, const void* y, const void* x, int dtype, int64_t rows, int c,
                                     const float* mean, const float* invstd, int relu, double* dgamma, double* dbeta,
                                     const uint8_t* relu_mask, void* stream) {
  IIT_REQUIRE(dy && x && mean && invstd && dgamma && dbeta && (!relu || y), "iit_bn_act_bwd_reduce: null pointer");
  if (bn_fast_ok(dtype, c, dy, y, x, mean, invstd))
    return bn_bwd_reduce_fast(dy, y, x, dtype, rows, c, mean, invstd, relu, dgamma, dbeta, relu_mask, as_stream(stream));
  IIT_REQUIRE(c % 4 == 0 && (c / 4 <= 64 ? 256 % (c / 4) == 0 : (c / 4) % 64 == 0), "iit_bn_act_bwd_reduce: unsupported channel count %d", c);
  RedPtrs2 o{{dgamma, dbeta}};
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_reduce_kernel<T><<<red_grid(rows, c), 256, red_smem(c), as_stream(stream)>>>((const T*)dy, (const T*)y, (const T*)x,
                                                                                        rows, c, mean, invstd, relu, o);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_bwd_reduce");
  return 0;
}

extern "C" int iit_bn_act_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows,
                                    int c, const float* mean, const float* invstd, const float* gamma, const double* dgamma,
                                    const double* dbeta, int relu, const uint8_t* relu_mask, void* stream) {
  IIT_REQUIRE(dy && x && dx && mean && invstd && gamma && dgamma && dbeta && (!relu || y), "iit_bn_act_bwd_apply: null pointer");
  if (bn_fast_ok(dtype, c, dy, y, x, dx, dres) && bn_fast_ok(dtype, c, mean, invstd, gamma, nullptr, nullptr))
    return bn_bwd_apply_fast(dy, y, x, dx, dres, dtype, rows, c, mean, invstd, gamma, dgamma, dbeta, relu, relu_mask,
                             as_stream(stream));
  IIT_REQUIRE(c % 4 == 0, "iit_bn_act_bwd_apply: channels must be a multiple of 4");
  int64_t n4 = rows * c / 4;
  IIT_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_apply_kernel<T><<<grid_for(n4, 256 * 2), 256, 0, as_stream(stream)>>>((const T*)dy, (const T*)y, (const T*)x, (T*)dx,
                                                                                 (T*)dres, n4, c, 1.f / (float)rows, mean, invstd,
                                                                                 gamma, dgamma, dbeta, relu);
  })
  IIT_LAUNCH_CHECK("iit_bn_act_bwd_apply");
  return 0;
}

extern "C" int iit_bn_act_bwd_apply_box(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows,
                                        int c, const float* mean, const float* invstd, const float* gamma, const double* dgamma,
                                        const double* dbeta, int relu, const uint8_t* relu_mask, int h, int w, const int32_t* box6,
                                        void* stream) {
  const char* name = "iit_bn_act_bwd_apply_box";
  IIT_REQUIRE(dy && x && dx && mean && invstd && gamma && dgamma && dbeta && (!relu || y) && box6, "%s: null pointer", name);
  IIT_REQUIRE(h >= 1 && w >= 1 && rows % ((int64_t)h * w) == 0, "%s: rows must be whole %d x %d images", name, h, w);
  const int np = dtype == IIT_BF16 ? 8 : 4;
  IIT_REQUIRE(box6[4] % np == 0 && box6[5] % np == 0, "%s: the channel range must be a multiple of %d", name, np);
  IIT_REQUIRE(bn_fast_ok(dtype, c, dy, y, x, dx, dres) && bn_fast_ok(dtype, c, mean, invstd, gamma, nullptr, nullptr),
              "%s: 16-byte aligned tensors with a channel count the vector kernels take only", name);
  const int zb[8] = {h, w, box6[0], box6[1], box6[2], box6[3], box6[4], box6[5]};
  return bn_bwd_apply_fast(dy, y, x, dx, dres, dtype, rows, c, mean, invstd, gamma, dgamma, dbeta, relu, relu_mask,
                           as_stream(stream), zb);
}

extern "C" int iit_relu_fwd(const void* x, void* y, int dtype, int64_t n, void* stream) {
  IIT_REQUIRE(x && y && n % 4 == 0, "iit_relu_fwd: bad arguments (n must be a multiple of 4)");
  IIT_DISPATCH_DTYPE(dtype, T, { relu_fwd_kernel<T><<<grid_for(n / 4, 512), 256, 0, as_stream(stream)>>>((const T*)x, (T*)y, n / 4); })
  IIT_LAUNCH_CHECK("iit_relu_fwd");
  return 0;
}
extern "C" int iit_div_scalar(const void* x, void* y, int dtype, int64_t n, float scale, void* stream) {
  IIT_REQUIRE(x && y && n % 4 == 0 && scale != 0.f, "iit_div_scalar: bad arguments (n must be a multiple of 4, scale non-zero)");
  if (n == 0) return 0;
  IIT_DISPATCH_DTYPE(dtype, T, {
    div_scalar_kernel<T><<<grid_for(n / 4, 512), 256, 0, as_stream(stream)>>>((const T*)x, (T*)y, n / 4, scale);
  })
  IIT_LAUNCH_CHECK("iit_div_scalar");
  return 0;
}
extern "C" int iit_relu_bwd(const void* dy, const void* y, void* dx, int dtype, int64_t n, void* stream) {
  IIT_REQUIRE(dy && y && dx && n % 4 == 0, "iit_relu_bwd: bad arguments (n must be a multiple of 4)");
  IIT_DISPATCH_DTYPE(dtype, T, {
    relu_bwd_kernel<T><<<grid_for(n / 4, 512), 256, 0, as_stream(stream)>>>((const T*)dy, (const T*)y, (T*)dx, n / 4);
  })
  IIT_LAUNCH_CHECK("iit_relu_bwd");
  return 0;
}
extern "C" int iit_add_relu_fwd(const void* a, const void* b, void* y, int dtype, int64_t n, int relu, void* stream) {
  IIT_REQUIRE(a && b && y && n % 4 == 0, "iit_add_relu_fwd: bad arguments (n must be a multiple of 4)");
  IIT_DISPATCH_DTYPE(dtype, T, {
    add_relu_kernel<T><<<grid_for(n / 4, 512), 256, 0, as_stream(stream)>>>((const T*)a, (const T*)b, (T*)y, n / 4, relu);
  })
  IIT_LAUNCH_CHECK("iit_add_relu_fwd");
  return 0;
}

extern "C" int iit_maxpool3x3s2_fwd(const void* x, void* y, int8_t* idx, int dtype, int n, int h, int w, int c, void* stream) {
  IIT_REQUIRE(x && y && idx && c % 4 == 0, "iit_maxpool3x3s2_fwd: bad arguments");
  int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  int64_t total = (int64_t)n * P * Q * (c / 4);
  IIT_DISPATCH_DTYPE(dtype, T, {
    maxpool_fwd_kernel<T><<<grid_for(total, 256), 256, 0, as_stream(stream)>>>((const T*)x, (T*)y, idx, n, h, w, c, P, Q);
  })
  IIT_LAUNCH_CHECK("iit_maxpool3x3s2_fwd");
  return 0;
}
extern "C" int iit_maxpool3x3s2_bwd(const void* dy, const int8_t* idx, void* dx, int dtype, int n, int h, int w, int c,
                                    void* stream) {
  IIT_REQUIRE(dy && dx && idx && c % 4 == 0, "iit_maxpool3x3s2_bwd: bad arguments");
  int P = (h + 2 - 3) / 2 + 1, Q = (w + 2 - 3) / 2 + 1;
  int64_t total = (int64_t)n * h * w * (c / 4);
  IIT_DISPATCH_DTYPE(dtype, T, {
    maxpool_bwd_kernel<T><<<grid_for(total, 256), 256, 0, as_stream(stream)>>>((const T*)dy, idx, (T*)dx, n, h, w, c, P, Q);
  })
  IIT_LAUNCH_CHECK("iit_maxpool3x3s2_bwd");
  return 0;
}

extern "C" int iit_avgpool_fwd(const void* x, float* y, int dtype, int n, int hw, int c, void* stream) {
  IIT_REQUIRE(x && y && hw >= 1, "iit_avgpool_fwd: bad arguments");
  IIT_DISPATCH_DTYPE(dtype, T, {
    avgpool_fwd_kernel<T><<<grid_for((int64_t)n * c, 256), 256, 0, as_stream(stream)>>>((const T*)x, y, n, hw, c);
  })
  IIT_LAUNCH_CHECK("iit_avgpool_fwd");
  return 0;
}
extern "C" int iit_avgpool_bwd(const float* dy, void* dx, int dtype, int n, int hw, int c, void* stream) {
  IIT_REQUIRE(dy && dx && hw >= 1, "iit_avgpool_bwd: bad arguments");
  IIT_DISPATCH_DTYPE(dtype, T, {
    avgpool_bwd_kernel<T><<<grid_for((int64_t)n * hw * c, 256), 256, 0, as_stream(stream)>>>(dy, (T*)dx, n, hw, c);
  })
  IIT_LAUNCH_CHECK("iit_avgpool_bwd");
  return 0;
}

extern "C" int iit_linear_fwd(const float* x, const float* w, const float* b, float* y, int n, int in, int out, void* stream) {
  IIT_REQUIRE(x && w && y, "iit_linear_fwd: null pointer");
  linear_fwd_kernel<<<grid_for((int64_t)n * out, 8), 256, 0, as_stream(stream)>>>(x, w, b, y, n, in, out);
  IIT_LAUNCH_CHECK("iit_linear_fwd");
  return 0;
}
extern "C" int iit_linear_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db, int n, int in,
                              int out, void* stream) {
  IIT_REQUIRE(x && w && dy, "iit_linear_bwd: null pointer");
  cudaStream_t st = as_stream(stream);
  if (dx) {
    linear_dx_kernel<<<grid_for((int64_t)n * in, 256), 256, 0, st>>>(dy, w, dx, n, in, out);
    IIT_LAUNCH_CHECK("iit_linear_bwd(dx)");
  }
  if (dw) {
    IIT_REQUIRE(out <= LIN_MAX_O, "iit_linear_bwd: at most %d output features (got %d)", LIN_MAX_O, out);
    IIT_CUDA(cudaMemsetAsync(dw, 0, (size_t)out * in * sizeof(float), st));
    if (db) IIT_CUDA(cudaMemsetAsync(db, 0, (size_t)out * sizeof(float), st));
    const int col_blocks = (in + 63) / 64;
    int row_blocks = (2 * kNumSMs + col_blocks - 1) / col_blocks;
    if (row_blocks > (n + 31) / 32) row_blocks = (n + 31) / 32;       // at least 32 rows per block
    if (row_blocks < 1) row_blocks = 1;
    const int rows_per_block = (n + row_blocks - 1) / row_blocks;
    linear_dw_kernel<<<dim3(col_blocks, (n + rows_per_block - 1) / rows_per_block), 256, 0, st>>>(dy, x, dw, db, n, in, out,
                                                                                               rows_per_block);
    IIT_LAUNCH_CHECK("iit_linear_bwd(dw)");
  }
  return 0;
}

extern "C" int iit_sqnorm(const float* g, int64_t n, double* sqnorm, void* stream) {
  IIT_REQUIRE(g && sqnorm && ((uintptr_t)g & 15) == 0, "iit_sqnorm: g must be 16-byte aligned");
  sqnorm_kernel<<<grid_for(n / 4 + 1, 256 * 4), 256, 0, as_stream(stream)>>>(g, n, sqnorm);
  IIT_LAUNCH_CHECK("iit_sqnorm");
  return 0;
}

extern "C" int iit_adam_step(float* p, const float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2,
                             double eps, int64_t step, const double* sqnorm, double max_norm, void* stream) {
  IIT_REQUIRE(p && g && m && v && step >= 1, "iit_adam_step: bad arguments");
  // bias corrections in double on the host, as torch.optim.Adam computes them in Python floats
  double bc1 = 1.0 - pow(beta1, (double)step);
  double bc2 = 1.0 - pow(beta2, (double)step);
  adam_kernel<<<grid_for(n, 256 * 4), 256, 0, as_stream(stream)>>>(p, g, m, v, n, (float)(lr / bc1), (float)beta1, (float)beta2,
                                                                 (float)eps, (float)sqrt(bc2), sqnorm, (float)max_norm);
  IIT_LAUNCH_CHECK("iit_adam_step");
  return 0;
}

extern "C" int iit_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev, double beta1,
                                 double beta2, double eps, long long* step_dev, const double* sqnorm, double max_norm,
                                 void* stream) {
  IIT_REQUIRE(p && g && m && v && lr_dev && step_dev, "iit_adam_step_dev: null pointer");
  cudaStream_t st = as_stream(stream);
  adam_tick_kernel<<<1, 1, 0, st>>>(step_dev);
  IIT_LAUNCH_CHECK("iit_adam_step_dev(tick)");
  adam_dev_kernel<<<grid_for(n, 256 * 4), 256, 0, st>>>(p, g, m, v, n, lr_dev, beta1, beta2, (float)eps, step_dev, sqnorm,
                                                       (float)max_norm);
  IIT_LAUNCH_CHECK("iit_adam_step_dev");
  return 0;
}
