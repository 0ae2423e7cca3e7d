// (c) high-level models + cross-entropy losses, (d) IIA reduction.
//
// All HBM-bound row reductions. Row kernels use GROUP cooperating threads per row: a warp when a
// row is short (PVR: 10 classes), a 256-thread CTA when long (IOI: 50257). Loss sums are reduced in
// a fixed order (deterministic), integer counts with one atomic per block.
#include "common.cuh"

namespace iit {

__constant__ int c_class_map[10] = {1, 1, 1, 1, 2, 2, 2, 3, 3, 3};  // iit/tasks/mnist_pvr/utils.py:10

// ---------------------------------------------------------------- PVR HL
// (tr, bl, br)[class_map[tl] - 1] for one pair's variables; tl outside [-10, 9] (what torch's indexing accepts) returns
// false and leaves *out = -1 -- the reference would raise IndexError there (pvr_hl.py:50-53)
__device__ __forceinline__ bool pvr_hl_eval(const int64_t (&v)[4], int64_t* out) {
  int64_t tl = v[0];
  tl = tl < 0 ? tl + 10 : tl;  // python-style negative index
  if (tl < 0 || tl > 9) { *out = -1; return false; }
  *out = v[c_class_map[(int)tl]];  // 1..3 -> tr, bl, br
  return true;
}

__global__ void hl_pvr_kernel(const int64_t* __restrict__ bv, const int64_t* __restrict__ sv, int var,
                              int64_t batch, int64_t* __restrict__ out, int* __restrict__ err_flag) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  int64_t v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = (i == var && sv != nullptr) ? sv[b * 4 + i] : bv[b * 4 + i];
  int64_t o;
  if (!pvr_hl_eval(v, &o) && err_flag != nullptr) *err_flag = 1;
  out[b] = o;
}

// ---------------------------------------------------------------- IOI HL (sparse, last position)
constexpr int kMaxSeq = 64;

__device__ __forceinline__ bool in_names(int64_t t, const int64_t* names, int n) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    int64_t v = names[mid];
    if (v == t) return true;
    if (v < t) lo = mid + 1; else hi = mid;
  }
  return false;
}

// duplicate head (ioi_hl.py:16-24): last earlier position with the same token, else -1
__device__ __forceinline__ void dup_head(const int64_t* tok, int seq, int* dup) {
  for (int j = 0; j < seq; ++j) {
    int d = -1;
    for (int i = 0; i < j; ++i) if (tok[i] == tok[j]) d = i;
    dup[j] = d;
  }
}

// Per sample: tokens feeding the name mover, and the s-inhibition flags, under one interchange.
//   ALL_NODES   : everything from the source tokens
//   DUPLICATE   : dup from source; s_inh = base token where dup_src != -1
//   S_INHIBITION: s_inh computed wholly on the source
//   NAME_MOVER  : the output itself is the source's -> same as ALL_NODES for the output
__device__ void ioi_sparse(const int64_t* bt, const int64_t* st, int node, int seq, int64_t* mover_tok,
                           int64_t* sinh) {
  const int64_t* t = (node == IIT_IOI_ALL_NODES || node == IIT_IOI_NAME_MOVER) ? st : bt;
  int dup[kMaxSeq];
  if (node == IIT_IOI_DUPLICATE) {
    dup_head(st, seq, dup);
    for (int j = 0; j < seq; ++j) { mover_tok[j] = bt[j]; sinh[j] = dup[j] == -1 ? -1 : bt[j]; }
  } else if (node == IIT_IOI_S_INHIBITION) {
    dup_head(st, seq, dup);
    for (int j = 0; j < seq; ++j) { mover_tok[j] = bt[j]; sinh[j] = dup[j] == -1 ? -1 : st[j]; }
  } else {
    dup_head(t, seq, dup);
    for (int j = 0; j < seq; ++j) { mover_tok[j] = t[j]; sinh[j] = dup[j] == -1 ? -1 : t[j]; }
  }
}

// argmax over the vocabulary of the IOI HL model's last-position logits for sample b, one hook interchanged
__device__ __noinline__ int64_t ioi_last_argmax_row(const int64_t* __restrict__ base_tok, const int64_t* __restrict__ src_tok, int node,
                                       const int64_t* __restrict__ names, int n_names, int64_t b, int seq, int64_t V) {
  int64_t bt[kMaxSeq], st[kMaxSeq], mt[kMaxSeq], sh[kMaxSeq];
  for (int j = 0; j < seq; ++j) { bt[j] = base_tok[b * seq + j]; st[j] = src_tok ? src_tok[b * seq + j] : bt[j]; }
  ioi_sparse(bt, st, node, seq, mt, sh);
  // sparse map (key -> value) of the cumulative logits at the last position (ioi_hl.py:59-67):
  // per position p:  L[p, tok] = 10*isname (assignment), then L[p, sinh] += -15 (if sinh != -1)
  int64_t key[2 * kMaxSeq];
  float val[2 * kMaxSeq];
  int n = 0;
  auto add = [&](int64_t k, float v) {
    for (int i = 0; i < n; ++i) if (key[i] == k) { val[i] += v; return; }
    key[n] = k; val[n] = v; ++n;
  };
  for (int p = 0; p < seq; ++p) {
    int64_t tk = mt[p] < 0 ? mt[p] + V : mt[p];
    add(tk, in_names(mt[p], names, n_names) ? 10.f : 0.f);
    if (sh[p] != -1) add(sh[p] < 0 ? sh[p] + V : sh[p], -15.f);
  }
  // first maximal index over the dense row: entries not in the map are 0
  float best = (int64_t)n < V ? 0.f : -INFINITY;
  for (int i = 0; i < n; ++i) best = fmaxf(best, val[i]);
  int64_t arg = -1;
  if (best != 0.f || (int64_t)n >= V) {
    for (int i = 0; i < n; ++i) if (val[i] == best && (arg < 0 || key[i] < arg)) arg = key[i];
  }
  if (best == 0.f) {
    // smallest index whose value is 0: skip indices mapped to a non-zero value
    for (int64_t c = 0; c < V; ++c) {
      bool nz = false;
      for (int i = 0; i < n; ++i) if (key[i] == c && val[i] != 0.f) nz = true;
      if (!nz) { arg = c; break; }
    }
  }
  return arg;
}

__global__ void hl_ioi_last_argmax_kernel(const int64_t* __restrict__ base_tok, const int64_t* __restrict__ src_tok,
                                          int node, const int64_t* __restrict__ names, int n_names, int64_t batch,
                                          int seq, int64_t V, int64_t* __restrict__ out) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  out[b] = ioi_last_argmax_row(base_tok, src_tok, node, names, n_names, b, seq, V);
}

// dense [B,S,V] fp32: zero fill is done by the caller-visible kernel below (grid-stride), then the
// sparse cumulative entries are scattered. One CTA per sample.
__global__ void hl_ioi_dense_kernel(const int64_t* __restrict__ base_tok, const int64_t* __restrict__ src_tok, int node,
                                    const int64_t* __restrict__ names, int n_names, int64_t batch, int seq, int64_t V,
                                    float* __restrict__ out) {
  int64_t b = blockIdx.x;
  float* ob = out + b * (int64_t)seq * V;
  // zero this sample's [S,V] slab
  int64_t total = (int64_t)seq * V;
  if ((total & 3) == 0 && ((uintptr_t)out & 15) == 0) {
    float4* o4 = reinterpret_cast<float4*>(ob);
    for (int64_t i = threadIdx.x; i < (total >> 2); i += blockDim.x) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    for (int64_t i = threadIdx.x; i < total; i += blockDim.x) ob[i] = 0.f;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  int64_t bt[kMaxSeq], st[kMaxSeq], mt[kMaxSeq], sh[kMaxSeq];
  for (int j = 0; j < seq; ++j) { bt[j] = base_tok[b * seq + j]; st[j] = src_tok ? src_tok[b * seq + j] : bt[j]; }
  ioi_sparse(bt, st, node, seq, mt, sh);
  // per position delta, then running sum over later positions (cumsum along seq)
  for (int p = 0; p < seq; ++p) {
    int64_t tk = mt[p] < 0 ? mt[p] + V : mt[p];
    float a = in_names(mt[p], names, n_names) ? 10.f : 0.f;
    float c = 0.f;
    int64_t sk = -1;
    if (sh[p] != -1) { sk = sh[p] < 0 ? sh[p] + V : sh[p]; c = -15.f; }
    for (int q = p; q < seq; ++q) {
      ob[(int64_t)q * V + tk] += a;
      if (sk >= 0) ob[(int64_t)q * V + sk] += c;
    }
  }
}

// ---------------------------------------------------------------- row reductions
struct ArgMax { float v; int64_t i; };
__device__ __forceinline__ bool better(float v, int64_t i, float bv, int64_t bi) {
  // torch.argmax: first maximal index; NaN is maximal
  bool vn = v != v, bn = bv != bv;
  if (bi < 0) return true;
  if (vn != bn) return vn;
  if (vn && bn) return i < bi;
  return v > bv || (v == bv && i < bi);
}
__device__ __forceinline__ ArgMax warp_argmax(ArgMax a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, a.v, o);
    int64_t oi = __shfl_xor_sync(0xffffffffu, a.i, o);
    if (oi >= 0 && better(ov, oi, a.v, a.i)) { a.v = ov; a.i = oi; }
  }
  return a;
}

// GROUP = 32: one warp per row, 8 rows per 256-thread CTA. GROUP = 256: one CTA per row.
template <typename T, int GROUP>
__global__ void __launch_bounds__(256) iia_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                  int64_t row_stride, const int64_t* __restrict__ hl,
                                                  const uint8_t* __restrict__ changed, int64_t* __restrict__ top1,
                                                  unsigned long long* __restrict__ counts) {
  constexpr int RPB = 256 / GROUP;
  __shared__ ArgMax s_part[8];
  __shared__ unsigned int s_hits, s_den;
  if (threadIdx.x == 0) { s_hits = 0; s_den = 0; }
  __syncthreads();
  const int g = threadIdx.x / GROUP, t = threadIdx.x % GROUP;
  for (int64_t r0 = (int64_t)blockIdx.x * RPB; r0 < rows; r0 += (int64_t)gridDim.x * RPB) {
    int64_t r = r0 + g;
    ArgMax a{0.f, -1};
    if (r < rows) {
      const T* row = logits + r * row_stride;
      for (int64_t c = t; c < classes; c += GROUP) {
        float v = to_f<T>(row[c]);
        if (better(v, c, a.v, a.i)) { a.v = v; a.i = c; }
      }
    }
    a = warp_argmax(a);
    if (GROUP == 256) {
      if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = a;
      __syncthreads();
      if (threadIdx.x < 32) {
        ArgMax p = threadIdx.x < 8 ? s_part[threadIdx.x] : ArgMax{0.f, -1};
        a = warp_argmax(p);
      }
      __syncthreads();
    }
    if (t == 0 && r < rows && (GROUP == 32 || threadIdx.x == 0)) {
      if (top1) top1[r] = a.i;
      if (hl && counts) {
        bool ch = changed ? changed[r] != 0 : true;
        if (ch) {
          atomicAdd(&s_den, 1u);
          if (a.i == hl[r]) atomicAdd(&s_hits, 1u);
        }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && counts && (s_hits | s_den)) {
    if (s_hits) atomicAdd(&counts[0], (unsigned long long)s_hits);
    if (s_den) atomicAdd(&counts[1], (unsigned long long)s_den);
  }
}

// online logsumexp of one row by GROUP threads; result valid in every thread of the group's warp 0
template <typename T, int GROUP>
__device__ __forceinline__ float row_lse(const T* __restrict__ row, int64_t classes, int t, float* s_m, float* s_s) {
  float m = -INFINITY, s = 0.f;
  for (int64_t c = t; c < classes; c += GROUP) {
    float v = to_f<T>(row[c]);
    if (v > m) { s = s * __expf(m - v) + 1.f; m = v; }
    else if (v != -INFINITY) s += __expf(v - m);
  }
  // combine (m, s) pairs
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    float nm = fmaxf(m, om);
    float a = (m == -INFINITY) ? 0.f : s * __expf(m - nm);
    float b = (om == -INFINITY) ? 0.f : os * __expf(om - nm);
    m = nm; s = a + b;
  }
  if (GROUP == 256) {
    int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_m[w] = m; s_s[w] = s; }
    __syncthreads();
    m = threadIdx.x < 8 ? s_m[threadIdx.x & 7] : -INFINITY;
    s = threadIdx.x < 8 ? s_s[threadIdx.x & 7] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
      float nm = fmaxf(m, om);
      float a = (m == -INFINITY) ? 0.f : s * __expf(m - nm);
      float b = (om == -INFINITY) ? 0.f : os * __expf(om - nm);
      m = nm; s = a + b;
    }
    __syncthreads();
  }
  return m + __logf(s);
}

// the same for a bf16 row whose start is 16-byte aligned: 8 values per load, one rescale per pack instead of per element
template <int GROUP>
__device__ __forceinline__ float row_lse_bf16x8(const __nv_bfloat16* __restrict__ row, int64_t classes, int t, float* s_m, float* s_s) {
  float m = -INFINITY, s = 0.f;
  const int packs = (int)(classes >> 3);
  auto fold = [&](const uint4* w, int n) {           // n packs: one running-maximum update and one rescale for all of them
    float pm = -INFINITY;
    for (int u = 0; u < n; ++u) {
      const uint32_t ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) pm = fmaxf(pm, fmaxf(__uint_as_float(ww[i] << 16), __uint_as_float(ww[i] & 0xffff0000u)));
    }
    if (pm == -INFINITY) return;
    const float nm = fmaxf(m, pm);
    float acc = (m == -INFINITY) ? 0.f : s * __expf(m - nm);
    for (int u = 0; u < n; ++u) {
      const uint32_t ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) acc += __expf(__uint_as_float(ww[i] << 16) - nm) + __expf(__uint_as_float(ww[i] & 0xffff0000u) - nm);
    }
    m = nm; s = acc;                                 // exp(-inf) = 0 for masked entries
  };
  int pk = t;
  for (; pk + 3 * GROUP < packs; pk += 4 * GROUP) {  // four 16-byte loads in flight per thread
    uint4 w[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) w[u] = *reinterpret_cast<const uint4*>(row + (int64_t)(pk + u * GROUP) * 8);
    fold(w, 4);
  }
  for (; pk < packs; pk += GROUP) {
    const uint4 w = *reinterpret_cast<const uint4*>(row + (int64_t)pk * 8);
    fold(&w, 1);
  }
  for (int64_t c = (int64_t)packs * 8 + t; c < classes; c += GROUP) {      // tail of fewer than 8 elements
    const float v = __bfloat162float(row[c]);
    if (v > m) { s = s * __expf(m - v) + 1.f; m = v; }
    else if (v != -INFINITY) s += __expf(v - m);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    float nm = fmaxf(m, om);
    float a = (m == -INFINITY) ? 0.f : s * __expf(m - nm);
    float b = (om == -INFINITY) ? 0.f : os * __expf(om - nm);
    m = nm; s = a + b;
  }
  if (GROUP == 256) {
    int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_m[w] = m; s_s[w] = s; }
    __syncthreads();
    m = threadIdx.x < 8 ? s_m[threadIdx.x & 7] : -INFINITY;
    s = threadIdx.x < 8 ? s_s[threadIdx.x & 7] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
      float nm = fmaxf(m, om);
      float a = (m == -INFINITY) ? 0.f : s * __expf(m - nm);
      float b = (om == -INFINITY) ? 0.f : os * __expf(om - nm);
      m = nm; s = a + b;
    }
    __syncthreads();
  }
  return m + __logf(s);
}

// (c) ONE launch for an IIT loss term's high-level side and its loss: per row the HL model's intervened output (the class
// index the LL logits are scored against), the row's logsumexp, its negative log-likelihood; the block that finishes last
// adds the rows' values in a fixed order (deterministic) and writes the mean. mode 0 takes the targets as given.
template <typename T, int GROUP>
__global__ void __launch_bounds__(256) hl_ce_fwd_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                        int64_t row_stride, iit_hl_spec_t hl, int64_t* __restrict__ target,
                                                        float* __restrict__ out_lse, float* __restrict__ row_nll,
                                                        float* __restrict__ loss, unsigned int* __restrict__ ticket,
                                                        int* __restrict__ err_flag, int vec_ok) {
  constexpr int RPB = 256 / GROUP;
  __shared__ float s_m[8], s_s[8], s_red[8];
  __shared__ int s_last;
  const int g = threadIdx.x / GROUP, t = threadIdx.x % GROUP;
  for (int64_t r0 = (int64_t)blockIdx.x * RPB; r0 < rows; r0 += (int64_t)gridDim.x * RPB) {
    const int64_t r = r0 + g;
    int64_t tgt = -1;
    if (t == 0 && r < rows) {
      if (hl.mode == 1) {            // MNIST-PVR: (tr, bl, br)[class_map[tl] - 1] with variable `node` taken from the source run
        int64_t v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = (i == hl.node && hl.src != nullptr) ? hl.src[r * 4 + i] : hl.base[r * 4 + i];
        if (!pvr_hl_eval(v, &tgt) && err_flag != nullptr) *err_flag = 1;
        target[r] = tgt;
      } else if (hl.mode == 2) {     // IOI: argmax of the HL logits at the last position
        tgt = ioi_last_argmax_row(hl.base, hl.src, hl.node, hl.names, hl.n_names, r, hl.seq, hl.d_vocab);
        target[r] = tgt;
      } else {
        tgt = target[r];
      }
    }
    float l = 0.f;
    if (GROUP == 256 || r < rows) {
      const T* row = logits + (r < rows ? r : 0) * row_stride;
      if (sizeof(T) == 2 && vec_ok) l = row_lse_bf16x8<GROUP>(reinterpret_cast<const __nv_bfloat16*>(row), classes, t, s_m, s_s);
      else l = row_lse<T, GROUP>(row, classes, t, s_m, s_s);
    }
    if (r < rows && t == 0) {
      out_lse[r] = l;
      float nll = l;
      if (tgt >= 0 && tgt < classes) nll -= to_f<T>(logits[r * row_stride + tgt]);
      else { nll = NAN; if (err_flag != nullptr) *err_flag = 1; }      // torch raises for a target outside [0, classes)
      row_nll[r] = nll;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float acc = 0.f;
  for (int64_t r = threadIdx.x; r < rows; r += 256) acc += __ldcg(row_nll + r);
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) v += s_red[i];
    *loss = v / (float)rows;
    *ticket = 0u;                    // ready for the next launch
  }
}

// d[r, c] = (softmax(logits[r])[c] - [c == target[r]]) * g / rows for c < classes, 0 for classes <= c < d_cols: bf16 rows whose
// starts are 16-byte aligned, 8 columns per thread
__global__ void __launch_bounds__(256) ce_index_bwd_bf16x8_kernel(const __nv_bfloat16* __restrict__ logits, int rows, int classes,
                                                                  int64_t row_stride, const int64_t* __restrict__ target,
                                                                  const float* __restrict__ lse, const float* __restrict__ gout,
                                                                  __nv_bfloat16* __restrict__ dl, int64_t d_row_stride, int d_cols) {
  // one block per row (grid-stride over rows), four 16-byte loads in flight per thread: a block streams its row
  const float g = (gout ? *gout : 1.f) / (float)rows;
  const int packs = (d_cols + 7) >> 3, full = classes >> 3;       // packs that lie wholly inside the classes
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const float l = lse[r];
    const int tg = (int)target[r];
    const __nv_bfloat16* lrow = logits + (int64_t)r * row_stride;
    __nv_bfloat16* drow = dl + (int64_t)r * d_row_stride;
    auto emit = [&](int pk, const uint4& w) {
      const int c0 = pk * 8;
      const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float p0 = __expf(__uint_as_float(ww[i] << 16) - l), p1 = __expf(__uint_as_float(ww[i] & 0xffff0000u) - l);
        if (c0 + 2 * i == tg) p0 -= 1.f;
        if (c0 + 2 * i + 1 == tg) p1 -= 1.f;
        __nv_bfloat162 h = __floats2bfloat162_rn(p0 * g, p1 * g);
        o[i] = *reinterpret_cast<uint32_t*>(&h);
      }
      *reinterpret_cast<uint4*>(drow + c0) = make_uint4(o[0], o[1], o[2], o[3]);
    };
    int pk = threadIdx.x;
    for (; pk + 3 * 256 < full; pk += 4 * 256) {
      uint4 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) w[u] = *reinterpret_cast<const uint4*>(lrow + (int64_t)(pk + u * 256) * 8);
#pragma unroll
      for (int u = 0; u < 4; ++u) emit(pk + u * 256, w[u]);
    }
    for (; pk < full; pk += 256) emit(pk, *reinterpret_cast<const uint4*>(lrow + (int64_t)pk * 8));
    // the pack that straddles `classes` and the padding packs: element by element (exp(-inf) = 0 gives the zero gradient)
    for (pk = full + threadIdx.x; pk < packs; pk += 256) {
      const int c0 = pk * 8;
      for (int i = 0; i < 8 && c0 + i < d_cols; ++i) {
        float pv = 0.f;
        if (c0 + i < classes) { pv = __expf(__bfloat162float(lrow[c0 + i]) - l); if (c0 + i == tg) pv -= 1.f; }
        drow[c0 + i] = __float2bfloat16_rn(pv * g);
      }
    }
  }
}

template <typename T, int GROUP>
__global__ void __launch_bounds__(256) ce_lse_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                     int64_t row_stride, float* __restrict__ out_lse) {
  constexpr int RPB = 256 / GROUP;
  __shared__ float s_m[8], s_s[8];
  const int g = threadIdx.x / GROUP, t = threadIdx.x % GROUP;
  for (int64_t r0 = (int64_t)blockIdx.x * RPB; r0 < rows; r0 += (int64_t)gridDim.x * RPB) {
    int64_t r = r0 + g;
    float l = 0.f;
    if (GROUP == 256 || r < rows) l = row_lse<T, GROUP>(logits + (r < rows ? r : 0) * row_stride, classes, t, s_m, s_s);
    if (r < rows && t == 0) out_lse[r] = l;
  }
}

// loss = sum_r (lse[r]*w[r] - picked[r]) / rows, fixed-order tree in one CTA (deterministic)
template <typename T>
__global__ void __launch_bounds__(1024) ce_index_loss_kernel(const T* __restrict__ logits, int64_t rows, int64_t row_stride,
                                                             const int64_t* __restrict__ target,
                                                             const float* __restrict__ lse, float* __restrict__ loss) {
  __shared__ float s[32];
  float acc = 0.f;
  for (int64_t r = threadIdx.x; r < rows; r += blockDim.x)
    acc += lse[r] - to_f<T>(logits[r * row_stride + target[r]]);
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = s[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) *loss += v / (float)rows;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) ce_index_bwd_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                           int64_t row_stride, const int64_t* __restrict__ target,
                                                           const float* __restrict__ lse, const float* __restrict__ gout,
                                                           T* __restrict__ dl, int64_t d_row_stride) {
  const float g = (gout ? *gout : 1.f) / (float)rows;
  const int64_t total = rows * classes;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / classes, c = i - r * classes;
    float p = __expf(to_f<T>(logits[r * row_stride + c]) - lse[r]);
    if (c == target[r]) p -= 1.f;
    dl[r * d_row_stride + c] = from_f<T>(p * g);
  }
}

// probability-target CE: per row psum = sum_c p, dot = sum_c p*l
template <typename T>
__global__ void __launch_bounds__(256) ce_prob_rows_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                           int64_t row_stride, const float* __restrict__ prob,
                                                           int64_t p_row_stride, const float* __restrict__ lse,
                                                           float* __restrict__ row_psum, float* __restrict__ row_nll) {
  __shared__ float s_a[8], s_b[8];
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    float ps = 0.f, dot = 0.f;
    for (int64_t c = threadIdx.x; c < classes; c += blockDim.x) {
      float p = prob[r * p_row_stride + c];
      if (p != 0.f) { ps += p; dot += p * to_f<T>(logits[r * row_stride + c]); }
    }
    ps = warp_sum(ps); dot = warp_sum(dot);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = ps; s_b[threadIdx.x >> 5] = dot; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float a = 0.f, b = 0.f;
      for (int i = 0; i < 8; ++i) { a += s_a[i]; b += s_b[i]; }
      row_psum[r] = a;
      row_nll[r] = lse[r] * a - b;
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(1024) sum_rows_kernel(const float* __restrict__ v, int64_t rows, float* __restrict__ loss) {
  __shared__ float s[32];
  float acc = 0.f;
  for (int64_t r = threadIdx.x; r < rows; r += blockDim.x) acc += v[r];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float x = warp_sum(s[threadIdx.x]);
    if (threadIdx.x == 0) *loss += x / (float)rows;
  }
}
template <typename T>
__global__ void __launch_bounds__(256) ce_prob_bwd_kernel(const T* __restrict__ logits, int64_t rows, int64_t classes,
                                                          int64_t row_stride, const float* __restrict__ prob,
                                                          int64_t p_row_stride, const float* __restrict__ lse,
                                                          const float* __restrict__ psum, const float* __restrict__ gout,
                                                          T* __restrict__ dl, int64_t d_row_stride) {
  const float g = (gout ? *gout : 1.f) / (float)rows;
  const int64_t total = rows * classes;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / classes, c = i - r * classes;
    float sm = __expf(to_f<T>(logits[r * row_stride + c]) - lse[r]);
    dl[r * d_row_stride + c] = from_f<T>((sm * psum[r] - prob[r * p_row_stride + c]) * g);
  }
}

}  // namespace iit

using namespace iit;

extern "C" int iit_hl_pvr(const int64_t* base_vars, const int64_t* src_vars, int var, int64_t batch, int64_t* hl_out,
                          int32_t* err_flag, void* stream) {
  IIT_REQUIRE(base_vars && hl_out && batch >= 0, "iit_hl_pvr: bad arguments");
  IIT_REQUIRE(var < 4, "iit_hl_pvr: var must be < 4 (got %d)", var);
  if (batch == 0) return 0;
  hl_pvr_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, as_stream(stream)>>>(base_vars, src_vars, var, batch, hl_out,
                                                                                  err_flag);
  IIT_LAUNCH_CHECK("iit_hl_pvr");
  return 0;
}

extern "C" int iit_hl_ioi_last_argmax(const int64_t* base_tok, const int64_t* src_tok, int node, const int64_t* names,
                                      int n_names, int64_t batch, int seq, int64_t d_vocab, int64_t* out_argmax,
                                      void* stream) {
  IIT_REQUIRE(base_tok && out_argmax && (names || n_names == 0), "iit_hl_ioi_last_argmax: null pointer");
  IIT_REQUIRE(seq >= 1 && seq <= kMaxSeq, "iit_hl_ioi_last_argmax: seq must be in [1,%d]", kMaxSeq);
  IIT_REQUIRE(node >= 0 && node <= 4, "iit_hl_ioi_last_argmax: unknown node %d", node);
  if (batch == 0) return 0;
  hl_ioi_last_argmax_kernel<<<(unsigned)((batch + 63) / 64), 64, 0, as_stream(stream)>>>(
      base_tok, src_tok, node, names, n_names, batch, seq, d_vocab, out_argmax);
  IIT_LAUNCH_CHECK("iit_hl_ioi_last_argmax");
  return 0;
}

extern "C" int iit_hl_ioi_dense(const int64_t* base_tok, const int64_t* src_tok, int node, const int64_t* names,
                                int n_names, int64_t batch, int seq, int64_t d_vocab, float* out, void* stream) {
  IIT_REQUIRE(base_tok && out && (names || n_names == 0), "iit_hl_ioi_dense: null pointer");
  IIT_REQUIRE(seq >= 1 && seq <= kMaxSeq, "iit_hl_ioi_dense: seq must be in [1,%d]", kMaxSeq);
  IIT_REQUIRE(node >= 0 && node <= 4, "iit_hl_ioi_dense: unknown node %d", node);
  if (batch == 0) return 0;
  hl_ioi_dense_kernel<<<(unsigned)batch, 256, 0, as_stream(stream)>>>(base_tok, src_tok, node, names, n_names, batch, seq,
                                                                     d_vocab, out);
  IIT_LAUNCH_CHECK("iit_hl_ioi_dense");
  return 0;
}

extern "C" int iit_iia_count(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                             const int64_t* hl, const uint8_t* changed, int64_t* top1, unsigned long long* counts,
                             void* stream) {
  IIT_REQUIRE(logits && rows >= 0 && classes >= 1, "iit_iia_count: bad arguments");
  IIT_REQUIRE(!hl || counts, "iit_iia_count: counts required when hl is given");
  if (rows == 0) return 0;
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (classes <= 1024) {
      iia_kernel<T, 32><<<grid_for(rows, 8), 256, 0, st>>>((const T*)logits, rows, classes, row_stride, hl, changed, top1, counts);
    } else {
      iia_kernel<T, 256><<<grid_for(rows, 1), 256, 0, st>>>((const T*)logits, rows, classes, row_stride, hl, changed, top1, counts);
    }
  })
  IIT_LAUNCH_CHECK("iit_iia_count");
  return 0;
}

template <typename T>
static int launch_lse(const T* logits, int64_t rows, int64_t classes, int64_t row_stride, float* lse, cudaStream_t st) {
  if (classes <= 1024) ce_lse_kernel<T, 32><<<grid_for(rows, 8), 256, 0, st>>>(logits, rows, classes, row_stride, lse);
  else ce_lse_kernel<T, 256><<<grid_for(rows, 1), 256, 0, st>>>(logits, rows, classes, row_stride, lse);
  IIT_LAUNCH_CHECK("ce_lse");
  return 0;
}

extern "C" int iit_ce_index_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                                const int64_t* target, float* row_lse_out, float* loss, void* stream) {
  IIT_REQUIRE(logits && target && row_lse_out && loss && rows >= 1 && classes >= 1, "iit_ce_index_fwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (launch_lse<T>((const T*)logits, rows, classes, row_stride, row_lse_out, st)) return 2;
    ce_index_loss_kernel<T><<<1, 1024, 0, st>>>((const T*)logits, rows, row_stride, target, row_lse_out, loss);
  })
  IIT_LAUNCH_CHECK("iit_ce_index_fwd");
  return 0;
}

extern "C" int iit_hl_ce_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                             const iit_hl_spec_t* hl, int64_t* target, float* row_lse_out, float* row_nll, float* loss,
                             unsigned int* ticket, int32_t* err_flag, void* stream) {
  IIT_REQUIRE(logits && hl && target && row_lse_out && row_nll && loss && ticket && rows >= 1 && classes >= 1,
              "iit_hl_ce_fwd: bad arguments");
  IIT_REQUIRE(hl->mode >= 0 && hl->mode <= 2, "iit_hl_ce_fwd: mode must be 0 (targets given), 1 (MNIST-PVR HL) or 2 (IOI HL)");
  IIT_REQUIRE(hl->mode == 0 || hl->base != nullptr, "iit_hl_ce_fwd: the HL model needs the base run's variables / tokens");
  IIT_REQUIRE(hl->mode != 1 || hl->node < 4, "iit_hl_ce_fwd: PVR variable index must be < 4 (got %d)", hl->node);
  IIT_REQUIRE(hl->mode != 2 || (hl->seq >= 1 && hl->seq <= kMaxSeq && hl->names != nullptr), "iit_hl_ce_fwd: IOI needs 1 <= seq <= %d "
              "and the sorted name ids", kMaxSeq);
  cudaStream_t st = as_stream(stream);
  const int vec_ok = dtype == IIT_BF16 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0 && row_stride % 8 == 0;
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (classes <= 1024)
      hl_ce_fwd_kernel<T, 32><<<grid_for(rows, 8), 256, 0, st>>>((const T*)logits, rows, classes, row_stride, *hl, target, row_lse_out,
                                                                 row_nll, loss, ticket, err_flag, vec_ok);
    else
      hl_ce_fwd_kernel<T, 256><<<grid_for(rows, 1), 256, 0, st>>>((const T*)logits, rows, classes, row_stride, *hl, target, row_lse_out,
                                                                  row_nll, loss, ticket, err_flag, vec_ok);
  })
  IIT_LAUNCH_CHECK("iit_hl_ce_fwd");
  return 0;
}

extern "C" int iit_ce_index_bwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                                const int64_t* target, const float* row_lse_in, const float* grad_out, void* dlogits,
                                int64_t d_row_stride, void* stream) {
  IIT_REQUIRE(logits && target && row_lse_in && dlogits && rows >= 1, "iit_ce_index_bwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  // bf16 rows with 16-byte aligned starts: 8 columns per thread; every column up to the row pitch is written (the gradient of
  // padding columns is zero), so a padded logits buffer gets a padded gradient with no slicing by the caller
  if (dtype == IIT_BF16 && ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(dlogits)) & 15) == 0 &&
      row_stride % 8 == 0 && d_row_stride % 8 == 0 && rows < 65536 && d_row_stride < (1ll << 31) && d_row_stride >= classes) {
    const int d_cols = (int)d_row_stride;
    const unsigned grid = (unsigned)(rows < (int64_t)kNumSMs * 8 ? rows : (int64_t)kNumSMs * 8);
    ce_index_bwd_bf16x8_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)logits, (int)rows, (int)classes, row_stride, target,
                                                     row_lse_in, grad_out, (__nv_bfloat16*)dlogits, d_row_stride, d_cols);
    IIT_LAUNCH_CHECK("iit_ce_index_bwd");
    return 0;
  }
  IIT_DISPATCH_DTYPE(dtype, T, {
    ce_index_bwd_kernel<T><<<grid_for(rows * classes, 256 * 4), 256, 0, st>>>((const T*)logits, rows, classes, row_stride,
                                                                             target, row_lse_in, grad_out, (T*)dlogits,
                                                                             d_row_stride);
  })
  IIT_LAUNCH_CHECK("iit_ce_index_bwd");
  return 0;
}

extern "C" int iit_ce_prob_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                               const float* prob, int64_t p_row_stride, float* row_lse_out, float* row_psum, float* loss,
                               void* stream) {
  IIT_REQUIRE(logits && prob && row_lse_out && row_psum && loss && rows >= 1, "iit_ce_prob_fwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  // row_psum doubles as scratch for the per-row nll: layout [rows] psum followed by [rows] nll
  IIT_DISPATCH_DTYPE(dtype, T, {
    if (launch_lse<T>((const T*)logits, rows, classes, row_stride, row_lse_out, st)) return 2;
    ce_prob_rows_kernel<T><<<grid_for(rows, 1), 256, 0, st>>>((const T*)logits, rows, classes, row_stride, prob,
                                                             p_row_stride, row_lse_out, row_psum, row_psum + rows);
    IIT_LAUNCH_CHECK("ce_prob_rows");
    sum_rows_kernel<<<1, 1024, 0, st>>>(row_psum + rows, rows, loss);
  })
  IIT_LAUNCH_CHECK("iit_ce_prob_fwd");
  return 0;
}

extern "C" int iit_ce_prob_bwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                               const float* prob, int64_t p_row_stride, const float* row_lse_in, const float* row_psum,
                               const float* grad_out, void* dlogits, int64_t d_row_stride, void* stream) {
  IIT_REQUIRE(logits && prob && row_lse_in && row_psum && dlogits && rows >= 1, "iit_ce_prob_bwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  IIT_DISPATCH_DTYPE(dtype, T, {
    ce_prob_bwd_kernel<T><<<grid_for(rows * classes, 256 * 4), 256, 0, st>>>((const T*)logits, rows, classes, row_stride,
                                                                            prob, p_row_stride, row_lse_in, row_psum,
                                                                            grad_out, (T*)dlogits, d_row_stride);
  })
  IIT_LAUNCH_CHECK("iit_ce_prob_bwd");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// KL(b || a) per row, the metric of the resample-ablation evaluation:
//   kl[r] = sum_c  B[r,c] * (log B[r,c] - log A[r,c]),   terms with B[r,c] == 0 are 0 (xlogy),
// where A / B are the rows themselves when they are probability vectors already, else softmax(row)
// (reference: iit/utils/eval_metrics.py:4-20 followed by F.kl_div(a.log(), b, reduction="none").sum(-1)).
// One block per row: pass 1 max and sum of exponentials of both rows, pass 2 the sum. Rows are read twice.
// ------------------------------------------------------------------------------------------------
namespace iit {

__device__ __forceinline__ float block_reduce(float v, bool is_max, float* s) {
  v = is_max ? warp_max(v) : warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) s[w] = v;
  __syncthreads();
  float r = is_max ? -INFINITY : 0.f;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) r = is_max ? fmaxf(r, s[i]) : r + s[i];
  return r;
}

__global__ void __launch_bounds__(256) kl_rows_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t classes,
                                                      int64_t a_stride, int64_t b_stride, int a_is_pmf, int b_is_pmf,
                                                      float* __restrict__ out) {
  __shared__ float s[8];
  const float* ar = a + (int64_t)blockIdx.x * a_stride;
  const float* br = b + (int64_t)blockIdx.x * b_stride;
  float ma = -INFINITY, mb = -INFINITY;
  for (int64_t c = threadIdx.x; c < classes; c += blockDim.x) { ma = fmaxf(ma, ar[c]); mb = fmaxf(mb, br[c]); }
  ma = block_reduce(ma, true, s);
  mb = block_reduce(mb, true, s);
  float sa = 0.f, sb = 0.f;
  for (int64_t c = threadIdx.x; c < classes; c += blockDim.x) { sa += expf(ar[c] - ma); sb += expf(br[c] - mb); }
  sa = block_reduce(sa, false, s);
  sb = block_reduce(sb, false, s);
  const float lza = ma + logf(sa), lzb = mb + logf(sb);      // log partition functions
  float kl = 0.f;
  for (int64_t c = threadIdx.x; c < classes; c += blockDim.x) {
    const float la = a_is_pmf ? logf(ar[c]) : ar[c] - lza;
    const float pb = b_is_pmf ? br[c] : expf(br[c] - lzb);
    const float lb = b_is_pmf ? logf(br[c]) : br[c] - lzb;
    if (pb > 0.f) kl += pb * (lb - la);
  }
  kl = block_reduce(kl, false, s);
  if (threadIdx.x == 0) out[blockIdx.x] = kl;
}

}  // namespace iit

extern "C" int iit_kl_div_rows(const float* a, const float* b, int64_t rows, int64_t classes, int64_t a_row_stride,
                               int64_t b_row_stride, int a_is_pmf, int b_is_pmf, float* kl, void* stream) {
  IIT_REQUIRE(a && b && kl && rows >= 0 && classes >= 1, "iit_kl_div_rows: bad arguments");
  if (rows == 0) return 0;
  IIT_REQUIRE(rows < (1ll << 31), "iit_kl_div_rows: too many rows");
  iit::kl_rows_kernel<<<(unsigned)rows, 256, 0, iit::as_stream(stream)>>>(a, b, classes, a_row_stride, b_row_stride, a_is_pmf,
                                                                        b_is_pmf, kl);
  IIT_LAUNCH_CHECK("iit_kl_div_rows");
  return 0;
}
