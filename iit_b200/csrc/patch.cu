// (a) TorchIndex box patch: dst[box] = src[box] / grad[box] = 0 / compact capture.
//
// HBM-bound byte movement. The host normalises the box (drops unit dims, sorts by stride, merges
// mergeable dims) and picks the widest vector unit (16/8/4/2 B) that every offset is aligned to.
// Device mapping: the two innermost box dims form a "plane" of e2*e1 units that lanes walk with unit
// stride (coalesced when the innermost stride is 1); the remaining (<=4) outer dims are decoded once
// per warp work item. Each lane moves 4 units per item: 4 loads in flight, then 4 stores.
#include <algorithm>

#include "common.cuh"

namespace iit {

struct PatchDesc {
  int n_outer;
  uint32_t oe[4];   // outer extents
  int64_t ods[4];   // outer dst strides (units)
  int64_t oss[4];   // outer src strides (units)
  uint32_t e2, e1;  // plane extents
  int64_t ds2, ss2, ds1, ss1;
  uint32_t chunks;  // chunks of kChunk units per plane
  uint64_t items;   // n_outer_total * chunks
};

constexpr int kLaneUnits = 4;
constexpr int kChunk = 32 * kLaneUnits;
enum { MODE_COPY = 0, MODE_ZERO = 1 };

template <typename U, int MODE>
__global__ void __launch_bounds__(256) patch_kernel(U* __restrict__ dst, const U* __restrict__ src, PatchDesc d) {
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint32_t plane = d.e2 * d.e1;
  for (uint64_t item = warp; item < d.items; item += nwarps) {
    uint64_t o = item / d.chunks;
    uint32_t chunk = (uint32_t)(item - o * d.chunks);
    int64_t doff = 0, soff = 0;
#pragma unroll
    for (int i = 3; i >= 0; --i) {
      if (i < d.n_outer) {
        uint64_t q = o / d.oe[i];
        uint32_t r = (uint32_t)(o - q * d.oe[i]);
        doff += (int64_t)r * d.ods[i];
        soff += (int64_t)r * d.oss[i];
        o = q;
      }
    }
    U v[kLaneUnits];
    int64_t dpos[kLaneUnits];
    bool ok[kLaneUnits];
#pragma unroll
    for (int j = 0; j < kLaneUnits; ++j) {
      uint32_t u = chunk * kChunk + j * 32 + lane;
      ok[j] = u < plane;
      uint32_t i2 = u / d.e1, i1 = u - i2 * d.e1;
      dpos[j] = doff + (int64_t)i2 * d.ds2 + (int64_t)i1 * d.ds1;
      if (MODE == MODE_COPY) {
        if (ok[j]) v[j] = src[soff + (int64_t)i2 * d.ss2 + (int64_t)i1 * d.ss1];
      }
    }
#pragma unroll
    for (int j = 0; j < kLaneUnits; ++j) {
      if (ok[j]) {
        if (MODE == MODE_COPY) dst[dpos[j]] = v[j];
        else { U z; memset(&z, 0, sizeof(U)); dst[dpos[j]] = z; }
      }
    }
  }
}

struct Dim { int64_t ext, ds, ss; };

// Normalise + choose unit. Returns unit bytes; fills desc and base element offsets.
static int plan(const iit_box_t* b, bool compact_dst, const void* dstp, const void* srcp, PatchDesc* pd,
                int64_t* dst_base, int64_t* src_base, bool* empty) {
  Dim dims[IIT_MAX_RANK];
  int n = 0;
  int64_t dbase = 0, sbase = 0;
  *empty = false;
  // compact destination strides: dense row-major over the box extents
  int64_t cstride[IIT_MAX_RANK];
  int64_t acc = 1;
  for (int i = b->rank - 1; i >= 0; --i) { cstride[i] = acc; acc *= b->extent[i]; }
  for (int i = 0; i < b->rank; ++i) {
    if (b->extent[i] <= 0) { *empty = true; return 0; }
    int64_t ds = compact_dst ? cstride[i] : b->dst_stride[i];
    if (!compact_dst) dbase += b->start[i] * b->dst_stride[i];
    int64_t ss = srcp ? b->src_stride[i] : ds;  // zero-fill has no source: mirror dst
    sbase += b->start[i] * ss;
    if (b->extent[i] == 1) continue;
    dims[n++] = {b->extent[i], ds, ss};
  }
  std::sort(dims, dims + n, [](const Dim& a, const Dim& c) { return a.ds > c.ds; });
  // merge (outer, inner) when the outer stride equals inner stride * inner extent on both sides
  int m = 0;
  for (int i = 0; i < n; ++i) {
    if (m > 0 && dims[m - 1].ds == dims[i].ds * dims[i].ext && dims[m - 1].ss == dims[i].ss * dims[i].ext) {
      dims[m - 1].ext *= dims[i].ext;
      dims[m - 1].ds = dims[i].ds;
      dims[m - 1].ss = dims[i].ss;
    } else {
      dims[m++] = dims[i];
    }
  }
  n = m;
  if (n == 0) { dims[0] = {1, 1, 1}; n = 1; }
  int eb = b->elem_bytes;
  int unit = eb;
  Dim& in = dims[n - 1];
  if (in.ds == 1 && in.ss == 1) {
    for (int cand = 16; cand > eb; cand >>= 1) {
      int f = cand / eb;
      bool ok = (in.ext % f) == 0;
      ok = ok && (((uintptr_t)dstp + (uintptr_t)(dbase * eb)) % cand == 0);
      if (srcp) ok = ok && (((uintptr_t)srcp + (uintptr_t)(sbase * eb)) % cand == 0);
      for (int i = 0; i < n - 1 && ok; ++i) ok = (dims[i].ds % f == 0) && (dims[i].ss % f == 0);
      if (ok) { unit = cand; break; }
    }
  }
  int f = unit / eb;
  if (f > 1) {
    in.ext /= f;
    for (int i = 0; i < n - 1; ++i) { dims[i].ds /= f; dims[i].ss /= f; }
  }
  *dst_base = dbase;  // in elements
  *src_base = sbase;
  // plane = last two dims; outer = the rest (<= 4)
  memset(pd, 0, sizeof(*pd));
  pd->e1 = (uint32_t)dims[n - 1].ext; pd->ds1 = dims[n - 1].ds; pd->ss1 = dims[n - 1].ss;
  if (n >= 2) { pd->e2 = (uint32_t)dims[n - 2].ext; pd->ds2 = dims[n - 2].ds; pd->ss2 = dims[n - 2].ss; }
  else { pd->e2 = 1; pd->ds2 = 0; pd->ss2 = 0; }
  // keep the plane below 2^31 units: if too large, split e1 (only when it is a single huge run)
  pd->n_outer = n > 2 ? n - 2 : 0;
  uint64_t outer_total = 1;
  for (int i = 0; i < pd->n_outer; ++i) {
    pd->oe[i] = (uint32_t)dims[i].ext; pd->ods[i] = dims[i].ds; pd->oss[i] = dims[i].ss;
    outer_total *= (uint64_t)dims[i].ext;
  }
  uint64_t plane = (uint64_t)pd->e2 * pd->e1;
  if (plane >= (1ull << 31)) {
    if (pd->n_outer < 4 && pd->e2 > 1) {  // move e2 out to the outer dims
      int k = pd->n_outer++;
      pd->oe[k] = pd->e2; pd->ods[k] = pd->ds2; pd->oss[k] = pd->ss2;
      outer_total *= pd->e2;
      pd->e2 = 1; pd->ds2 = 0; pd->ss2 = 0;
      plane = pd->e1;
    }
    if (plane >= (1ull << 31)) {
      // single run: fold into (e2, e1) with e1 = 2^20 units when divisible
      uint32_t blk = 1u << 20;
      if (pd->e2 == 1 && pd->e1 % blk == 0) {
        pd->e2 = pd->e1 / blk; pd->ds2 = pd->ds1 * blk; pd->ss2 = pd->ss1 * blk; pd->e1 = blk;
      }
      plane = (uint64_t)pd->e2 * pd->e1;
      if (plane >= (1ull << 32)) return -1;
    }
  }
  pd->chunks = (uint32_t)((plane + kChunk - 1) / kChunk);
  pd->items = outer_total * pd->chunks;
  return unit;
}

template <int MODE>
static int launch(void* dst, const void* src, const iit_box_t* box, bool compact, void* stream, const char* name) {
  IIT_REQUIRE(box && box->rank >= 0 && box->rank <= IIT_MAX_RANK, "%s: rank must be in [0,%d]", name, IIT_MAX_RANK);
  IIT_REQUIRE(box->elem_bytes == 2 || box->elem_bytes == 4 || box->elem_bytes == 8 || box->elem_bytes == 1,
              "%s: elem_bytes must be 1, 2, 4 or 8", name);
  for (int i = 0; i < box->rank; ++i)
    if (box->extent[i] <= 0) return 0;  // empty box: nothing to do (pointers may be null)
  IIT_REQUIRE(dst != nullptr && (MODE == MODE_ZERO || src != nullptr), "%s: null pointer", name);
  PatchDesc pd;
  int64_t db, sb;
  bool empty;
  int unit = plan(box, compact, dst, src, &pd, &db, &sb, &empty);
  if (empty) return 0;
  IIT_REQUIRE(unit > 0, "%s: box too large for one launch", name);
  char* d = (char*)dst + db * box->elem_bytes;
  const char* s = src ? (const char*)src + sb * box->elem_bytes : nullptr;
  int grid = grid_for((int64_t)std::min<uint64_t>(pd.items, (uint64_t)1 << 40), 8);
  cudaStream_t st = as_stream(stream);
  switch (unit) {
    case 16: patch_kernel<uint4, MODE><<<grid, 256, 0, st>>>((uint4*)d, (const uint4*)s, pd); break;
    case 8: patch_kernel<uint2, MODE><<<grid, 256, 0, st>>>((uint2*)d, (const uint2*)s, pd); break;
    case 4: patch_kernel<uint32_t, MODE><<<grid, 256, 0, st>>>((uint32_t*)d, (const uint32_t*)s, pd); break;
    case 2: patch_kernel<uint16_t, MODE><<<grid, 256, 0, st>>>((uint16_t*)d, (const uint16_t*)s, pd); break;
    default: patch_kernel<uint8_t, MODE><<<grid, 256, 0, st>>>((uint8_t*)d, (const uint8_t*)s, pd); break;
  }
  IIT_LAUNCH_CHECK(name);
  return 0;
}

}  // namespace iit

extern "C" int iit_patch_fwd(void* dst, const void* src, const iit_box_t* box, void* stream) {
  return iit::launch<iit::MODE_COPY>(dst, src, box, false, stream, "iit_patch_fwd");
}
extern "C" int iit_patch_bwd(void* grad, const iit_box_t* box, void* stream) {
  return iit::launch<iit::MODE_ZERO>(grad, nullptr, box, false, stream, "iit_patch_bwd");
}
extern "C" int iit_capture_box(void* dst_compact, const void* src, const iit_box_t* box, void* stream) {
  return iit::launch<iit::MODE_COPY>(dst_compact, src, box, true, stream, "iit_capture_box");
}
