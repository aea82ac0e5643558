// K1 — encode raw feature rows into tiled one-byte bin codes.
//
// The reference discovers candidate thresholds with np.unique at every node
// (mpitree/tree/decision_tree.py:73).  Here each feature gets an ascending edge
// table once per fit and every value is replaced by the index of the first edge
// >= value, so `x <= threshold` on raw values equals `code <= bin` on codes.
// Output layout: codes[g][row][k] for feature f = 16*g + k (one 128-bit word per
// row per feature group), the layout the histogram kernel streams with TMA.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

template <typename T>
__global__ void __launch_bounds__(kThreads) k_encode_float(const T* __restrict__ X, int64_t n, int F, int64_t ldx,
                                                           const double* __restrict__ edges,
                                                           const int* __restrict__ n_bins, uint4* __restrict__ codes,
                                                           int64_t n_cap, int64_t row0,
                                                           unsigned long long* mismatch) {
  __shared__ double s_edges[16][MTB_MAX_BINS];
  __shared__ int s_nb[16];
  const int g = blockIdx.y;
  for (int i = threadIdx.x; i < 16 * MTB_MAX_BINS; i += kThreads) {
    const int k = i / MTB_MAX_BINS, b = i % MTB_MAX_BINS;
    const int f = g * 16 + k;
    s_edges[k][b] = f < F ? edges[(int64_t)f * MTB_MAX_BINS + b] : INFINITY;
  }
  if (threadIdx.x < 16) {
    const int f = g * 16 + threadIdx.x;
    s_nb[threadIdx.x] = f < F ? n_bins[f] : 1;
  }
  __syncthreads();
  unsigned int bad = 0;
  for (int64_t row = (int64_t)blockIdx.x * kThreads + threadIdx.x; row < n; row += (int64_t)gridDim.x * kThreads) {
    uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int f = g * 16 + k;
      if (f >= F) break;
      const double x = (double)X[row * ldx + f];
      // first edge >= x (edges beyond n_bins are +inf)
      int lo = 0, hi = s_nb[k] - 1;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_edges[k][mid] >= x) hi = mid; else lo = mid + 1;
      }
      bad += s_edges[k][lo] != x;
      w[k >> 2] |= (uint32_t)lo << ((k & 3) * 8);
    }
    codes[(int64_t)g * n_cap + row0 + row] = make_uint4(w[0], w[1], w[2], w[3]);
  }
  if (mismatch) {
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatch, (unsigned long long)bad);
  }
}

// Variant for float32 rows whose 4-feature quads are 16-byte aligned (opt-in, see
// mtb_encode_f32_quads).  One lane owns one (row, quad): it loads the quad with a single 128-bit
// streaming load, runs four fixed-length (8-step) branch-free lower-bound searches in a float32
// copy of the edge table and stores one 32-bit word of codes.  Four neighbouring lanes cover the
// 64 input bytes and the 16 output bytes of one row of the feature group, so a warp reads 2 KB
// and writes 128 B of contiguous memory per instruction.  The edge table is converted with
// round-toward-minus-infinity: for a float32 x and a float64 edge e, `e >= x` holds exactly when
// `rd(e) >= x`, so the codes equal k_encode_float's.  Rows of the table are padded to 257 words so
// that the same search step of different features falls into different banks.
constexpr int kEdgeStride = MTB_MAX_BINS + 1;
static_assert(MTB_MAX_BINS <= 256, "k_encode_f32_slots packs two bin codes into a 16-bit table entry");

__global__ void __launch_bounds__(kThreads) k_encode_f32_quads(const float4* __restrict__ X, int64_t n, int F,
                                                               int64_t ldx4, const double* __restrict__ edges,
                                                               const int* __restrict__ n_bins,
                                                               uint32_t* __restrict__ codes, int64_t n_cap,
                                                               int64_t row0) {
  __shared__ float s_edges[16 * kEdgeStride];
  __shared__ int s_last[16];
  const int g = blockIdx.y;
  for (int i = threadIdx.x; i < 16 * MTB_MAX_BINS; i += kThreads) {
    const int k = i / MTB_MAX_BINS, b = i % MTB_MAX_BINS;
    const int f = g * 16 + k;
    s_edges[k * kEdgeStride + b] = f < F ? __double2float_rd(edges[(int64_t)f * MTB_MAX_BINS + b]) : INFINITY;
  }
  if (threadIdx.x < 16) {
    const int f = g * 16 + threadIdx.x;
    s_last[threadIdx.x] = (f < F ? n_bins[f] : 1) - 1;
  }
  __syncthreads();
  const int quads = min(4, (F - g * 16 + 3) / 4);  // quads of this group that exist in X
  const int64_t total = n * 4;
  for (int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x; c < total; c += (int64_t)gridDim.x * kThreads) {
    const int64_t row = c >> 2;
    const int q = (int)(c & 3);
    uint32_t word = 0;
    if (q < quads) {
      const uint4 raw = mtb::ld_nc_u4(X + row * ldx4 + g * 4 + q);
      const float x[4] = {__uint_as_float(raw.x), __uint_as_float(raw.y), __uint_as_float(raw.z),
                          __uint_as_float(raw.w)};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = q * 4 + j;
        const float* e = s_edges + k * kEdgeStride;
        int lo = 0;  // number of edges that are not >= x (NaN compares false: it ends in the last bin)
#pragma unroll
        for (int step = MTB_MAX_BINS / 2; step > 0; step >>= 1) lo += !(e[lo + step - 1] >= x[j]) ? step : 0;
        lo = min(lo, s_last[k]);
        word |= (uint32_t)lo << (j * 8);
      }
    }
    codes[(((int64_t)g * n_cap + row0 + row) << 2) + q] = word;
  }
}

// Same mapping as k_encode_f32_quads (one lane per row x 4 features, one 128-bit load, one 32-bit store)
// with a cheaper search: the value range of every feature is cut into 512 equal slots and a table
// gives, per slot, the index range of the edges that can be the answer, so a value costs one
// subtract-multiply-truncate, one table load (both ends of the range in one word; the feature's
// constants sit in registers: a lane always serves the same quad) and a probe of typically one or two
// edges instead of eight dependent steps of binary search (26.7 ms at 1.2 TB/s for 100M x 64 with the
// latter).
// Exactness: slot() is monotone in x and is applied to the edges with the very same float32
// operations, so an edge whose slot is below x's slot is below x, and one whose slot is above is
// >= x: the first edge >= x lies in [lut[s], lut[s + 1]]; ranges longer than a few edges (skewed data)
// are narrowed by bisection first.
constexpr int kSlots = 512;

__device__ __forceinline__ int slot_of(float x, float mn, float inv) {
  float t = __fmul_rn(__fsub_rn(x, mn), inv);
  t = fminf(fmaxf(t, 0.0f), (float)(kSlots - 1));  // (NaN -> 0; NaN values are handled by the caller)
  return (int)t;
}

__global__ void __launch_bounds__(kThreads) k_encode_f32_slots(const float4* __restrict__ X, int64_t n, int F,
                                                               int64_t ldx4, const double* __restrict__ edges,
                                                               const int* __restrict__ n_bins,
                                                               uint32_t* __restrict__ codes, int64_t n_cap,
                                                               int64_t row0) {
  __shared__ float s_edges[16 * kEdgeStride];
  __shared__ uint16_t s_lut[16][kSlots + 2];  // 16-bit entries: 33 KB per block in all, six blocks per SM (the kernel
                                              // lives on loads in flight: 1,024 slots / 82 KB ran at half the speed)
  __shared__ float s_mn[16], s_inv[16];
  __shared__ int s_last[16];
  const int g = blockIdx.y;
  for (int i = threadIdx.x; i < 16 * MTB_MAX_BINS; i += kThreads) {
    const int k = i / MTB_MAX_BINS, b = i % MTB_MAX_BINS;
    const int f = g * 16 + k;
    s_edges[k * kEdgeStride + b] = f < F ? __double2float_rd(edges[(int64_t)f * MTB_MAX_BINS + b]) : INFINITY;
  }
  if (threadIdx.x < 16) {
    const int f = g * 16 + threadIdx.x;
    s_last[threadIdx.x] = (f < F ? n_bins[f] : 1) - 1;
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    const float* e = s_edges + threadIdx.x * kEdgeStride;
    const float lo = e[0], hi = e[s_last[threadIdx.x]];
    const float w = (hi - lo) / (float)kSlots;
    s_mn[threadIdx.x] = lo;
    s_inv[threadIdx.x] = (w > 0.0f && w < INFINITY) ? 1.0f / w : 0.0f;
  }
  __syncthreads();
  // below(k, s) = number of edges whose slot is below s (the edges' slots ascend: bisection per entry)
  auto below = [&](int k, int sl) {
    const float* e = s_edges + k * kEdgeStride;
    const float mn = s_mn[k], inv = s_inv[k];
    int lo = 0, hi = s_last[k] + 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (slot_of(e[mid], mn, inv) >= sl) hi = mid; else lo = mid + 1;
    }
    return (uint32_t)lo;
  };
  // one 16-bit entry per slot: first candidate code in the low byte, last one (capped at the last bin) in the
  // high byte, so a value costs one table load before its probes (codes are < 256 = MTB_MAX_BINS)
  for (int i = threadIdx.x; i < 16 * (kSlots + 1); i += kThreads) {
    const int k = i / (kSlots + 1), sl = i % (kSlots + 1);
    const uint32_t lo = below(k, sl);
    const uint32_t hi = min(sl < kSlots ? below(k, sl + 1) : (uint32_t)s_last[k] + 1u, (uint32_t)s_last[k]);
    s_lut[k][sl] = (uint16_t)(lo | (hi << 8));
  }
  __syncthreads();
  const int quads = min(4, (F - g * 16 + 3) / 4);  // quads of this group that exist in X
  const int64_t total = n * 4;
  // the grid stride is a multiple of four, so a lane serves the same quad (the same four features) throughout
  const int q = (int)(((int64_t)blockIdx.x * kThreads + threadIdx.x) & 3);
  float mn[4], inv[4];
  int last[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    mn[j] = s_mn[q * 4 + j];
    inv[j] = s_inv[q * 4 + j];
    last[j] = s_last[q * 4 + j];
  }
  auto encode4 = [&](const uint4& raw) -> uint32_t {
    const float x[4] = {__uint_as_float(raw.x), __uint_as_float(raw.y), __uint_as_float(raw.z), __uint_as_float(raw.w)};
    uint32_t word = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = q * 4 + j;
      const float* e = s_edges + k * kEdgeStride;
      const int sl = slot_of(x[j], mn[j], inv[j]);
      const uint32_t ent = s_lut[k][sl];
      int lo = (int)(ent & 255u), hi = (int)(ent >> 8);
      while (hi - lo > 4) {
        const int mid = (lo + hi) >> 1;
        if (e[mid] >= x[j]) hi = mid; else lo = mid + 1;
      }
      while (lo < hi && !(e[lo] >= x[j])) ++lo;
      if (x[j] != x[j]) lo = last[j];  // NaN compares false everywhere: the last bin, like the other encoders
      word |= (uint32_t)lo << (j * 8);
    }
    return word;
  };
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  // two rows per iteration: both 128-bit loads are in flight before the first search starts
  for (int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x; c < total; c += 2 * stride) {
    const int64_t c1 = c + stride;
    const int64_t row = c >> 2, row1 = c1 >> 2;
    uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
    if (q < quads) {
      r0 = mtb::ld_nc_u4(X + row * ldx4 + g * 4 + q);
      if (c1 < total) r1 = mtb::ld_nc_u4(X + row1 * ldx4 + g * 4 + q);
    }
    codes[(((int64_t)g * n_cap + row0 + row) << 2) + q] = q < quads ? encode4(r0) : 0u;
    if (c1 < total) codes[(((int64_t)g * n_cap + row0 + row1) << 2) + q] = q < quads ? encode4(r1) : 0u;
  }
}

__global__ void __launch_bounds__(kThreads) k_encode_u8(const uint8_t* __restrict__ X, int64_t n, int F, int64_t ldx,
                                                        const uint8_t* __restrict__ lut, uint4* __restrict__ codes,
                                                        int64_t n_cap, int64_t row0, int vec_ok) {
  __shared__ uint8_t s_lut[16][256];
  const int g = blockIdx.y;
  if (lut) {
    for (int i = threadIdx.x; i < 16 * 256; i += kThreads) {
      const int k = i >> 8, f = g * 16 + k;
      s_lut[k][i & 255] = f < F ? lut[(int64_t)f * 256 + (i & 255)] : 0;
    }
    __syncthreads();
  }
  const int kmax = min(16, F - g * 16);
  for (int64_t row = (int64_t)blockIdx.x * kThreads + threadIdx.x; row < n; row += (int64_t)gridDim.x * kThreads) {
    const uint8_t* src = X + row * ldx + g * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (vec_ok && kmax == 16) {
      v = mtb::ld_nc_u4(src);
    } else {
      uint32_t w[4] = {0, 0, 0, 0};
      for (int k = 0; k < kmax; ++k) w[k >> 2] |= (uint32_t)src[k] << ((k & 3) * 8);
      v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    if (lut) {
      uint32_t w[4] = {v.x, v.y, v.z, v.w};
      uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const uint32_t c = (w[k >> 2] >> ((k & 3) * 8)) & 0xffu;
        o[k >> 2] |= (uint32_t)(k < kmax ? s_lut[k][c] : 0) << ((k & 3) * 8);
      }
      v = make_uint4(o[0], o[1], o[2], o[3]);
    }
    codes[(int64_t)g * n_cap + row0 + row] = v;
  }
}

// identity codes, F a multiple of 16, 16-byte aligned rows: consecutive threads read consecutive
// 16-byte chunks of a row (fully coalesced) and write them to their feature-group planes
__global__ void __launch_bounds__(kThreads) k_retile_u8(const uint4* __restrict__ X, int64_t n, int G, int64_t ldx16,
                                                        uint4* __restrict__ codes, int64_t n_cap, int64_t row0) {
  const int64_t total = n * G;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int64_t row = i / G;
    const int g = (int)(i - row * G);
    codes[(int64_t)g * n_cap + row0 + row] = mtb::ld_nc_u4(X + row * ldx16 + g);
  }
}

}  // namespace

extern "C" int mtb_encode_float(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx, const double* edges,
                                const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0,
                                unsigned long long* mismatch, void* stream) {
  if (n <= 0) return 0;
  const int G = (int)((F + 15) / 16);
  int64_t bx = (n + kThreads - 1) / kThreads;
  if (bx > 148 * 16) bx = 148 * 16;
  dim3 grid((unsigned)bx, (unsigned)G);
  if (dtype == 0)
    k_encode_float<float><<<grid, kThreads, 0, (cudaStream_t)stream>>>((const float*)X, n, (int)F, ldx, edges, n_bins,
                                                                       (uint4*)codes, n_cap, row0, mismatch);
  else
    k_encode_float<double><<<grid, kThreads, 0, (cudaStream_t)stream>>>((const double*)X, n, (int)F, ldx, edges,
                                                                        n_bins, (uint4*)codes, n_cap, row0, mismatch);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_encode_f32_quads(const float* X, int64_t n, int64_t F, int64_t ldx, const double* edges,
                                    const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0,
                                    void* stream) {
  if (n <= 0) return 0;
  if (F % 4 != 0 || ldx % 4 != 0 || (uintptr_t)X % 16 != 0) {
    mtb_set_error("encode_f32_quads", "needs F % 4 == 0, ldx % 4 == 0 and 16-byte aligned rows", __FILE__, __LINE__);
    return -1;
  }
  const int G = (int)((F + 15) / 16);
  int64_t bx = (n * 4 + kThreads - 1) / kThreads;
  if (bx > 148 * 8) bx = 148 * 8;
  dim3 grid((unsigned)bx, (unsigned)G);
  k_encode_f32_quads<<<grid, kThreads, 0, (cudaStream_t)stream>>>((const float4*)X, n, (int)F, ldx / 4, edges, n_bins,
                                                                  (uint32_t*)codes, n_cap, row0);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_encode_f32_slots(const float* X, int64_t n, int64_t F, int64_t ldx, const double* edges,
                                    const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0,
                                    void* stream) {
  if (n <= 0) return 0;
  if (F % 4 != 0 || ldx % 4 != 0 || (uintptr_t)X % 16 != 0) {
    mtb_set_error("encode_f32_slots", "needs F % 4 == 0, ldx % 4 == 0 and 16-byte aligned rows", __FILE__, __LINE__);
    return -1;
  }
  const int G = (int)((F + 15) / 16);
  int64_t bx = (n * 4 + kThreads - 1) / kThreads;
  if (bx > 148 * 3) bx = 148 * 3;  // few, long-lived blocks: every block builds the slot tables of its feature group
  dim3 grid((unsigned)bx, (unsigned)G);
  k_encode_f32_slots<<<grid, kThreads, 0, (cudaStream_t)stream>>>((const float4*)X, n, (int)F, ldx / 4, edges, n_bins,
                                                                  (uint32_t*)codes, n_cap, row0);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_encode_u8(const uint8_t* X, int64_t n, int64_t F, int64_t ldx, const uint8_t* lut, uint8_t* codes,
                             int64_t n_cap, int64_t row0, void* stream) {
  if (n <= 0) return 0;
  const int G = (int)((F + 15) / 16);
  int64_t bx = (n + kThreads - 1) / kThreads;
  if (bx > 148 * 16) bx = 148 * 16;
  dim3 grid((unsigned)bx, (unsigned)G);
  const int vec_ok = (ldx % 16 == 0) && ((uintptr_t)X % 16 == 0);
  if (vec_ok && lut == nullptr && F % 16 == 0) {
    int64_t blocks = (n * G + kThreads - 1) / kThreads;
    if (blocks > 148 * 32) blocks = 148 * 32;
    k_retile_u8<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>((const uint4*)X, n, G, ldx / 16, (uint4*)codes,
                                                                        n_cap, row0);
    MTB_CHECK(cudaGetLastError());
    return 0;
  }
  k_encode_u8<<<grid, kThreads, 0, (cudaStream_t)stream>>>(X, n, (int)F, ldx, lut, (uint4*)codes, n_cap, row0, vec_ok);
  MTB_CHECK(cudaGetLastError());
  return 0;
}
