// K4 — batched prediction: every row walks the flattened node array.
//
// Replaces the reference's per-row recursive Python walk driven by
// np.apply_along_axis and the arg-max over leaf counts
// (mpitree/tree/decision_tree.py:208-227, :248).  Nodes are packed into 16-byte
// records {feature, threshold, left child, leaf label}; children are numbered
// consecutively (right = left + 1), so one shared-memory load per level decides the
// step.  Trees up to kSmemNodes nodes are staged in shared memory, larger ones are
// read through L1/L2.
#include <stdio.h>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kSmemNodes = 8192;  // 128 KB of records

template <typename T>
struct Rec {
  int feature;
  int left;
  int label;
  int pad;
};

template <typename T, typename TH, bool kSmem>
__global__ void __launch_bounds__(kThreads) k_predict(const T* __restrict__ X, int64_t n, int64_t ldx,
                                                      const int4* __restrict__ nodes,
                                                      const TH* __restrict__ thr, int n_nodes,
                                                      int* __restrict__ out_leaf, int* __restrict__ out_label) {
  extern __shared__ __align__(16) uint8_t smem[];
  int4* s_nodes = reinterpret_cast<int4*>(smem);
  TH* s_thr = reinterpret_cast<TH*>(smem + (size_t)(kSmem ? n_nodes : 0) * sizeof(int4));
  if (kSmem) {
    for (int i = threadIdx.x; i < n_nodes; i += kThreads) {
      s_nodes[i] = nodes[i];
      s_thr[i] = thr[i];
    }
    __syncthreads();
  }
  const int4* N = kSmem ? s_nodes : nodes;
  const TH* TT = kSmem ? s_thr : thr;
  for (int64_t row = (int64_t)blockIdx.x * kThreads + threadIdx.x; row < n; row += (int64_t)gridDim.x * kThreads) {
    const T* x = X + row * ldx;
    int node = 0;
    int4 r = N[0];
    while (r.x >= 0) {
      const TH v = (TH)x[r.x];
      node = r.y + (v <= TT[node] ? 0 : 1);
      r = N[node];
    }
    if (out_leaf) out_leaf[row] = node;
    if (out_label) out_label[row] = r.z;
  }
}

// Staged variant: each warp copies 32 whole rows into shared memory with coalesced 128-bit
// loads (every input byte is read exactly once at full-line efficiency), then each lane walks
// the tree on its own row.  Rows are padded to an odd number of words so that lanes reading the
// same feature hit different banks.  Needs contiguous rows whose size is a multiple of 16 bytes.
template <typename T>
__device__ __forceinline__ T staged_elem(const uint32_t* row, int f);
template <>
__device__ __forceinline__ float staged_elem<float>(const uint32_t* row, int f) { return __uint_as_float(row[f]); }
template <>
__device__ __forceinline__ uint8_t staged_elem<uint8_t>(const uint32_t* row, int f) {
  return (uint8_t)((row[f >> 2] >> ((f & 3) * 8)) & 0xffu);
}
template <>
__device__ __forceinline__ double staged_elem<double>(const uint32_t* row, int f) {
  return __hiloint2double((int)row[2 * f + 1], (int)row[2 * f]);
}

constexpr int kStagedThreads = 1024;  // the walk is latency bound: many warps per SM

template <typename T, typename TH>
__global__ void __launch_bounds__(kStagedThreads) k_predict_staged(const T* __restrict__ X, int64_t n, int F,
                                                             const int4* __restrict__ nodes,
                                                             const TH* __restrict__ thr, int n_nodes, int stage_nodes,
                                                             int* __restrict__ out_leaf, int* __restrict__ out_label) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int rowbytes = F * (int)sizeof(T);
  const int cpr = rowbytes / 16;             // 16-byte chunks per row
  const int stride = (rowbytes / 4) | 1;     // words per staged row (odd)
  // node records staged in shared memory carry the threshold in .w when it is a 4-byte type
  constexpr bool kPacked = sizeof(TH) == 4;
  int4* s_nodes = reinterpret_cast<int4*>(smem);
  const size_t node_bytes = stage_nodes ? (size_t)n_nodes * (sizeof(int4) + (kPacked ? 0 : sizeof(TH))) : 0;
  TH* s_thr = reinterpret_cast<TH*>(smem + (size_t)(stage_nodes ? n_nodes : 0) * sizeof(int4));
  uint32_t* s_rows = reinterpret_cast<uint32_t*>(smem + ((node_bytes + 15) & ~(size_t)15));
  if (stage_nodes) {
    for (int i = threadIdx.x; i < n_nodes; i += kStagedThreads) {
      int4 r = nodes[i];
      if (kPacked) {
        const TH t = thr[i];
        r.w = *reinterpret_cast<const int*>(&t);
      } else {
        s_thr[i] = thr[i];
      }
      s_nodes[i] = r;
    }
    __syncthreads();
  }
  const int4* N = stage_nodes ? s_nodes : nodes;
  const TH* TT = stage_nodes ? s_thr : thr;
  const bool packed = kPacked && stage_nodes;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t* mine = s_rows + (size_t)w * 32 * stride;
  const int64_t tiles = (n + 31) / 32;
  const int64_t wstride = (int64_t)gridDim.x * (kStagedThreads / 32);
  for (int64_t tile = (int64_t)blockIdx.x * (kStagedThreads / 32) + w; tile < tiles; tile += wstride) {
    const int64_t row0 = tile * 32;
    const int rows = (int)min((int64_t)32, n - row0);
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(X) + row0 * rowbytes);
    const int chunks = rows * cpr;
    for (int q = lane; q < chunks; q += 32) {
      const uint4 v = mtb::ld_nc_u4(src + q);
      const int r = q / cpr, c = q - r * cpr;
      uint32_t* d = mine + r * stride + c * 4;
      d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    __syncwarp();
    if (lane < rows) {
      const uint32_t* x = mine + lane * stride;
      int node = 0;
      int4 r = N[0];
      while (r.x >= 0) {
        const TH v = (TH)staged_elem<T>(x, r.x);
        TH t;
        if (packed) t = *reinterpret_cast<const TH*>(&r.w); else t = TT[node];
        node = r.y + (v <= t ? 0 : 1);
        r = N[node];
      }
      if (out_leaf) out_leaf[row0 + lane] = node;
      if (out_label) out_label[row0 + lane] = r.z;
    }
    __syncwarp();
  }
}

// Compact variant for one-byte codes: a node is ONE 32-bit word
//   [31:24] feature (255 = leaf)   [23:16] threshold code   [15:0] left child id (leaf: unused)
// so every level costs one 4-byte shared-memory load per lane instead of a 16-byte one (random
// 16-byte loads take at least four wavefronts per warp).  Needs F <= 255 and <= 65,536 nodes;
// labels are looked up once at the end.
__global__ void __launch_bounds__(kStagedThreads) k_predict_u8_compact(const uint8_t* __restrict__ X, int64_t n,
                                                                       int F, const int4* __restrict__ nodes,
                                                                       const int* __restrict__ thr, int n_nodes,
                                                                       int* __restrict__ out_leaf,
                                                                       int* __restrict__ out_label) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int cpr = F / 16;
  const int stride = (F / 4) | 1;
  uint32_t* s_rec = reinterpret_cast<uint32_t*>(smem);
  uint32_t* s_rows = s_rec + ((n_nodes + 3) & ~3);
  for (int i = threadIdx.x; i < n_nodes; i += kStagedThreads) {
    const int4 r = nodes[i];
    const uint32_t f = r.x < 0 ? 255u : (uint32_t)r.x;
    // a negative threshold sends every code right: encode it as "always <=" into the right child
    const bool all_right = r.x >= 0 && thr[i] < 0;
    const uint32_t t = (r.x < 0 || all_right) ? 255u : (uint32_t)min(thr[i], 255);
    s_rec[i] = (f << 24) | (t << 16) | ((uint32_t)(max(r.y, 0) + (all_right ? 1 : 0)) & 0xffffu);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t* mine = s_rows + (size_t)w * 32 * stride;
  const int64_t tiles = (n + 31) / 32;
  const int64_t wstride = (int64_t)gridDim.x * (kStagedThreads / 32);
  for (int64_t tile = (int64_t)blockIdx.x * (kStagedThreads / 32) + w; tile < tiles; tile += wstride) {
    const int64_t row0 = tile * 32;
    const int rows = (int)min((int64_t)32, n - row0);
    const uint4* src = reinterpret_cast<const uint4*>(X + row0 * F);
    const int chunks = rows * cpr;
    for (int q = lane; q < chunks; q += 32) {
      const uint4 v = mtb::ld_nc_u4(src + q);
      const int r = q / cpr, c = q - r * cpr;
      uint32_t* d = mine + r * stride + c * 4;
      d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    __syncwarp();
    if (lane < rows) {
      const uint32_t* x = mine + lane * stride;
      int node = 0;
      uint32_t rec = s_rec[0];
      while ((rec >> 24) != 255u) {
        const uint32_t f = rec >> 24;
        const uint32_t v = (x[f >> 2] >> ((f & 3) * 8)) & 0xffu;
        node = (int)(rec & 0xffffu) + (v <= ((rec >> 16) & 0xffu) ? 0 : 1);
        rec = s_rec[node];
      }
      if (out_leaf) out_leaf[row0 + lane] = node;
      if (out_label) out_label[row0 + lane] = nodes[node].z;
    }
    __syncwarp();
  }
}

__global__ void k_pack_nodes(const int* feature, const int* left, const int* label, int n, int4* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = make_int4(feature[i], left[i], label ? label[i] : 0, 0);
}

template <typename T, typename TH>
int launch(const void* X, int64_t n, int64_t F, int64_t ldx, const int4* nodes, const void* thr, int n_nodes,
           int* out_leaf, int* out_label, cudaStream_t stream) {
  int64_t blocks = (n + kThreads - 1) / kThreads;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const int64_t rowbytes = F * (int64_t)sizeof(T);
  if constexpr (sizeof(T) == 1) {
    if (ldx == F && F % 16 == 0 && F <= 240 && n_nodes <= 65536 && ((uintptr_t)X % 16) == 0) {
      const int stride = (int)(F / 4) | 1;
      const int smem = ((n_nodes + 3) & ~3) * 4 + (kStagedThreads / 32) * 32 * stride * 4 + 16;
      if (smem <= 220 * 1024) {
        MTB_CHECK(cudaFuncSetAttribute(k_predict_u8_compact, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       smem > 48 * 1024 ? smem : 48 * 1024));
        int64_t sblocks = ((n + 31) / 32 + (kStagedThreads / 32) - 1) / (kStagedThreads / 32);
        if (sblocks > 148 * 2) sblocks = 148 * 2;
        k_predict_u8_compact<<<(unsigned)sblocks, kStagedThreads, smem, stream>>>(
            (const uint8_t*)X, n, (int)F, nodes, (const int*)thr, n_nodes, out_leaf, out_label);
        MTB_CHECK(cudaGetLastError());
        return 0;
      }
    }
  }
  if (ldx == F && rowbytes % 16 == 0 && rowbytes <= 1024 && ((uintptr_t)X % 16) == 0) {
    const int stride = (int)(rowbytes / 4) | 1;
    const int rows_smem = (kStagedThreads / 32) * 32 * stride * 4;
    const int per_node = (int)sizeof(int4) + (sizeof(TH) == 4 ? 0 : (int)sizeof(TH));
    const int node_smem = (n_nodes * per_node + 15) & ~15;
    const int stage_nodes = (node_smem + rows_smem <= 220 * 1024) ? 1 : 0;
    const int smem = (stage_nodes ? node_smem : 0) + rows_smem + 16;
    if (smem <= 220 * 1024) {
      auto kern = k_predict_staged<T, TH>;
      MTB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem > 48 * 1024 ? smem : 48 * 1024));
      int64_t sblocks = ((n + 31) / 32 + (kStagedThreads / 32) - 1) / (kStagedThreads / 32);
      if (sblocks > 148 * 2) sblocks = 148 * 2;
      kern<<<(unsigned)sblocks, kStagedThreads, smem, stream>>>((const T*)X, n, (int)F, nodes, (const TH*)thr, n_nodes,
                                                                stage_nodes, out_leaf, out_label);
      MTB_CHECK(cudaGetLastError());
      return 0;
    }
  }
  if (n_nodes <= kSmemNodes) {
    const int smem = n_nodes * (int)(sizeof(int4) + sizeof(TH));
    auto kern = k_predict<T, TH, true>;
    MTB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem > 48 * 1024 ? smem : 48 * 1024));
    kern<<<(unsigned)blocks, kThreads, smem, stream>>>((const T*)X, n, ldx, nodes, (const TH*)thr, n_nodes, out_leaf,
                                                       out_label);
  } else {
    k_predict<T, TH, false><<<(unsigned)blocks, kThreads, 0, stream>>>((const T*)X, n, ldx, nodes, (const TH*)thr,
                                                                       n_nodes, out_leaf, out_label);
  }
  MTB_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace

// packs {feature, left, label} into 16-byte node records (scratch: int4[n_nodes])
extern "C" int mtb_pack_nodes(const int32_t* feature, const int32_t* left, const int32_t* leaf_label, int64_t n_nodes,
                              void* packed_out, void* stream) {
  k_pack_nodes<<<(unsigned)((n_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>(feature, left, leaf_label,
                                                                                    (int)n_nodes, (int4*)packed_out);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_predict(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx, const void* packed_nodes,
                           const void* threshold, int64_t n_nodes, int32_t* out_leaf, int32_t* out_label,
                           void* stream) {
  if (n <= 0) return 0;
  const int4* nodes = (const int4*)packed_nodes;
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case 0: return launch<float, float>(X, n, F, ldx, nodes, threshold, (int)n_nodes, out_leaf, out_label, s);
    case 1: return launch<uint8_t, int>(X, n, F, ldx, nodes, threshold, (int)n_nodes, out_leaf, out_label, s);
    case 2: return launch<double, double>(X, n, F, ldx, nodes, threshold, (int)n_nodes, out_leaf, out_label, s);
    default:
      mtb_set_error("predict", "dtype must be 0 (f32), 1 (u8) or 2 (f64)", __FILE__, __LINE__);
      return -2;
  }
}

// ---------------------------------------------------------------------------------------
// mtb_predict_noalloc — prediction for callers that cannot obtain device memory.
//
// Who needs this.  mtb_predict() wants a 16-byte-per-node scratch buffer that the caller
// has to allocate.  Some callers have no way to do that: a serving process that carved up
// the whole GPU at start-up and forbids any later cudaMalloc (an allocation may fail because
// the pool already owns every byte, and cudaMalloc/cudaFree synchronise the device, which
// stalls the other streams of a latency-bound server); or a C plugin invoked from inside a
// host application's operator, where every device byte belongs to that application's
// allocator and this library is only handed input and output pointers.
//
// What it does instead.  The node-record scratch lives in the fixed array below, which is
// part of the library image: every GPU that loads the library holds these
// MTB_NOALLOC_MAX_NODES * 16 bytes (1 MiB) until the process exits.  Trees with more nodes
// are rejected with an error; nothing is truncated.
//
// Rules for callers.
//  * Inputs (X, feature, threshold, left, leaf_label) and outputs (out_leaf, out_label) are
//    still the caller's buffers; the device that owns them must be current.
//  * The first call on each device looks up the array's device address (one
//    cudaGetSymbolAddress, no allocation).  Make that first call before starting any stream
//    capture on that device.
//  * All calls on one device share the one array, and it keeps whatever the previous call
//    packed into it.  Each call rewrites records [0, n_nodes) before reading them and never
//    reads past n_nodes, so nothing needs clearing between calls.  Two calls whose GPU work
//    overlaps (different streams, or different host threads) race on the array: serialise
//    them, e.g. by using one stream.
//  * No cudaMalloc, cudaFree or other allocating routine is reached from this entry point.
// ---------------------------------------------------------------------------------------
#include <mutex>

#define MTB_NOALLOC_MAX_NODES 65536
#define MTB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_nodes[(size_t)MTB_NOALLOC_MAX_NODES * 16];

static std::mutex g_noalloc_mutex;
static void* g_noalloc_addr[MTB_NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of g_noalloc_nodes on the current device.  In host code the symbol's name is
// only a handle, never a pointer a kernel may use, so the address is resolved once per device.
static int noalloc_scratch(void** out) {
  int dev = -1;
  MTB_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MTB_NOALLOC_MAX_DEVICES) {
    mtb_set_error("mtb_predict_noalloc", "device ordinal out of range (0..63)", __FILE__, __LINE__);
    return -2;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_addr[dev] == nullptr) {
    void* ptr = nullptr;
    MTB_CHECK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_nodes));
    g_noalloc_addr[dev] = ptr;
  }
  *out = g_noalloc_addr[dev];
  return 0;
}

extern "C" int mtb_predict_noalloc(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx,
                                   const int32_t* feature, const void* threshold, const int32_t* left,
                                   const int32_t* leaf_label, int64_t n_nodes, int32_t* out_leaf,
                                   int32_t* out_label, void* stream) {
  if (n_nodes < 1 || n_nodes > MTB_NOALLOC_MAX_NODES) {
    char why[160];
    snprintf(why, sizeof(why), "tree has %lld nodes; this entry point holds at most %d (use mtb_predict with your own scratch)",
             (long long)n_nodes, MTB_NOALLOC_MAX_NODES);
    mtb_set_error("mtb_predict_noalloc", why, __FILE__, __LINE__);
    return -3;
  }
  if (n <= 0) return 0;
  void* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc != 0) return rc;
  k_pack_nodes<<<(unsigned)((n_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>(feature, left, leaf_label,
                                                                                    (int)n_nodes, (int4*)scratch);
  MTB_CHECK(cudaGetLastError());
  return mtb_predict(X, dtype, n, F, ldx, scratch, threshold, n_nodes, out_leaf, out_label, stream);
}
