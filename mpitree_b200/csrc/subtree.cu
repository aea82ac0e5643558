// K6 (compute half) — finish whole subtrees of small nodes inside one CTA each.
//
// The reference's parallel strategy hands subtrees to halves of the communicator and lets
// each finish its part serially (mpitree/tree/decision_tree.py:430-466).  Here that idea is
// phase 2 of the scheduler: once a node's global row count is <= small_rows the level-wise
// machinery (one F x B x C slab per node, grid-wide phases) is the wrong tool; the node's
// rows are contiguous (after K3, or after the K6 gather onto the owner rank), so one CTA
// keeps a row-index permutation in shared memory and grows the complete subtree depth-first:
//
//   per node: for every feature tile  -> shared-memory histogram of the node's rows (128-bit
//             gathers, atomic increments), in-place prefix over bins, float64 cost with the
//             same device functions as the level-wise gain phase (impurity.cuh), first
//             arg-min over bins, first arg-max over features;
//             then partition the index range in shared memory and push the children.
//
// Output: flat node records per subtree (feature, bin, left child, depth, gain, counts),
// stitched into the trunk on the host with deterministic ids.
#include "common.cuh"
#include "impurity.cuh"

using namespace mtb;

typedef struct MtbSubtreeArgs {
  int64_t codes[2], labels[2];
  int64_t n_cap, F, G, B, C, FT, T;
  int64_t max_depth, min_samples_split, criterion, dot_mode;
  int64_t plogp, plogp_n, klogk, klogk_n;
  int64_t n_items;
  int64_t it_side, it_begin, it_end, it_depth, it_off;  // i32 [n_items]
  int64_t it_counts;                                     // u32 [n_items][C]
  int64_t o_feature, o_bin, o_left, o_depth;             // i32 [total_nodes]
  int64_t o_gain;                                        // f64
  int64_t o_counts;                                      // u32 [total_nodes][C]
  int64_t o_nn;                                          // i32 [n_items] nodes written per item
  int64_t stack;                                         // i32 [grid][T][3]
  int64_t counter;                                       // i32 work counter (zeroed by the caller)
} MtbSubtreeArgs;

namespace {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;

struct Best {
  double gain;
  int feature, bin;
};

__device__ __forceinline__ bool better(double g, int f, const Best& b) {
  return b.feature < 0 || g > b.gain || (g == b.gain && f < b.feature);
}

// Evaluate one feature whose [B][C] counts of the current node are in `h` (shared memory).
// The prefix over bins is done in place; returns the minimal cost and its first bin.
__device__ void eval_feature(const MtbSubtreeArgs& a, uint32_t* h, uint32_t m, double& cost_out, int& bin_out) {
  const int B = (int)a.B, C = (int)a.C, BPL = B / 32;
  const int lane = lane_id();
  uint32_t present = 0;
  for (int c = 0; c < C; ++c) {
    uint32_t carry = 0;
    for (int j = 0; j < BPL; ++j) {
      const int b = j * 32 + lane;
      const uint32_t x = h[b * C + c];
      if (x) present |= 1u << j;
      const uint32_t inc = warp_incl_scan(x);
      h[b * C + c] = inc + carry;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
  }
  __syncwarp();
  const uint32_t* last = h + (B - 1) * C;  // inclusive prefix at the last bin == class totals
  const double* plogp = ptr_of<const double>(a.plogp);
  const bool exact = plogp == nullptr || m <= (uint32_t)a.plogp_n;
  KlogK klogk;
  klogk.bind(ptr_of<const double>(a.klogk), ptr_of<const double>(a.klogk));
  const int crit = (int)a.criterion, dm = (int)a.dot_mode;
  double best = INFINITY;
  int best_bin = 0x7fffffff;
  for (int j = 0; j < BPL; ++j) {
    if (!(present >> j & 1u)) continue;
    const int b = j * 32 + lane;
    const uint32_t* L = h + b * C;
    uint32_t nl = 0;
    for (int c = 0; c < C; ++c) nl += L[c];
    const uint32_t nr = m - nl;
    if (nr == 0) continue;
    double cost;
    if (exact) {
      const double hl = impurity(C, [&](int c) { return L[c]; }, nl, crit, plogp, (uint32_t)a.plogp_n);
      const double hr = impurity(C, [&](int c) { return last[c] - L[c]; }, nr, crit, plogp, (uint32_t)a.plogp_n);
      cost = split_cost(nl, nr, hl, hr, dm);
    } else {
      cost = fast_split_cost(C, [&](int c) { return L[c]; }, [&](int c) { return last[c] - L[c]; }, nl, nr, crit, klogk);
    }
    if (cost < best || (cost == best && b < best_bin)) {
      best = cost;
      best_bin = b;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_xor_sync(0xffffffffu, best, o);
    const int ob = __shfl_xor_sync(0xffffffffu, best_bin, o);
    if (oc < best || (oc == best && ob < best_bin)) {
      best = oc;
      best_bin = ob;
    }
  }
  cost_out = best;
  bin_out = best_bin == 0x7fffffff ? -1 : best_bin;
}

template <int FT>
__global__ void __launch_bounds__(kThreads, 1) k_subtree(const __grid_constant__ MtbSubtreeArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int B = (int)a.B, C = (int)a.C, F = (int)a.F, G = (int)a.G, T = (int)a.T;
  const int tile_words = FT * B * C;
  uint32_t* tile = reinterpret_cast<uint32_t*>(smem);
  uint16_t* idx = reinterpret_cast<uint16_t*>(smem + (size_t)tile_words * 4);
  uint16_t* idx2 = idx + T;
  uint8_t* lab = reinterpret_cast<uint8_t*>(idx2 + T);
  __shared__ Best s_best[kWarps];
  __shared__ uint32_t s_cnt[MTB_MAX_CLASSES];
  __shared__ uint32_t s_cm[MTB_MAX_CLASSES];  // tiny nodes: lanes (rows) of each class
  __shared__ int s_ctl[8];  // 0: item, 1: rel, 2: lo, 3: hi, 4: curL, 5: curR, 6: sp, 7: nn

  const int tid = threadIdx.x, lane = lane_id(), w = warp_id();
  const int* it_side = ptr_of<const int>(a.it_side);
  const int* it_begin = ptr_of<const int>(a.it_begin);
  const int* it_end = ptr_of<const int>(a.it_end);
  const int* it_depth = ptr_of<const int>(a.it_depth);
  const int* it_off = ptr_of<const int>(a.it_off);
  const uint32_t* it_counts = ptr_of<const uint32_t>(a.it_counts);
  int* o_feature = ptr_of<int>(a.o_feature);
  int* o_bin = ptr_of<int>(a.o_bin);
  int* o_left = ptr_of<int>(a.o_left);
  int* o_depth = ptr_of<int>(a.o_depth);
  double* o_gain = ptr_of<double>(a.o_gain);
  uint32_t* o_counts = ptr_of<uint32_t>(a.o_counts);
  int* stack = ptr_of<int>(a.stack) + (int64_t)blockIdx.x * T * 3;
  const int S = 16 / FT;

  for (;;) {
    if (tid == 0) s_ctl[0] = atomicAdd(ptr_of<int>(a.counter), 1);
    __syncthreads();
    const int item = s_ctl[0];
    __syncthreads();
    if (item >= (int)a.n_items) break;
    const int side = it_side[item], begin = it_begin[item], n = it_end[item] - begin;
    const int base = it_off[item];
    const uint4* codes = reinterpret_cast<const uint4*>(ptr_of<const uint8_t>(a.codes[side]));
    const uint8_t* labels = ptr_of<const uint8_t>(a.labels[side]);
    for (int i = tid; i < n; i += kThreads) {
      idx[i] = (uint16_t)i;
      lab[i] = labels[begin + i];
    }
    if (tid == 0) {
      o_depth[base] = it_depth[item];
      o_feature[base] = -1; o_bin[base] = -1; o_left[base] = -1; o_gain[base] = NAN;
      stack[0] = 0; stack[1] = 0; stack[2] = n;
      s_ctl[6] = 1;  // stack entries
      s_ctl[7] = 1;  // nodes
    }
    for (int c = tid; c < C; c += kThreads) o_counts[(int64_t)base * C + c] = it_counts[(int64_t)item * C + c];
    __syncthreads();

    while (s_ctl[6] > 0) {
      __syncthreads();
      if (tid == 0) {
        const int sp = --s_ctl[6];
        s_ctl[1] = stack[sp * 3]; s_ctl[2] = stack[sp * 3 + 1]; s_ctl[3] = stack[sp * 3 + 2];
      }
      __syncthreads();
      const int rel = s_ctl[1], lo = s_ctl[2], hi = s_ctl[3];
      const int m = hi - lo;
      const int depth = o_depth[base + rel];
      const uint32_t* cnt = o_counts + (int64_t)(base + rel) * C;
      // stop rules (same as the level-wise finalize phase)
      int present_classes = 0;
      for (int c = 0; c < C; ++c) present_classes += cnt[c] != 0u;
      const bool terminal = present_classes <= 1 || (a.max_depth >= 0 && depth >= a.max_depth) ||
                            (int64_t)m < a.min_samples_split;
      if (terminal) continue;  // stays a leaf (record already initialised)

      const bool exact = a.plogp == 0 || (uint32_t)m <= (uint32_t)a.plogp_n;
      KlogK klogk;
      klogk.bind(ptr_of<const double>(a.klogk), ptr_of<const double>(a.klogk));
      const double hp = exact ? impurity(C, [&](int c) { return cnt[c]; }, (uint32_t)m, (int)a.criterion,
                                         ptr_of<const double>(a.plogp), (uint32_t)a.plogp_n)
                              : fast_impurity(C, [&](int c) { return cnt[c]; }, (uint32_t)m, (int)a.criterion, klogk);
      Best mine{0.0, -1, -1};
      if (m <= 32) {
        // ---- tiny node: no histogram at all.  Lane j holds row j; a warp evaluates one feature from
        // registers: the "goes left" set of candidate i is a ballot, class counts are popcounts
        // against per-class lane masks.  Same candidates, costs and tie-breaks as the tile path.
        const bool have = lane < m;
        const int r = have ? idx[lo + lane] : 0;
        const int l = have ? lab[r] : -1;
        if (w == 0)
          for (int c = 0; c < C; ++c) {
            const uint32_t cm = __ballot_sync(0xffffffffu, l == c);
            if (lane == 0) s_cm[c] = cm;
          }
        __syncthreads();
        const double* plogp = ptr_of<const double>(a.plogp);
        for (int f = w; f < F; f += kWarps) {
          int code = -1;
          if (have) code = reinterpret_cast<const uint8_t*>(codes + (int64_t)(f >> 4) * a.n_cap + begin + r)[f & 15];
          uint32_t left_set = 0;
          for (int i = 0; i < m; ++i) {
            const int t = __shfl_sync(0xffffffffu, code, i);
            const uint32_t bm = __ballot_sync(0xffffffffu, have && code <= t);
            if (lane == i) left_set = bm;
          }
          double cost = INFINITY;
          int bin = 0x7fffffff;
          const uint32_t nl = __popc(left_set), nr = (uint32_t)m - nl;
          if (have && nr != 0) {
            const double hl = impurity(C, [&](int c) { return (uint32_t)__popc(left_set & s_cm[c]); }, nl,
                                       (int)a.criterion, plogp, (uint32_t)a.plogp_n);
            const double hr = impurity(C, [&](int c) { return cnt[c] - (uint32_t)__popc(left_set & s_cm[c]); }, nr,
                                       (int)a.criterion, plogp, (uint32_t)a.plogp_n);
            cost = split_cost(nl, nr, hl, hr, (int)a.dot_mode);
            bin = code;
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double oc = __shfl_xor_sync(0xffffffffu, cost, o);
            const int ob = __shfl_xor_sync(0xffffffffu, bin, o);
            if (oc < cost || (oc == cost && ob < bin)) {
              cost = oc;
              bin = ob;
            }
          }
          if (bin != 0x7fffffff) {
            const double gain = __dsub_rn(hp, cost);
            if (better(gain, f, mine)) mine = Best{gain, f, bin};
          }
        }
        __syncthreads();
      } else {
      for (int g = 0; g < G; ++g) {
          for (int s = 0; s < S; ++s) {
            const int f0 = g * 16 + s * FT;
            if (f0 >= F) break;
            for (int i = tid; i < tile_words; i += kThreads) tile[i] = 0;
            __syncthreads();
            for (int i = lo + tid; i < hi; i += kThreads) {
              const int r = idx[i];
              const uint4 v = codes[(int64_t)g * a.n_cap + begin + r];
              const int l = lab[r];
#pragma unroll
              for (int k = 0; k < FT; ++k)
                if (f0 + k < F) atomicAdd(&tile[(k * B + (int)code_byte(v, s * FT + k)) * C + l], 1u);
            }
            __syncthreads();
            for (int k = w; k < FT && f0 + k < F; k += kWarps) {
              double cost;
              int bin;
              eval_feature(a, tile + k * B * C, (uint32_t)m, cost, bin);
              if (bin >= 0) {
                const double gain = __dsub_rn(hp, cost);
                if (better(gain, f0 + k, mine)) mine = Best{gain, f0 + k, bin};
              }
            }
            __syncthreads();
          }
        }
      }
      if (lane == 0) s_best[w] = mine;
      __syncthreads();
      Best best{0.0, -1, -1};
      for (int i = 0; i < kWarps; ++i)
        if (s_best[i].feature >= 0 && better(s_best[i].gain, s_best[i].feature, best)) best = s_best[i];
      if (best.feature < 0) continue;  // every feature constant in this node: leaf

      // partition idx[lo, hi) by code <= bin (left part first); count the left classes
      if (tid == 0) { s_ctl[4] = 0; s_ctl[5] = 0; }
      for (int c = tid; c < C; c += kThreads) s_cnt[c] = 0;
      __syncthreads();
      const int gq = best.feature >> 4, kq = best.feature & 15;
      for (int i0 = lo; i0 < hi; i0 += kThreads) {
        const int i = i0 + tid;
        bool valid = i < hi, left = false;
        int r = 0;
        if (valid) {
          r = idx[i];
          const uint8_t code = reinterpret_cast<const uint8_t*>(codes + (int64_t)gq * a.n_cap + begin + r)[kq];
          left = (int)code <= best.bin;
          if (left) atomicAdd(&s_cnt[lab[r]], 1u);
        }
        const uint32_t ml = __ballot_sync(0xffffffffu, valid && left);
        const uint32_t mr = __ballot_sync(0xffffffffu, valid && !left);
        int bl = 0, br = 0;
        if (lane == 0) {
          bl = atomicAdd(&s_ctl[4], __popc(ml));
          br = atomicAdd(&s_ctl[5], __popc(mr));
        }
        bl = __shfl_sync(0xffffffffu, bl, 0);
        br = __shfl_sync(0xffffffffu, br, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (valid) {
          if (left) idx2[lo + bl + __popc(ml & below)] = (uint16_t)r;
          else idx2[hi - 1 - (br + __popc(mr & below))] = (uint16_t)r;
        }
      }
      __syncthreads();
      for (int i = lo + tid; i < hi; i += kThreads) idx[i] = idx2[i];
      const int nl = s_ctl[4];
      if (tid == 0) {
        const int lid = s_ctl[7], rid = lid + 1;
        s_ctl[7] += 2;
        o_feature[base + rel] = best.feature;
        o_bin[base + rel] = best.bin;
        o_left[base + rel] = lid;
        o_gain[base + rel] = best.gain;
        for (int q = 0; q < 2; ++q) {
          const int id = base + lid + q;
          o_feature[id] = -1; o_bin[id] = -1; o_left[id] = -1; o_gain[id] = NAN;
          o_depth[id] = depth + 1;
        }
        int sp = s_ctl[6];
        stack[sp * 3] = rid; stack[sp * 3 + 1] = lo + nl; stack[sp * 3 + 2] = hi;
        ++sp;
        stack[sp * 3] = lid; stack[sp * 3 + 1] = lo; stack[sp * 3 + 2] = lo + nl;
        s_ctl[6] = sp + 1;
      }
      __syncthreads();
      {
        const int lid = s_ctl[7] - 2;
        for (int c = tid; c < C; c += kThreads) {
          const uint32_t lc = s_cnt[c];
          o_counts[(int64_t)(base + lid) * C + c] = lc;
          o_counts[(int64_t)(base + lid + 1) * C + c] = cnt[c] - lc;
        }
      }
      __syncthreads();
    }
    if (tid == 0) ptr_of<int>(a.o_nn)[item] = s_ctl[7];
    __syncthreads();
  }
}

int subtree_ft(int B, int C) {
  for (int ft = 16; ft >= 1; ft >>= 1)
    if (ft * B * C * 4 <= 128 * 1024) return ft;
  return 0;
}

template <int FT>
int launch(const MtbSubtreeArgs* a, int grid, cudaStream_t stream) {
  const int smem = FT * (int)a->B * (int)a->C * 4 + (int)a->T * 5 + 64;
  MTB_CHECK(cudaFuncSetAttribute(k_subtree<FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  k_subtree<FT><<<grid, kThreads, smem, stream>>>(*a);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace

// Features per shared-memory tile the subtree kernel will use for (B, C); 0 = does not fit.
extern "C" int mtb_subtree_config(int64_t B, int64_t C) { return subtree_ft((int)B, (int)C); }

// Finish every listed subtree.  `grid` CTAs (also the number of stack slices).  The caller
// sizes the output arrays with sum(2 * rows - 1) nodes and zeroes *counter.
extern "C" int mtb_subtree_build(const MtbSubtreeArgs* a, int grid, void* stream) {
  if (a->n_items <= 0) return 0;
  if (a->T > 65535 || a->T < 1) {
    mtb_set_error("subtree", "T (max rows per subtree) must be in [1, 65535]", __FILE__, __LINE__);
    return -2;
  }
  switch ((int)a->FT) {
    case 16: return launch<16>(a, grid, (cudaStream_t)stream);
    case 8: return launch<8>(a, grid, (cudaStream_t)stream);
    case 4: return launch<4>(a, grid, (cudaStream_t)stream);
    case 2: return launch<2>(a, grid, (cudaStream_t)stream);
    case 1: return launch<1>(a, grid, (cudaStream_t)stream);
    default:
      mtb_set_error("subtree", "bins x classes too large for the subtree kernel's shared-memory tile", __FILE__, __LINE__);
      return -2;
  }
}

// ------------------------------------------------------------------------------------------
// K6 (gather half) — move the rows of handed-off nodes to their owner.
// Replaces the data movement implied by the reference's rank-parity split, where every rank
// already holds all rows (mpitree/tree/decision_tree.py:447-454): here rows are sharded, so
// each rank writes its local segment of a node straight into the owner's (peer-mapped)
// receive planes at an offset agreed on the host.  One CTA per (item, plane) stripe.
namespace {

__global__ void __launch_bounds__(256) k_gather_rows(const uint4* c0, const uint4* c1, const uint8_t* l0,
                                                     const uint8_t* l1, int64_t src_n_cap, int G, const int* side,
                                                     const int* sbegin, const int* send, const int64_t* dcodes,
                                                     const int64_t* dlabels, const int64_t* dncap, const int* dbegin,
                                                     int n_items) {
  for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
    const int n = send[it] - sbegin[it];
    if (n <= 0) continue;
    const uint4* sc = side[it] ? c1 : c0;
    const uint8_t* sl = side[it] ? l1 : l0;
    uint4* dc = reinterpret_cast<uint4*>(static_cast<uintptr_t>(dcodes[it]));
    uint8_t* dl = reinterpret_cast<uint8_t*>(static_cast<uintptr_t>(dlabels[it]));
    const int64_t dn = dncap[it];
    for (int g = 0; g < G; ++g)
      for (int i = threadIdx.x; i < n; i += blockDim.x)
        st_na_u4(dc + (int64_t)g * dn + dbegin[it] + i, ld_nc_u4(sc + (int64_t)g * src_n_cap + sbegin[it] + i));
    for (int i = threadIdx.x; i < n; i += blockDim.x) dl[dbegin[it] + i] = sl[sbegin[it] + i];
  }
}

}  // namespace

extern "C" int mtb_gather_rows(const void* src_codes0, const void* src_codes1, const void* src_labels0,
                               const void* src_labels1, int64_t src_n_cap, int64_t G, const int32_t* src_side,
                               const int32_t* src_begin, const int32_t* src_end, const int64_t* dst_codes,
                               const int64_t* dst_labels, const int64_t* dst_n_cap, const int32_t* dst_begin,
                               int64_t n_items, void* stream) {
  if (n_items <= 0) return 0;
  int blocks = (int)(n_items < 148 * 8 ? n_items : 148 * 8);
  k_gather_rows<<<blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)src_codes0, (const uint4*)src_codes1,
                                                         (const uint8_t*)src_labels0, (const uint8_t*)src_labels1,
                                                         src_n_cap, (int)G, src_side, src_begin, src_end, dst_codes,
                                                         dst_labels, dst_n_cap, dst_begin, (int)n_items);
  MTB_CHECK(cudaGetLastError());
  return 0;
}
