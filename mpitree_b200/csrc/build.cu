// K2 (split search) + K3 (row partition) — level-wise tree construction on sm_100a.
//
// Replaces the reference's per-node Python loops (mpitree/tree/decision_tree.py):
// feature loop :133-138, threshold loop :76-86, entropy :49-51, arg-min/arg-max
// :88-91 / :140-141, node counts :146, mask-and-copy partition :150-164.
//
// Data layout (chosen for 128-bit loads and TMA bulk tiles):
//   codes  [G][n_cap][16] u8   feature-group planes, rows of a node are contiguous
//   labels [n_cap] u8
//   hist   [slot][G*16][C][B] u32 (class-major: a class row of B bins is contiguous), one slab per
//          frontier slot, ping-pong by level; B is a power of two, 32..256
//
// Phases of one level (each also callable on its own; mtb_level_fused runs
// plan+hist+gain+finalize in a single cooperative launch):
//   plan      work decomposition (prefix over slots), cursor reset
//   hist      persistent CTAs; a producer warp streams 1024-row x 16-byte tiles with
//             cp.async.bulk (TMA) through an mbarrier ring; 16 consumer warps do
//             shared-memory atomic increments into an [FT][B][C] tile; tiles are
//             written to the slot's local slab (plain stores when one CTA saw all of the tile's
//             rows, red.global.add otherwise); with world > 1 the slabs are partial sums that the
//             push phase copies to the owners' staging areas
//   gain      one warp per (slot, feature): sibling = parent - built child,
//             prefix over bins, float64 entropy cost with the reference's rounding,
//             first arg-min
//   finalize  per slot first arg-max over features, stop rules, child creation with
//             deterministic breadth-first ids, next frontier
//   partition unstable two-sided stream compaction of a node's rows into the other
//             ping-pong side (left rows grow up from the segment start, right rows
//             grow down from its end)
//   commit    row segments of the next frontier from the partition cursors
#include <cooperative_groups.h>
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "impurity.cuh"

namespace cg = cooperative_groups;
using namespace mtb;

namespace {

constexpr int kRows = 1024;            // rows per TMA stage
constexpr int kStageBytes = kRows * 16;
constexpr int kConsumerWarps = 16;
constexpr int kConsumers = kConsumerWarps * 32;
constexpr int kHistThreads = kConsumers + 32;  // + producer warp
constexpr int kMaxStages = 8;
constexpr int kSmemBudget = 225 * 1024;
constexpr int kPartRows = 1024;
constexpr int kPartThreads = 256;
constexpr int kGainWarps = 8;
constexpr int kFusedGainWarps = 16;  // the fused kernel's consumer warps all evaluate gains

struct HistGeom {
  int tile_words;  // FT*B*C
  int stages;
  int smem_bytes;
};

__host__ __device__ inline HistGeom hist_geom(int FT, int B, int C) {
  HistGeom g;
  g.tile_words = FT * B * C;
  int tile_bytes = g.tile_words * 4;
  int left = kSmemBudget - tile_bytes - 256;
  int st = left / kStageBytes;
  if (st > kMaxStages) st = kMaxStages;
  g.stages = st;
  g.smem_bytes = tile_bytes + st * kStageBytes + 256;
  return g;
}

// ------------------------------------------------------------------ device view
struct View {
  const MtbCtx& c;
  __device__ explicit View(const MtbCtx& ctx) : c(ctx) {}
  __device__ int* scalars() const { return ptr_of<int>(c.scalars); }
  __device__ int n_slots() const { return scalars()[0]; }
  __device__ const uint8_t* codes_cur() const { return ptr_of<const uint8_t>(c.codes[c.cur]); }
  __device__ uint8_t* codes_nxt() const { return ptr_of<uint8_t>(c.codes[c.cur ^ 1]); }
  __device__ const uint8_t* labels_cur() const { return ptr_of<const uint8_t>(c.labels[c.cur]); }
  __device__ uint8_t* labels_nxt() const { return ptr_of<uint8_t>(c.labels[c.cur ^ 1]); }
  __device__ uint32_t* hist_cur() const { return ptr_of<uint32_t>(c.hist[c.hcur]); }
  __device__ const uint32_t* hist_prev() const { return ptr_of<const uint32_t>(c.hist[c.hcur ^ 1]); }
  __device__ int64_t slab_words() const { return c.G * 16 * c.B * c.C; }
};

__device__ __forceinline__ int find_slot(const int* prefix, int n, int u) {
  // largest s with prefix[s] <= u
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (prefix[mid] <= u) lo = mid; else hi = mid;
  }
  return lo;
}

// block-wide exclusive scan of one int per thread (blockDim multiple of 32, <= 1024)
__device__ int block_excl_scan(int v, int* smem_warp /*[33]*/, int& total) {
  uint32_t inc = warp_incl_scan((uint32_t)v);
  int lane = lane_id(), w = warp_id();
  int nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 31) smem_warp[w] = (int)inc;
  __syncthreads();
  if (w == 0) {
    int x = lane < nw ? smem_warp[lane] : 0;
    uint32_t s = warp_incl_scan((uint32_t)x);
    smem_warp[lane] = (int)s - x;
    if (lane == 31) smem_warp[32] = (int)s;
  }
  __syncthreads();
  total = smem_warp[32];
  return (int)inc - v + smem_warp[w];
}

// ------------------------------------------------------------------------ plan
// Depth 0: the frontier is the root slot over all local rows.  The kernel writes it (and the scalars)
// itself, so a fit starts without any host-side fill or copy of the builder's tables.
__device__ void init_root(const MtbCtx& ctx) {
  if (threadIdx.x == 0) {
    ptr_of<int>(ctx.slot_node)[0] = 0;
    ptr_of<int>(ctx.slot_begin)[0] = 0;
    ptr_of<int>(ctx.slot_end)[0] = (int)ctx.n_local;
    ptr_of<int>(ctx.slot_mode)[0] = MTB_MODE_BUILD;
    ptr_of<int>(ctx.slot_parent)[0] = -1;
    ptr_of<int>(ctx.slot_sibling)[0] = -1;
    int* sc = ptr_of<int>(ctx.scalars);
    sc[0] = 1; sc[1] = 0; sc[2] = 1; sc[3] = 0; sc[4] = 0; sc[5] = 0; sc[6] = 0; sc[7] = 0;
    ptr_of<int>(ctx.t_left)[0] = -1;
    ptr_of<int>(ctx.t_parent)[0] = -1;
    ptr_of<int>(ctx.t_depth)[0] = 0;
  }
  __syncthreads();
}

__device__ void plan_phase(const MtbCtx& ctx, int* smem_warp) {
  View v(ctx);
  if (ctx.depth == 0) init_root(ctx);
  const int n = v.n_slots();
  const int* mode = ptr_of<const int>(ctx.slot_mode);
  const int* sb = ptr_of<const int>(ctx.slot_begin);
  const int* se = ptr_of<const int>(ctx.slot_end);
  int* prefix = ptr_of<int>(ctx.unit_prefix);
  uint32_t* cursors = ptr_of<uint32_t>(ctx.cursors);
  const int GS = (int)(ctx.G * (16 / ctx.FT));
  int carry = 0;
  for (int base = 0; base < n; base += blockDim.x) {
    int s = base + threadIdx.x;
    int units = 0;
    if (s < n) {
      if (mode[s] == MTB_MODE_BUILD) units = ((se[s] - sb[s] + kRows - 1) / kRows) * GS;
      cursors[2 * s] = 0;
      cursors[2 * s + 1] = 0;
    }
    int total;
    int ex = block_excl_scan(units, smem_warp, total);
    if (s < n) prefix[s] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    prefix[n] = carry;
    v.scalars()[3] = carry;
    v.scalars()[6] = 0;  // gain-phase work counters (pass 0 / -1, pass 1)
    v.scalars()[7] = 0;
  }
}

__global__ void k_plan(const __grid_constant__ MtbCtx ctx) {
  __shared__ int smem_warp[33];
  if (ctx.depth > 0 && ptr_of<const int>(ctx.scalars)[0] == 0) {  // empty frontier: leave a consistent "nothing to do"
    if (threadIdx.x == 0) {
      int* sc = ptr_of<int>(ctx.scalars);
      sc[1] = 0; sc[3] = 0; sc[4] = 0; sc[6] = 0; sc[7] = 0;
    }
    return;
  }
  plan_phase(ctx, smem_warp);
}

// ------------------------------------------------------------------------ hist
// Walks the CTA's contiguous range of work units (1024-row chunks ordered by slot, then
// feature tile `gs`, then chunk) without a search per unit: the slot lookup is redone only
// when the range crosses into the next slot.
struct UnitCursor {
  const int* prefix;
  const int* sb;
  const int* se;
  int n_slots, GS;
  int slot, slot_u0, slot_u1, chunks, begin, end;
  int gs, chunk, row0, nrows;

  __device__ void init(const MtbCtx& ctx, const int* prefix_, int n_slots_) {
    prefix = prefix_;
    sb = ptr_of<const int>(ctx.slot_begin);
    se = ptr_of<const int>(ctx.slot_end);
    n_slots = n_slots_;
    GS = (int)(ctx.G * (16 / ctx.FT));
    slot = -1;
    slot_u0 = slot_u1 = 0;
  }
  // position on unit u: a full search/divide only when entering a new slot, otherwise the
  // (gs, chunk) pair of the previous unit is stepped by one
  __device__ __forceinline__ void seek(int u) {
    if (u >= slot_u1 || slot < 0) {
      slot = find_slot(prefix, n_slots, u);
      slot_u0 = prefix[slot];
      slot_u1 = prefix[slot + 1];
      begin = sb[slot];
      end = se[slot];
      chunks = (end - begin + kRows - 1) / kRows;
      const int local = u - slot_u0;
      gs = local / chunks;
      chunk = local - gs * chunks;
    } else if (++chunk == chunks) {
      chunk = 0;
      ++gs;
    }
    row0 = begin + chunk * kRows;
    nrows = min(kRows, end - row0);
  }
  // true when units [u0, u1) contain every chunk of the current (slot, gs) tile
  __device__ __forceinline__ bool owns_whole_tile(int u0, int u1) const {
    const int t0 = slot_u0 + gs * chunks;
    return u0 <= t0 && t0 + chunks <= u1;
  }
};

// Interleaved-tile histogram update of one row (16 codes in `x`, label `lab`).  Lane l takes the
// features in the order l ^ 0, l ^ 1, ... l ^ 15 (l taken mod 16): in every one of the 16
// instructions the 32 lanes of a warp address 16 different features, two lanes each, and a
// feature's counters live in banks f and 16 + f only.  Worst case per instruction is therefore a
// 2-way conflict (the two lanes of a feature hitting the same bank with different cells) instead
// of the ~3.6-way average that random (bin, class) addresses give the [f][B][C] tile.  The row's
// 16 bytes are permuted once by the lane's offset so that step k reads the statically indexed
// byte k.
//
// Three instructions per increment, one per pipe: with the exclusive-or order the lane's feature
// offset is (4 l) ^ (4 k), so the per-row term (tile base + class offset + 4 l, 64-byte aligned
// apart from the 4 l) needs one LOP3 with the immediate 4 k; a byte dot product (IDP.4A with the
// multiplier 64 in byte k & 3) extracts the code, scales it and adds that term in one
// instruction; the shared-memory increment is the third.  (The additive order l + k mod 16 took
// six: add, mask, add, byte extract, multiply-add, increment.  Both address terms as integer
// multiply-adds with run-time multipliers: 3.06 instead of 2.87 ms per level.)
__device__ __forceinline__ void tile_count_il(uint32_t* tile, uint4 x, uint32_t lab, int B, int lane) {
  const uint32_t r = (uint32_t)lane & 15u;
  uint32_t w0 = x.x, w1 = x.y, w2 = x.z, w3 = x.w;
  if (r & 4u) { uint32_t t = w0; w0 = w1; w1 = t; t = w2; w2 = w3; w3 = t; }
  if (r & 8u) { uint32_t t = w0; w0 = w2; w2 = t; t = w1; w1 = w3; w3 = t; }
  const uint32_t sel = 0x3210u ^ (0x1111u * (r & 3u));
  // byte k of y = byte k ^ r of x
  const uint32_t yw[4] = {__byte_perm(w0, 0u, sel), __byte_perm(w1, 0u, sel), __byte_perm(w2, 0u, sel),
                          __byte_perm(w3, 0u, sel)};
  // cell = lab * B + code (class-major, like the slabs); shared-window byte addresses
  const uint32_t row_base = smem_u32(tile) + lab * (uint32_t)B * 64u + 4u * r;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const uint32_t addr = __dp4a(yw[k >> 2], 0x40u << (8 * (k & 3)), row_base ^ (uint32_t)(4 * k));
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(1u) : "memory");
  }
}

// Flush a finished [FT][B][C] tile into the slot's slab.  When this CTA processed every
// row of the tile it simply stores it (128-bit, no atomics, no zero-skipping); otherwise
// several CTAs share the tile and merge with red.global.add.
//
// IL (interleaved tile, FT == 16 only, opt-in): the tile is kept as [B*C][16] with the feature
// index in the low address bits, so the 16 features own two banks each and the lanes of one
// histogram instruction, which work on 16 different features (see tile_count_il), cannot collide
// across features.  The slab in global memory keeps the [f][B][C] layout, so the flush
// transposes: a warp takes 32 (bin, class) cells x 16 features at a time, reads them with four
// conflict-free 128-bit loads per lane and writes, per store instruction, four fully used
// 32-byte sectors (8 consecutive cells of 4 features).
template <int FT, bool IL = false>
__device__ void flush_tile(const MtbCtx& ctx, uint32_t* tile, int slot, int gs, int tid, bool exclusive) {
  View v(ctx);
  const int S = 16 / FT;
  const int g = gs / S, s = gs - g * S;
  const int64_t f0 = (int64_t)g * 16 + (int64_t)s * FT;
  const int words = FT * (int)ctx.B * (int)ctx.C;
  uint32_t* dst = v.hist_cur() + (int64_t)slot * v.slab_words() + f0 * ctx.B * ctx.C;
  if constexpr (IL) {
    static_assert(FT == 16, "the interleaved tile covers one whole feature group");
    const int BC = (int)ctx.B * (int)ctx.C;
    const int lane = tid & 31, w = tid >> 5;
    const int q = lane & 3, r = lane >> 2;
    const uint4 zero = make_uint4(0, 0, 0, 0);
    for (int c0 = w * 32; c0 < BC; c0 += kConsumerWarps * 32) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int cell = c0 + r + 8 * i;
        if (cell < BC) {
          uint4* src = reinterpret_cast<uint4*>(tile + cell * 16 + 4 * q);
          const uint4 x = *src;
          *src = zero;
          const uint32_t e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t* d = dst + (int64_t)(4 * q + j) * BC + cell;
            if (exclusive) *d = e[j];
            else if (e[j]) red_add_u32(d, e[j]);
          }
        }
      }
    }
    return;
  }
  if (exclusive) {
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    uint4* t4 = reinterpret_cast<uint4*>(tile);
    const uint4 zero = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < words / 4; i += kConsumers) {
      d4[i] = t4[i];
      t4[i] = zero;
    }
  } else {
    for (int i = tid; i < words; i += kConsumers) {
      uint32_t x = tile[i];
      if (x) {
        red_add_u32(dst + i, x);
        tile[i] = 0;
      }
    }
  }
}

// Interleaved tile with two 16-bit counters per word (cells 2i and 2i+1 share word i): half the
// shared memory of the 32-bit tile, which the root cluster kernel turns into row stages.  The caller
// flushes before any counter can pass 65,535.
__device__ __forceinline__ void tile_count_il16(uint32_t* tile, uint4 x, uint32_t lab, int B, int lane) {
  const int rot = lane & 15;
  uint32_t w0 = x.x, w1 = x.y, w2 = x.z, w3 = x.w;
  if (rot & 4) { const uint32_t t = w0; w0 = w1; w1 = w2; w2 = w3; w3 = t; }
  if (rot & 8) { uint32_t t = w0; w0 = w2; w2 = t; t = w1; w1 = w3; w3 = t; }
  const uint32_t sel = 0x3210u + 0x1111u * (uint32_t)(rot & 3);
  const uint4 y = make_uint4(__byte_perm(w0, w1, sel), __byte_perm(w1, w2, sel), __byte_perm(w2, w3, sel),
                             __byte_perm(w3, w0, sel));  // byte k of y = byte (k + rot) mod 16 of x
  const uint32_t row_base = smem_u32(tile) + lab * (uint32_t)(B / 2) * 64u;  // cell = lab * B + code, word = cell >> 1
  const uint32_t yw[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const uint32_t code = __byte_perm(yw[k >> 2], 0u, 0x4440u | (uint32_t)(k & 3));
    const uint32_t addr = (code >> 1) * 64u + (row_base + (uint32_t)(((k + rot) & 15) * 4));
    const uint32_t one = 1u << ((code & 1u) * 16u);
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(one) : "memory");
  }
}

// Merge (and clear) the packed tile of feature group `g` into the slab of `slot` with red.global.add.
__device__ void flush_tile_il16(const MtbCtx& ctx, uint32_t* tile, int g, int tid, int slot = 0) {
  View v(ctx);
  const int BC = (int)ctx.B * (int)ctx.C;
  uint32_t* dst = v.hist_cur() + (int64_t)slot * v.slab_words() + (int64_t)g * 16 * BC;
  const int lane = tid & 31, w = tid >> 5;
  const int q = lane & 3, r = lane >> 2;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  for (int c0 = w * 32; c0 < BC / 2; c0 += kConsumerWarps * 32) {  // word (= cell pair) index
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pair = c0 + r + 8 * i;
      if (pair < BC / 2) {
        uint4* src = reinterpret_cast<uint4*>(tile + pair * 16 + 4 * q);
        const uint4 x = *src;
        *src = zero;
        const uint32_t e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t* d = dst + (int64_t)(4 * q + j) * BC + 2 * pair;
          if (e[j] & 0xffffu) red_add_u32(d, e[j] & 0xffffu);
          if (e[j] >> 16) red_add_u32(d + 1, e[j] >> 16);
        }
      }
    }
  }
}

template <int FT, bool IL = false>
__device__ void hist_phase(const MtbCtx& ctx, uint8_t* smem) {
  View v(ctx);
  const int B = (int)ctx.B, C = (int)ctx.C;
  const HistGeom geom = hist_geom(FT, B, C);
  uint32_t* tile = reinterpret_cast<uint32_t*>(smem);
  uint8_t* stage_base = smem + (size_t)geom.tile_words * 4;
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_base + (size_t)geom.stages * kStageBytes);
  uint64_t* empty = full + kMaxStages;
  const int stages = geom.stages;

  const int total_units = v.scalars()[3];
  const int n_slots = v.n_slots();
  const int* prefix = ptr_of<const int>(ctx.unit_prefix);
  const int u0 = (int)((int64_t)total_units * blockIdx.x / gridDim.x);
  const int u1 = (int)((int64_t)total_units * (blockIdx.x + 1) / gridDim.x);

  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], kConsumerWarps);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < geom.tile_words; i += blockDim.x) tile[i] = 0;
  __syncthreads();

  const uint8_t* codes = v.codes_cur();
  const int S = 16 / FT;
  UnitCursor cur;
  cur.init(ctx, prefix, n_slots);

  if (tid >= kConsumers) {
    // ===== producer warp: one elected lane issues the bulk copies =====
    if (tid == kConsumers) {
      int st = 0;
      uint32_t ph = 0;
      for (int u = u0; u < u1; ++u) {
        cur.seek(u);
        const int g = cur.gs / S;
        mbar_wait_sleepy(&empty[st], ph ^ 1, 1000u);
        const uint32_t bytes = (uint32_t)cur.nrows * 16u;
        mbar_arrive_expect_tx(&full[st], bytes);
        tma_load_1d(stage_base + (size_t)st * kStageBytes, codes + ((int64_t)g * ctx.n_cap + cur.row0) * 16, bytes,
                    &full[st]);
        if (++st == stages) { st = 0; ph ^= 1; }
      }
    }
  } else {
    // ===== consumer warps =====
    const uint8_t* labels = v.labels_cur();
    int st = 0;
    uint32_t ph = 0;
    int cur_slot = -1, cur_gs = -1;
    bool cur_excl = false;
    for (int u = u0; u < u1; ++u) {
      cur.seek(u);
      if (cur.slot != cur_slot || cur.gs != cur_gs) {
        if (cur_slot >= 0) {
          named_bar_sync(1, kConsumers);
          flush_tile<FT, IL>(ctx, tile, cur_slot, cur_gs, tid, cur_excl);
          named_bar_sync(1, kConsumers);
        }
        cur_slot = cur.slot;
        cur_gs = cur.gs;
        cur_excl = cur.owns_whole_tile(u0, u1);
      }
      const int sub = (cur.gs % S) * FT;
      // labels first (global), then wait for the tile
      const int r_a = tid, r_b = tid + kConsumers;
      const int nrows = cur.nrows;
      uint32_t lab_a = 0, lab_b = 0;
      if (r_a < nrows) lab_a = labels[cur.row0 + r_a];
      if (r_b < nrows) lab_b = labels[cur.row0 + r_b];
      mbar_wait_sleepy(&full[st], ph, 500u);
      const uint32_t rows = smem_u32(stage_base) + (uint32_t)st * kStageBytes;  // (explicit shared-space loads: see lds_u4)
      uint4 va = make_uint4(0, 0, 0, 0), vb = va;
      if (r_a < nrows) va = lds_u4(rows + 16 * r_a);
      if (r_b < nrows) vb = lds_u4(rows + 16 * r_b);
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&empty[st]);  // tile is in registers: release the stage
      if constexpr (IL) {
        if (r_a < nrows) tile_count_il(tile, va, lab_a, B, tid);
        if (r_b < nrows) tile_count_il(tile, vb, lab_b, B, tid);
      } else {
        if (r_a < nrows) {
#pragma unroll
          for (int k = 0; k < FT; ++k) atomicAdd(&tile[(k * C + (int)lab_a) * B + (int)code_byte(va, sub + k)], 1u);
        }
        if (r_b < nrows) {
#pragma unroll
          for (int k = 0; k < FT; ++k) atomicAdd(&tile[(k * C + (int)lab_b) * B + (int)code_byte(vb, sub + k)], 1u);
        }
      }
      if (++st == stages) { st = 0; ph ^= 1; }
    }
    if (cur_slot >= 0) {
      named_bar_sync(1, kConsumers);
      flush_tile<FT, IL>(ctx, tile, cur_slot, cur_gs, tid, cur_excl);
    }
  }
}

template <int FT, bool IL = false>
__global__ void __launch_bounds__(kHistThreads, 1) k_hist(const __grid_constant__ MtbCtx ctx) {
  extern __shared__ __align__(128) uint8_t smem[];
  hist_phase<FT, IL>(ctx, smem);
}

// ------------------------------------------------------------------------ gain
// ------------------------------------------------------------- peer helpers
// Record mailbox layout (identical on every rank): fb_cost | fb_bin | fb_left | node_cnt.
struct Mailbox {
  double* cost;
  int* bin;
  uint32_t* left;
  uint32_t* node_cnt;
};
__device__ __forceinline__ Mailbox mailbox_at(const MtbCtx& ctx, int64_t base) {
  Mailbox m;
  uint8_t* p = ptr_of<uint8_t>(base);
  const int64_t MF = ctx.m_cap * ctx.F;
  m.cost = reinterpret_cast<double*>(p);
  m.bin = reinterpret_cast<int*>(p + MF * 8);
  m.left = reinterpret_cast<uint32_t*>(p + MF * 12);
  m.node_cnt = reinterpret_cast<uint32_t*>(p + MF * 12 + MF * ctx.C * 4);
  return m;
}

// Cross-GPU barrier inside a cooperative kernel.  Every CTA calls it.  Data written to peers before
// the call is visible to them after it: grid.sync orders the CTAs' writes before CTA 0's system-scope
// fence and release stores of the flags; every CTA then polls this rank's own flag words itself (no
// second grid-wide synchronisation) and its bar.sync extends the acquire to the whole CTA.
//
// Failure handling: a rank that never arrives (its process died or raised between two launches)
// would leave the others spinning for ever.  Every spin therefore also watches an abort word in the
// own flag block and a deadline on %globaltimer; a rank that gives up records the failure in
// scalars[7], raises the abort word of every peer, and from then on all barriers of the fit fall
// through, so the launches still queued finish at once (their results are garbage and the host
// raises on every rank once it has read the scalars back).
constexpr int kAbortWord = MTB_MAX_WORLD;  // u32 index in a rank's flag block

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ void peer_barrier(const MtbCtx& ctx, cg::grid_group& grid, uint32_t epoch) {
  grid.sync();
  const uint32_t* mine = ptr_of<const uint32_t>(ctx.peer_flags[ctx.rank]);
  if (threadIdx.x < (unsigned)ctx.world) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      st_release_sys_u32(ptr_of<uint32_t>(ctx.peer_flags[threadIdx.x]) + ctx.rank, epoch);
    }
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys_u32(mine + threadIdx.x) - epoch) < 0) {
      if ((++spins & 63u) == 0u) {
        if (ld_acquire_sys_u32(mine + kAbortWord) != 0u) break;
        if (ctx.barrier_timeout_ns > 0 && globaltimer_ns() - t0 > (unsigned long long)ctx.barrier_timeout_ns) {
          ptr_of<int>(ctx.scalars)[7] = 1;  // this rank gave up waiting for peer `threadIdx.x`
          for (int q = 0; q < (int)ctx.world; ++q)
            st_release_sys_u32(ptr_of<uint32_t>(ctx.peer_flags[q]) + kAbortWord, 1u);
          break;
        }
      }
    }
    __threadfence_system();
  }
  __syncthreads();
}

// world > 1: copy this rank's partial counts of every BUILD slot to the owners' staging areas with
// 128-bit stores over NVLink (the own features stay in the local slab).  A slot with fewer than
// 65,536 local rows cannot hold a count that needs more than 16 bits, so its cells travel as u16
// (half the bytes on the wire: the deep levels, whose pushes are the largest, consist of such slots);
// a format word per (source, slot) behind the staging cells tells the owner which one it got.
// `rows_from_cursors`: the slots' row segments are not committed yet (fused partition), their local
// row counts are the partition cursors.
__device__ __forceinline__ uint32_t* stage_format_table(const MtbCtx& ctx, int owner) {
  return ptr_of<uint32_t>(ctx.peer_hist[owner]) + ctx.world * ctx.h_cap * (ctx.fpo * ctx.B * ctx.C);
}

__device__ void push_phase(const MtbCtx& ctx, int n_slots, bool rows_from_cursors) {
  View v(ctx);
  const int P = (int)ctx.world, r = (int)ctx.rank;
  if (P <= 1) return;
  const int F = (int)ctx.F, fpo = (int)ctx.fpo;
  const int64_t BC = ctx.B * ctx.C;
  const int64_t region = (int64_t)fpo * BC;  // words per (source, slot) staging cell
  const int* mode = ptr_of<const int>(ctx.slot_mode);
  const int* sb = ptr_of<const int>(ctx.slot_begin);
  const int* se = ptr_of<const int>(ctx.slot_end);
  const uint32_t* cursors = ptr_of<const uint32_t>(ctx.cursors);
  const int* src_of = ptr_of<const int>(ctx.nslot_src);  // after the swap: this level's slots as children
  const int64_t items = (int64_t)n_slots * (P - 1);
  // one (slot, owner) region per CTA iteration
  for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
    const int slot = (int)(it / (P - 1));
    int o = (int)(it % (P - 1));
    if (o >= r) ++o;
    if (mode[slot] != MTB_MODE_BUILD) continue;
    const int f0 = o * fpo;
    if (f0 >= F) continue;
    const int nf = min(fpo, F - f0);
    const uint32_t rows = rows_from_cursors ? cursors[src_of[slot]] : (uint32_t)(se[slot] - sb[slot]);
    const bool narrow = rows < 65536u;
    const uint4* src = reinterpret_cast<const uint4*>(v.hist_cur() + (int64_t)slot * v.slab_words() + (int64_t)f0 * BC);
    uint4* dst = reinterpret_cast<uint4*>(ptr_of<uint32_t>(ctx.peer_hist[o]) + ((int64_t)r * ctx.h_cap + slot) * region);
    if (threadIdx.x == 0) stage_format_table(ctx, o)[(int64_t)r * ctx.h_cap + slot] = narrow ? 1u : 0u;
    if (narrow) {
      const int64_t n16 = (int64_t)nf * BC / 8;  // eight counts per 128-bit store
      for (int64_t i = threadIdx.x; i < n16; i += blockDim.x) {
        const uint4 a = ld_cg_u4(src + 2 * i), b = ld_cg_u4(src + 2 * i + 1);
        st_na_u4(dst + i, make_uint4(a.x | (a.y << 16), a.z | (a.w << 16), b.x | (b.y << 16), b.z | (b.w << 16)));
      }
    } else {
      const int64_t n16 = (int64_t)nf * BC / 4;
      for (int64_t i = threadIdx.x; i < n16; i += blockDim.x) st_na_u4(dst + i, ld_cg_u4(src + i));
    }
  }
}

__global__ void __launch_bounds__(512) k_push(const __grid_constant__ MtbCtx ctx) {
  push_phase(ctx, ptr_of<const int>(ctx.scalars)[0], false);
}

// Gain phase.  One warp per (slot, owned feature) pair, no block-level synchronisation: the warps of
// an SM work on independent pairs.  The slabs are class-major ([C][B]); lane L owns the BPL = B/32
// consecutive bins [L*BPL, (L+1)*BPL) of every class row, so a class row is one fully coalesced
// 128-bit load per lane (two for 256 bins).  Per pair:
//   pass 1   for every class (two at a time in flight): load the lane's bins - BUILD = own partial
//            (+ the peers' staged partials, or one multimem.ld_reduce when the slabs are
//            multicast-mapped), DERIVE = parent - sibling - write the reduced counts back to global
//            memory for the next level's subtraction, form the inclusive prefix over bins *in
//            registers* (lane-local adds + one warp scan of the lane totals) and park it in the
//            warp's shared-memory slab [C][32*(BPL+1)] (the +1 keeps lane-strided accesses
//            conflict-free); per-bin row counts and left totals accumulate in registers
//   pass 2   one candidate per (lane, bin): float64 cost - the reference's operation order for small
//            nodes, the k log2 k form for large ones with all of a lane's bins evaluated together
//            (2*BPL independent fp64 chains per class: an fp64 operation has ~40 cycles of latency
//            on this part) - then the first arg-min by (cost, bin) across the warp
//   publish  best-split record of the pair into every rank's mailbox
// History (100M x 64, deepest level, one B200): one warp per pair with [B][C] slabs scattered into
// shared memory and scanned there 0.97 ms (global-table gathers for k log2 k) / 0.77 ms (arithmetic
// k log2 k); a 256-thread group per pair with five group barriers 1.30 ms.
struct GainGeom {
  int warps;       // warps per CTA that take part
  int slab_words;  // shared-memory words per warp
  int cs;          // words per class row in the slab
};

constexpr int kGainTabBytes = ((kKlogkTabDoubles * 8 + 127) / 128) * 128;

__host__ __device__ inline GainGeom gain_geom(int B, int C, int warps, int smem_bytes) {
  GainGeom g;
  g.cs = 32 * (B / 32 + 1);
  g.slab_words = (C * g.cs + C + 3) & ~3;
  int by_smem = (smem_bytes - kGainTabBytes) / (g.slab_words * 4);
  g.warps = warps < by_smem ? warps : by_smem;
  return g;
}

// the lane's BPL consecutive words of a class row
template <int BPL>
struct Bins {
  uint32_t v[BPL];
};
template <int BPL>
__device__ __forceinline__ Bins<BPL> ld_bins(const uint32_t* p) {
  Bins<BPL> r;
  if constexpr (BPL >= 4) {
#pragma unroll
    for (int i = 0; i < BPL / 4; ++i) {
      const uint4 x = ld_cg_u4(p + 4 * i);
      r.v[4 * i] = x.x; r.v[4 * i + 1] = x.y; r.v[4 * i + 2] = x.z; r.v[4 * i + 3] = x.w;
    }
  } else if constexpr (BPL == 2) {
    asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(r.v[0]), "=r"(r.v[1]) : "l"(p) : "memory");
  } else {
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(r.v[0]) : "l"(p) : "memory");
  }
  return r;
}
// the same bins from a u16-packed staging cell (`p16` = the class row's first u16)
template <int BPL>
__device__ __forceinline__ Bins<BPL> ld_bins_u16(const uint16_t* p16) {
  Bins<BPL> r;
  if constexpr (BPL == 8) {
    const uint4 x = ld_cg_u4(p16);
    r.v[0] = x.x & 0xffffu; r.v[1] = x.x >> 16; r.v[2] = x.y & 0xffffu; r.v[3] = x.y >> 16;
    r.v[4] = x.z & 0xffffu; r.v[5] = x.z >> 16; r.v[6] = x.w & 0xffffu; r.v[7] = x.w >> 16;
  } else if constexpr (BPL == 4) {
    uint32_t a, b;
    asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p16) : "memory");
    r.v[0] = a & 0xffffu; r.v[1] = a >> 16; r.v[2] = b & 0xffffu; r.v[3] = b >> 16;
  } else if constexpr (BPL == 2) {
    uint32_t a;
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(a) : "l"(p16) : "memory");
    r.v[0] = a & 0xffffu; r.v[1] = a >> 16;
  } else {
    uint16_t a;
    asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(a) : "l"(p16) : "memory");
    r.v[0] = a;
  }
  return r;
}
template <int BPL>
__device__ __forceinline__ void st_bins(uint32_t* p, const Bins<BPL>& r) {
  if constexpr (BPL >= 4) {
#pragma unroll
    for (int i = 0; i < BPL / 4; ++i)
      *reinterpret_cast<uint4*>(p + 4 * i) = make_uint4(r.v[4 * i], r.v[4 * i + 1], r.v[4 * i + 2], r.v[4 * i + 3]);
  } else if constexpr (BPL == 2) {
    *reinterpret_cast<uint2*>(p) = make_uint2(r.v[0], r.v[1]);
  } else {
    *p = r.v[0];
  }
}
template <int BPL>
__device__ __forceinline__ Bins<BPL> ld_bins_mc(const uint32_t* mc) {  // sum over all ranks, formed in the switch
  Bins<BPL> r;
  if constexpr (BPL >= 2) {
#pragma unroll
    for (int i = 0; i < BPL / 2; ++i) {
      uint64_t x;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(x) : "l"(mc + 2 * i) : "memory");
      r.v[2 * i] = (uint32_t)x;
      r.v[2 * i + 1] = (uint32_t)(x >> 32);
    }
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(r.v[0]) : "l"(mc) : "memory");
  }
  return r;
}

// Cost of one candidate from the parked prefixes (L[c * CS] = left count of class c), for the paths
// that are not worth inlining eight times: the reference's exact operation order (small nodes) and gini.
__device__ __noinline__ double slow_pair_cost(const MtbCtx& ctx, const KlogK& klogk, const uint32_t* L, const uint32_t* tot,
                                              int CS, uint32_t nl, uint32_t nr, bool exact) {
  const int C = (int)ctx.C, crit = (int)ctx.criterion;
  if (exact) {
    const double* plogp = ptr_of<const double>(ctx.plogp);
    const uint32_t plogp_n = (uint32_t)ctx.plogp_n;
    const double hl = impurity(C, [&](int c) { return L[c * CS]; }, nl, crit, plogp, plogp_n);
    const double hr = impurity(C, [&](int c) { return tot[c] - L[c * CS]; }, nr, crit, plogp, plogp_n);
    return split_cost(nl, nr, hl, hr, (int)ctx.dot_mode);
  }
  return fast_split_cost(C, [&](int c) { return L[c * CS]; }, [&](int c) { return tot[c] - L[c * CS]; }, nl, nr, crit, klogk);
}

template <int BPL, bool NV>
__device__ void gain_pair(const MtbCtx& ctx, const KlogK& klogk, uint32_t* slab, const GainGeom& gg, int slot, int f) {
  View v(ctx);
  const int B = (int)ctx.B, C = (int)ctx.C, F = (int)ctx.F;
  const int P = (int)ctx.world, r = (int)ctx.rank;
  const int lane = lane_id();
  const int mode = ptr_of<const int>(ctx.slot_mode)[slot];
  const int64_t rec = (int64_t)slot * F + f;
  if (mode == MTB_MODE_NONE || mode == MTB_MODE_HANDOFF) {
    if (lane < P) {
      Mailbox m = mailbox_at(ctx, P > 1 ? ctx.peer_rec[lane] : ctx.fb_cost);
      m.cost[rec] = INFINITY;
      m.bin[rec] = -1;
    }
    return;
  }
  const int CS = gg.cs;
  uint32_t* tot = slab + C * CS;               // [C] class totals of the node
  uint32_t* mine = slab + lane * (BPL + 1);    // + c * CS + j
  const bool derive = mode == MTB_MODE_DERIVE;
  const int64_t BC = (int64_t)B * C;
  const int64_t lane_off = (int64_t)lane * BPL;

  // ---- sources / sinks of pass 1 (word pointers at class 0, this lane's bins)
  const uint32_t* own;         // local partial (or the single-rank slab)
  const uint32_t* par = nullptr;  // DERIVE: the parent's reduced slab
  uint32_t* out = nullptr;        // where the reduced counts go (nullptr: nowhere)
  const uint32_t* stage_cell = nullptr;  // staging cell of (source 0, slot), first word
  int64_t fl_off = 0;                    // counts of the owned feature within a cell
  uint32_t narrow_mask = 0;              // bit p: source p sent this slot as u16
  const uint32_t* mc = nullptr;
  int64_t src_stride = 0;
  if (P > 1) {
    // Multi-rank owner path.  Reduced slabs of the own features live in rstore (ping-pong by level);
    // the local partial slabs are never modified.
    const int fl = f - r * (int)ctx.fpo;
    const int64_t region = ctx.fpo * BC;
    const int src = derive ? ptr_of<const int>(ctx.slot_sibling)[slot] : slot;
    own = v.hist_cur() + (int64_t)src * v.slab_words() + (int64_t)f * BC + lane_off;
    stage_cell = ptr_of<const uint32_t>(ctx.peer_hist[r]) + (int64_t)src * region;
    fl_off = (int64_t)fl * BC;
    src_stride = ctx.h_cap * region;
    if constexpr (!NV) {
      const uint32_t* fmt = stage_format_table(ctx, r);
      if (lane < P && lane != r) narrow_mask = fmt[(int64_t)lane * ctx.h_cap + src] ? 1u << lane : 0u;
      narrow_mask = __reduce_or_sync(0xffffffffu, narrow_mask);
    }
    if constexpr (NV)
      mc = ptr_of<const uint32_t>(ctx.mc_hist[ctx.hcur]) + (int64_t)src * v.slab_words() + (int64_t)f * BC + lane_off;
    out = ptr_of<uint32_t>(ctx.rstore[ctx.hcur]) + (int64_t)slot * region + (int64_t)fl * BC + lane_off;
    if (derive)
      par = ptr_of<const uint32_t>(ctx.rstore[ctx.hcur ^ 1]) + (int64_t)ptr_of<const int>(ctx.slot_parent)[slot] * region +
            (int64_t)fl * BC + lane_off;
  } else if (derive) {
    own = v.hist_cur() + (int64_t)ptr_of<const int>(ctx.slot_sibling)[slot] * v.slab_words() + (int64_t)f * BC + lane_off;
    par = v.hist_prev() + (int64_t)ptr_of<const int>(ctx.slot_parent)[slot] * v.slab_words() + (int64_t)f * BC + lane_off;
    out = v.hist_cur() + (int64_t)slot * v.slab_words() + (int64_t)f * BC + lane_off;
  } else {
    own = v.hist_cur() + (int64_t)slot * v.slab_words() + (int64_t)f * BC + lane_off;
  }

  // ---- pass 1: load, reduce, prefix over bins in registers, park in shared memory
  uint32_t rows_here[BPL], nl[BPL];
#pragma unroll
  for (int j = 0; j < BPL; ++j) rows_here[j] = nl[j] = 0;
  constexpr int kCB = 2;  // classes in flight together
  for (int c0 = 0; c0 < C; c0 += kCB) {
    Bins<BPL> h[kCB];
#pragma unroll
    for (int u = 0; u < kCB; ++u) {
      const int c = min(c0 + u, C - 1);  // a class past the end repeats the last one (discarded below)
      if constexpr (NV) h[u] = ld_bins_mc<BPL>(mc + (int64_t)c * B);
      else h[u] = ld_bins<BPL>(own + (int64_t)c * B);
    }
    if (P > 1 && !NV) {
      // staged partials of the peers, two sources (x kCB classes) in flight together: the pair's
      // latency is a chain of L2 round trips, one per group of loads
      constexpr int kPS = 2;  // (four overflow the 96-register budget of the fused kernel into local memory)
      for (int p0 = 0; p0 < P; p0 += kPS) {
        Bins<BPL> y[kPS][kCB];
#pragma unroll
        for (int i = 0; i < kPS; ++i) {
          const int p = p0 + i;
          const bool live = p < P && p != r;
          const uint32_t* cell = stage_cell + (int64_t)(live ? p : r) * src_stride;  // this source's cell of the slot
          if (live && (narrow_mask >> p & 1u)) {
#pragma unroll
            for (int u = 0; u < kCB; ++u)
              y[i][u] = ld_bins_u16<BPL>(reinterpret_cast<const uint16_t*>(cell) + (int64_t)fl_off + (int64_t)min(c0 + u, C - 1) * B + lane_off);
          } else if (live) {
#pragma unroll
            for (int u = 0; u < kCB; ++u) y[i][u] = ld_bins<BPL>(cell + fl_off + (int64_t)min(c0 + u, C - 1) * B + lane_off);
          } else {
#pragma unroll
            for (int u = 0; u < kCB; ++u)
#pragma unroll
              for (int j = 0; j < BPL; ++j) y[i][u].v[j] = 0;
          }
        }
#pragma unroll
        for (int i = 0; i < kPS; ++i)
#pragma unroll
          for (int u = 0; u < kCB; ++u)
#pragma unroll
            for (int j = 0; j < BPL; ++j) h[u].v[j] += y[i][u].v[j];
      }
    }
    if (derive) {
      Bins<BPL> a[kCB];
#pragma unroll
      for (int u = 0; u < kCB; ++u) a[u] = ld_bins<BPL>(par + (int64_t)min(c0 + u, C - 1) * B);
#pragma unroll
      for (int u = 0; u < kCB; ++u)
#pragma unroll
        for (int j = 0; j < BPL; ++j) h[u].v[j] = a[u].v[j] - h[u].v[j];
    }
#pragma unroll
    for (int u = 0; u < kCB; ++u) {
      const int c = c0 + u;
      if (c < C) {
        if (out) st_bins<BPL>(out + (int64_t)c * B, h[u]);
        uint32_t run = 0;
#pragma unroll
        for (int j = 0; j < BPL; ++j) {
          rows_here[j] += h[u].v[j];
          run += h[u].v[j];
          h[u].v[j] = run;  // lane-local inclusive prefix
        }
        const uint32_t inc = warp_incl_scan(run);
        const uint32_t off = inc - run;
#pragma unroll
        for (int j = 0; j < BPL; ++j) {
          const uint32_t L = h[u].v[j] + off;
          nl[j] += L;
          mine[c * CS + j] = L;
        }
        if (lane == 31) tot[c] = inc;
      }
    }
  }
  __syncwarp();
  uint32_t n = 0;
  for (int c = 0; c < C; ++c) n += tot[c];

  // ---- pass 2: one candidate per (lane, bin)
  double best = INFINITY;
  int best_bin = 0x7fffffff;
  const int crit = (int)ctx.criterion;
  // small nodes: the reference's exact float64 operation order; large nodes: the cheap form
  const bool exact = ctx.plogp == 0 || n <= (uint32_t)ctx.plogp_n;
  if (!exact && crit == 0) {
    // summation order of fast_split_cost: s = g(nL) + g(nR), then s -= g(kL_c) + g(kR_c) for ascending c
    double s[BPL];
#pragma unroll
    for (int j = 0; j < BPL; ++j) s[j] = __dadd_rn(klogk(nl[j]), klogk(n - nl[j]));
    for (int c = 0; c < C; ++c) {
      const uint32_t tc = tot[c];
      double a[BPL], b[BPL];
#pragma unroll
      for (int j = 0; j < BPL; ++j) {
        const uint32_t l = mine[c * CS + j];
        a[j] = klogk(l);
        b[j] = klogk(tc - l);
      }
#pragma unroll
      for (int j = 0; j < BPL; ++j) s[j] = __dsub_rn(s[j], __dadd_rn(a[j], b[j]));
    }
    const double dn = u2d(n);
#pragma unroll
    for (int j = 0; j < BPL; ++j) {
      if (rows_here[j] != 0u && nl[j] != n) {  // empty bins repeat a candidate; an empty right side is no split
        const double cost = __ddiv_rn(s[j], dn);
        if (cost < best) {  // strict: first (lowest) bin wins ties inside the lane
          best = cost;
          best_bin = lane * BPL + j;
        }
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < BPL; ++j) {
      if (rows_here[j] == 0u || nl[j] == n) continue;
      const double cost = slow_pair_cost(ctx, klogk, mine + j, tot, CS, nl[j], n - nl[j], exact);
      if (cost < best) {
        best = cost;
        best_bin = lane * BPL + j;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double oc = __shfl_xor_sync(0xffffffffu, best, o);
    int ob = __shfl_xor_sync(0xffffffffu, best_bin, o);
    if (oc < best || (oc == best && ob < best_bin)) {
      best = oc;
      best_bin = ob;
    }
  }
  const bool ok = best_bin != 0x7fffffff;
  const uint32_t* Lb = slab + (ok ? (best_bin / BPL) * (BPL + 1) + best_bin % BPL : 0);  // + c * CS
  // publish the record to every rank's mailbox (the own one when world == 1)
  for (int q = 0; q < P; ++q) {
    Mailbox m = mailbox_at(ctx, P > 1 ? ctx.peer_rec[q] : ctx.fb_cost);
    if (lane == 0) {
      m.cost[rec] = ok ? best : INFINITY;
      m.bin[rec] = ok ? best_bin : -1;
    }
    if (ok)
      for (int c = lane; c < C; c += 32) m.left[rec * C + c] = Lb[c * CS];
    if (f == 0)
      for (int c = lane; c < C; c += 32) m.node_cnt[(int64_t)slot * C + c] = tot[c];
  }
  __syncwarp();
}

// `warps`: warps of the CTA that may take part (the first ones); `smem_bytes`: dynamic shared memory
// available to the phase (fewer warps work when their slabs do not all fit).  pass -1: every pair;
// 0 / 1: BUILD+NONE / DERIVE slots only (kept for experiments: pairs never depend on one another).
template <bool NV = false>
__device__ void gain_phase(const MtbCtx& ctx, uint8_t* smem, int warps, int smem_bytes, int pass, int n_slots) {
  const int F = (int)ctx.F;
  const int f_lo = (int)(ctx.rank * ctx.fpo);
  const int f_hi = min(F, f_lo + (int)ctx.fpo);
  const int n_own = max(0, f_hi - f_lo);
  const int n_pairs = n_slots * n_own;
  const GainGeom gg = gain_geom((int)ctx.B, (int)ctx.C, warps, smem_bytes);
  // constants of g(k) = k log2 k into shared memory (33 KB), coefficients into registers
  double* tab = reinterpret_cast<double*>(smem);
  const double* gtab = ptr_of<const double>(ctx.klogk);
  for (int i = threadIdx.x; i < kKlogkTabDoubles; i += blockDim.x) tab[i] = gtab[i];
  __syncthreads();
  const int w = warp_id();
  if (w >= gg.warps) return;
  KlogK klogk;
  klogk.bind(tab, gtab);
  uint32_t* slab = reinterpret_cast<uint32_t*>(smem + kGainTabBytes) + (size_t)w * gg.slab_words;
  const int* mode = ptr_of<const int>(ctx.slot_mode);
  const int total_warps = gg.warps * (int)gridDim.x;
  const int bpl = (int)ctx.B / 32;
  // pairs are dealt round-robin, CTAs first: a level with few pairs spreads over the SMs, and the
  // pairs of one slot (equal cost) spread over the warps
  for (int p = w * (int)gridDim.x + (int)blockIdx.x; p < n_pairs; p += total_warps) {
    const int slot = p / n_own;
    if (pass >= 0 && (mode[slot] == MTB_MODE_DERIVE) != (pass == 1)) continue;
    const int f = f_lo + p % n_own;
    switch (bpl) {
      case 8: gain_pair<8, NV>(ctx, klogk, slab, gg, slot, f); break;
      case 4: gain_pair<4, NV>(ctx, klogk, slab, gg, slot, f); break;
      case 2: gain_pair<2, NV>(ctx, klogk, slab, gg, slot, f); break;
      default: gain_pair<1, NV>(ctx, klogk, slab, gg, slot, f); break;
    }
  }
}

__global__ void __launch_bounds__(kGainWarps * 32) k_gain(const __grid_constant__ MtbCtx ctx, int pass, int smem_bytes) {
  extern __shared__ __align__(128) uint8_t smem[];
  gain_phase(ctx, smem, kGainWarps, smem_bytes, pass, ptr_of<const int>(ctx.scalars)[0]);
}

// -------------------------------------------------------------------- finalize
template <typename Cnt>
__device__ __forceinline__ bool is_terminal(Cnt cnt, int C, int depth, const MtbCtx& ctx, uint32_t& n) {
  int present = 0;
  n = 0;
  for (int c = 0; c < C; ++c) {
    const uint32_t x = cnt(c);
    n += x;
    present += x != 0u;
  }
  return present <= 1 || (ctx.max_depth >= 0 && depth >= ctx.max_depth) || (int64_t)n < ctx.min_samples_split;
}

// `ktab`: the k log2 k constants (the shared-memory copy left by the gain phase in the fused kernels)
__device__ void finalize_phase(const MtbCtx& ctx, int* smem_warp, const double* ktab) {
  View v(ctx);
  const int n = v.n_slots();
  const int F = (int)ctx.F, C = (int)ctx.C;
  const int depth = (int)ctx.depth;
  int* scal = v.scalars();
  const int nodes0 = scal[2];

  const int* slot_node = ptr_of<const int>(ctx.slot_node);
  const int* slot_mode = ptr_of<const int>(ctx.slot_mode);
  const int* sb = ptr_of<const int>(ctx.slot_begin);
  const int* se = ptr_of<const int>(ctx.slot_end);
  const double* fb_cost = ptr_of<const double>(ctx.fb_cost);
  const int* fb_bin = ptr_of<const int>(ctx.fb_bin);
  const uint32_t* fb_left = ptr_of<const uint32_t>(ctx.fb_left);
  const uint32_t* node_cnt = ptr_of<const uint32_t>(ctx.node_cnt);
  int* rec_feature = ptr_of<int>(ctx.rec_feature);
  int* rec_bin = ptr_of<int>(ctx.rec_bin);
  int* part_flags = ptr_of<int>(ctx.part_flags);
  int* t_feature = ptr_of<int>(ctx.t_feature);
  int* t_bin = ptr_of<int>(ctx.t_bin);
  int* t_left = ptr_of<int>(ctx.t_left);
  int* t_parent = ptr_of<int>(ctx.t_parent);
  int* t_depth = ptr_of<int>(ctx.t_depth);
  double* t_gain = ptr_of<double>(ctx.t_gain);
  uint32_t* t_counts = ptr_of<uint32_t>(ctx.t_counts);
  int* ns_node = ptr_of<int>(ctx.nslot_node);
  int* ns_mode = ptr_of<int>(ctx.nslot_mode);
  int* ns_parent = ptr_of<int>(ctx.nslot_parent);
  int* ns_sibling = ptr_of<int>(ctx.nslot_sibling);
  int* ns_src = ptr_of<int>(ctx.nslot_src);
  int* pprefix = ptr_of<int>(ctx.unit_prefix) + (ctx.m_cap + 1);

  int carry_split = 0, carry_pair = 0, carry_units = 0;
  for (int base = 0; base < n; base += blockDim.x) {
    const int s = base + threadIdx.x;
    int split = 0, pair = 0, units = 0;
    int bf = -1, bb = -1;
    double bgain = 0.0;
    bool tl = true, tr = true;   // child is a leaf by the stop rules
    bool hl = false, hr = false; // child leaves the level-wise build (handed to the subtree phase)
    uint32_t nl = 0, nr = 0;
    if (s < n && (slot_mode[s] == MTB_MODE_BUILD || slot_mode[s] == MTB_MODE_DERIVE)) {
      const uint32_t* cnt = node_cnt + (int64_t)s * C;
      const int node = slot_node[s];
      for (int c = 0; c < C; ++c) t_counts[(int64_t)node * C + c] = cnt[c];
      uint32_t ntot;
      if (!is_terminal([&](int c) { return cnt[c]; }, C, depth, ctx, ntot)) {
        const bool exact = ctx.plogp == 0 || ntot <= (uint32_t)ctx.plogp_n;
        KlogK klogk;
        klogk.bind(ktab, ptr_of<const double>(ctx.klogk));
        const double hp = exact ? impurity(C, [&](int c) { return cnt[c]; }, ntot, (int)ctx.criterion,
                                           ptr_of<const double>(ctx.plogp), (uint32_t)ctx.plogp_n)
                                : fast_impurity(C, [&](int c) { return cnt[c]; }, ntot, (int)ctx.criterion, klogk);
        bool have = false;
        // costs of sixteen features are loaded together: the loop is bound by load latency otherwise
        for (int f0 = 0; f0 < F; f0 += 16) {
          double cst[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) cst[j] = f0 + j < F ? fb_cost[(int64_t)s * F + f0 + j] : INFINITY;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (!(cst[j] < INFINITY)) continue;
            const double g = __dsub_rn(hp, cst[j]);
            if (!have || g > bgain) {  // strict: first (lowest) feature wins ties
              have = true;
              bgain = g;
              bf = f0 + j;
            }
          }
        }
        if (have) bb = fb_bin[(int64_t)s * F + bf];
        if (have) {
          split = 1;
          const uint32_t* L = fb_left + ((int64_t)s * F + bf) * C;
          tl = is_terminal([&](int c) { return L[c]; }, C, depth + 1, ctx, nl);
          tr = is_terminal([&](int c) { return cnt[c] - L[c]; }, C, depth + 1, ctx, nr);
          hl = !tl && (int64_t)nl <= ctx.small_rows;
          hr = !tr && (int64_t)nr <= ctx.small_rows;
          pair = !(tl && tr);
          if (pair) units = (se[s] - sb[s] + kPartRows - 1) / kPartRows;
        }
      }
      rec_feature[s] = bf;
      rec_bin[s] = bb;
      t_feature[node] = bf;
      t_bin[node] = bb;
      t_gain[node] = split ? bgain : NAN;
      // Which children's rows must be written by the partition.  With the fused partition +
      // histogram path (part_child set) a child's histogram is accumulated while its parent's rows
      // stream by, so its own rows are only needed if *its* children can be split again
      // (depth + 2 < max_depth) or if it is handed to the subtree kernel.
      const bool rows_needed = ctx.part_child == 0 || ctx.max_depth < 0 || depth + 2 < ctx.max_depth;
      const bool keep_l = !tl && (rows_needed || hl), keep_r = !tr && (rows_needed || hr);
      part_flags[s] = pair ? ((keep_l ? 1 : 0) | (keep_r ? 2 : 0)) : 0;
      if (ctx.part_child) ptr_of<int>(ctx.part_child)[s] = -1;
    } else if (s < n) {
      rec_feature[s] = -1;
      rec_bin[s] = -1;
      part_flags[s] = 0;
      if (ctx.part_child) ptr_of<int>(ctx.part_child)[s] = -1;
    }
    if (s < n) {  // the partition of these slots' rows counts into fresh cursors
      ptr_of<uint32_t>(ctx.cursors)[2 * s] = 0;
      ptr_of<uint32_t>(ctx.cursors)[2 * s + 1] = 0;
    }
    int tot_split, tot_pair, tot_units;
    const int ex_split = block_excl_scan(split, smem_warp, tot_split);
    const int ex_pair = block_excl_scan(pair, smem_warp, tot_pair);
    const int ex_units = block_excl_scan(units, smem_warp, tot_units);
    if (s < n) pprefix[s] = carry_units + ex_units;
    if (split) {
      const int node = slot_node[s];
      const int lid = nodes0 + 2 * (carry_split + ex_split), rid = lid + 1;
      t_left[node] = lid;
      const uint32_t* cnt = node_cnt + (int64_t)s * C;
      const uint32_t* L = fb_left + ((int64_t)s * F + bf) * C;
      for (int c = 0; c < C; ++c) {
        t_counts[(int64_t)lid * C + c] = L[c];
        t_counts[(int64_t)rid * C + c] = cnt[c] - L[c];
      }
      t_parent[lid] = node; t_parent[rid] = node;
      t_depth[lid] = depth + 1; t_depth[rid] = depth + 1;
      t_feature[lid] = -1; t_feature[rid] = -1;
      t_bin[lid] = -1; t_bin[rid] = -1;
      t_left[lid] = -1; t_left[rid] = -1;
      t_gain[lid] = NAN; t_gain[rid] = NAN;
      if (pair) {
        const int a = 2 * (carry_pair + ex_pair), b = a + 1;
        if (ctx.part_child) ptr_of<int>(ctx.part_child)[s] = a;
        ns_node[a] = lid; ns_node[b] = rid;
        ns_parent[a] = s; ns_parent[b] = s;
        ns_sibling[a] = b; ns_sibling[b] = a;
        ns_src[a] = 2 * s; ns_src[b] = 2 * s + 1;
        int ma, mb;
        if (hl || hr) {
          // a handed-off child has no histogram, so its sibling (if it continues) is built directly
          ma = tl ? MTB_MODE_NONE : (hl ? MTB_MODE_HANDOFF : MTB_MODE_BUILD);
          mb = tr ? MTB_MODE_NONE : (hr ? MTB_MODE_HANDOFF : MTB_MODE_BUILD);
        } else if (!tl && !tr) {
          // both children continue: build the smaller (global count, ties -> left), derive the other
          const bool left_small = nl <= nr;
          ma = left_small ? MTB_MODE_BUILD : MTB_MODE_DERIVE;
          mb = left_small ? MTB_MODE_DERIVE : MTB_MODE_BUILD;
        } else {
          ma = tl ? MTB_MODE_NONE : MTB_MODE_BUILD;
          mb = tr ? MTB_MODE_NONE : MTB_MODE_BUILD;
        }
        ns_mode[a] = ma; ns_mode[b] = mb;
      }
    }
    carry_split += tot_split;
    carry_pair += tot_pair;
    carry_units += tot_units;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    pprefix[n] = carry_units;
    scal[1] = 2 * carry_pair;
    scal[2] = nodes0 + 2 * carry_split;
    scal[4] = carry_units;
  }
}

__global__ void __launch_bounds__(1024) k_finalize(const __grid_constant__ MtbCtx ctx) {
  __shared__ int smem_warp[33];
  if (ptr_of<const int>(ctx.scalars)[0] == 0) return;
  finalize_phase(ctx, smem_warp, ptr_of<const double>(ctx.klogk));
}

// ------------------------------------------------------------------- partition
__global__ void __launch_bounds__(kPartThreads) k_partition(const __grid_constant__ MtbCtx ctx) {
  View v(ctx);
  __shared__ uint32_t wl[4][kPartThreads / 32], wr[4][kPartThreads / 32];
  __shared__ uint32_t base_l, base_r;
  const int n_slots = v.n_slots();
  const int total = v.scalars()[4];
  const int* pprefix = ptr_of<const int>(ctx.unit_prefix) + (ctx.m_cap + 1);
  const int* sb = ptr_of<const int>(ctx.slot_begin);
  const int* se = ptr_of<const int>(ctx.slot_end);
  const int* rec_feature = ptr_of<const int>(ctx.rec_feature);
  const int* rec_bin = ptr_of<const int>(ctx.rec_bin);
  const int* part_flags = ptr_of<const int>(ctx.part_flags);
  uint32_t* cursors = ptr_of<uint32_t>(ctx.cursors);
  const uint4* src = reinterpret_cast<const uint4*>(v.codes_cur());
  uint4* dst = reinterpret_cast<uint4*>(v.codes_nxt());
  const uint8_t* lsrc = v.labels_cur();
  uint8_t* ldst = v.labels_nxt();
  const int G = (int)ctx.G;
  const int64_t plane = ctx.n_cap;
  const int tid = threadIdx.x, lane = lane_id(), w = warp_id();

  const int u0 = (int)((int64_t)total * blockIdx.x / gridDim.x);
  const int u1 = (int)((int64_t)total * (blockIdx.x + 1) / gridDim.x);
  for (int u = u0; u < u1; ++u) {
    const int slot = find_slot(pprefix, n_slots, u);
    const int chunk = u - pprefix[slot];
    const int b = sb[slot], e = se[slot];
    const int r0 = b + chunk * kPartRows;
    const int r1 = min(e, r0 + kPartRows);
    const int f = rec_feature[slot], bin = rec_bin[slot], flags = part_flags[slot];
    const int g = f >> 4, k = f & 15;

    bool valid[4], left[4];
    uint32_t rank[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int row = r0 + it * kPartThreads + tid;
      valid[it] = row < r1;
      uint32_t code = 0;
      if (valid[it]) code = reinterpret_cast<const uint8_t*>(src + (int64_t)g * plane + row)[k];
      left[it] = valid[it] && (int)code <= bin;
      const uint32_t ml = __ballot_sync(0xffffffffu, left[it]);
      const uint32_t mr = __ballot_sync(0xffffffffu, valid[it] && !left[it]);
      const uint32_t below = (1u << lane) - 1u;
      rank[it] = left[it] ? __popc(ml & below) : __popc(mr & below);
      if (lane == 0) {
        wl[it][w] = __popc(ml);
        wr[it][w] = __popc(mr);
      }
    }
    __syncthreads();
    if (w == 0) {
      // exclusive scan over the 4*NW (iteration, warp) cells, in row order
      const int cell = lane;  // 4*NW == 32
      uint32_t cl = (&wl[0][0])[cell], cr = (&wr[0][0])[cell];
      uint32_t il = warp_incl_scan(cl), ir = warp_incl_scan(cr);
      (&wl[0][0])[cell] = il - cl;
      (&wr[0][0])[cell] = ir - cr;
      if (lane == 31) {
        base_l = atomicAdd(&cursors[2 * slot], il);
        base_r = atomicAdd(&cursors[2 * slot + 1], ir);
      }
    }
    __syncthreads();
    const bool keep_l = flags & 1, keep_r = flags & 2;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      if (!valid[it]) continue;
      const int row = r0 + it * kPartThreads + tid;
      int64_t d;
      if (left[it]) {
        if (!keep_l) continue;
        d = (int64_t)b + base_l + wl[it][w] + rank[it];
      } else {
        if (!keep_r) continue;
        d = (int64_t)e - 1 - (base_r + wr[it][w] + rank[it]);
      }
      for (int gg = 0; gg < G; ++gg) st_na_u4(dst + (int64_t)gg * plane + d, ld_nc_u4(src + (int64_t)gg * plane + row));
      ldst[d] = lsrc[row];
    }
    __syncthreads();
  }
}
static_assert(4 * (kPartThreads / 32) == 32, "partition scan assumes 32 cells");

// `first`, `stride`: the calling thread's index and the thread count of whoever runs the phase (the
// whole grid in k_commit, CTA 0 alone in the fused kernel)
__device__ void commit_phase(const MtbCtx& ctx, int first, int stride) {
  View v(ctx);
  const int n_next = v.scalars()[1];
  const int* ns_src = ptr_of<const int>(ctx.nslot_src);
  const int* ns_mode = ptr_of<const int>(ctx.nslot_mode);
  const int* sb = ptr_of<const int>(ctx.slot_begin);
  const int* se = ptr_of<const int>(ctx.slot_end);
  const uint32_t* cursors = ptr_of<const uint32_t>(ctx.cursors);
  int* nb = ptr_of<int>(ctx.nslot_begin);
  int* ne = ptr_of<int>(ctx.nslot_end);
  for (int j = first; j < n_next; j += stride) {
    const int p = ns_src[j] >> 1, side = ns_src[j] & 1;
    int b = 0, e = 0;
    if (ns_mode[j] != MTB_MODE_NONE) {
      if (side == 0) {
        b = sb[p];
        e = b + (int)cursors[2 * p];
      } else {
        e = se[p];
        b = e - (int)cursors[2 * p + 1];
      }
    }
    nb[j] = b;
    ne[j] = e;
    if (ns_mode[j] == MTB_MODE_HANDOFF) {
      // rows are in place (side cur^1): list the node for the subtree phase.  Order of the
      // entries is arbitrary; the host sorts them by node id.
      const int k = atomicAdd(v.scalars() + 5, 1);
      if (k < (int)ctx.ho_cap) {
        ptr_of<int>(ctx.ho_node)[k] = ptr_of<const int>(ctx.nslot_node)[j];
        ptr_of<int>(ctx.ho_begin)[k] = b;
        ptr_of<int>(ctx.ho_end)[k] = e;
        ptr_of<int>(ctx.ho_side)[k] = (int)(ctx.cur ^ 1);
      }
    }
  }
}

__global__ void k_commit(const __grid_constant__ MtbCtx ctx) {
  commit_phase(ctx, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
  // the next level's slot count becomes current (nothing in this kernel reads scalars[0])
  if (blockIdx.x == 0 && threadIdx.x == 0) ptr_of<int>(ctx.scalars)[0] = ptr_of<int>(ctx.scalars)[1];
}

// ------------------------------------------------ partition fused with the next level's histograms
// Levels >= 1.  P is the finished parent level, C the same buffers seen from the children's
// level.  Two grid-wide steps replace the separate partition kernel and the histogram phase:
//
//   move A   the (parent slot, split feature's plane) tiles.  Rows, labels arrive by TMA; three
//            *ranker* warps read the split feature's byte of each unit's 1,024 rows (from global
//            memory, up to twelve units ahead), form the "goes left" bit words and their prefix,
//            reserve the unit's destination ranges with one atomic pair per unit, and hand the unit
//            record to the consumer warps through a shared-memory ring (and to step B through global
//            memory).  The
//            consumers rank rows from the bit words, write them to their child's segment and - for
//            rows of the child that is built directly - increment the child's histogram tile;
//   move B   all other (parent slot, feature tile) pairs: same consumers, the unit record now
//            arrives by TMA next to the rows.
//   commit   children's row segments (+ hand-off list).
//
// Every plane is read exactly once per level (round 1 had a separate pre-pass over the split
// feature's plane: 0.4 ms of a 3.4 ms level on 100M x 64).
constexpr int kMetaWords = 68;  // per partition unit: 32 "goes left" words, 32 exclusive left-count prefixes, base_l, base_r, 2 pad (272 B)
constexpr int kRankers = 3;  // ranker warps: a unit's record waits ~2 us for its atomics, units are dealt round-robin
constexpr int kFusedThreads = kConsumers + 32 + 32 * kRankers;  // consumers + producer warp + ranker warps (20 warps:
                                                                // five per scheduler, the same 96-register budget as 17)

// cursor over (parent slot, tile ordinal j, 1024-row unit) triples, slot-major; `tiles` ordinals per slot
struct MoveCursor {
  const int* pprefix;
  const int* sb;
  const int* se;
  int n_slots, tiles;
  int slot, slot_u0, slot_u1, chunks, begin, end;
  int j, chunk, row0, nrows;
  __device__ __forceinline__ int find(int u) const {
    int lo = 0, hi = n_slots;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (pprefix[mid] * tiles <= u) lo = mid; else hi = mid;
    }
    return lo;
  }
  __device__ void init(const MtbCtx& P, int n_slots_, int tiles_) {
    pprefix = ptr_of<const int>(P.unit_prefix) + (P.m_cap + 1);
    sb = ptr_of<const int>(P.slot_begin);
    se = ptr_of<const int>(P.slot_end);
    n_slots = n_slots_;
    tiles = tiles_;
    slot = -1;
    slot_u0 = slot_u1 = 0;
  }
  __device__ __forceinline__ void seek(int u) {
    if (u >= slot_u1 || slot < 0) {
      slot = find(u);
      slot_u0 = pprefix[slot] * tiles;
      slot_u1 = pprefix[slot + 1] * tiles;
      begin = sb[slot];
      end = se[slot];
      chunks = pprefix[slot + 1] - pprefix[slot];
      const int local = u - slot_u0;
      j = local / chunks;
      chunk = local - j * chunks;
    } else if (++chunk == chunks) {
      chunk = 0;
      ++j;
    }
    row0 = begin + chunk * kPartRows;
    nrows = min(kPartRows, end - row0);
  }
  // position on an arbitrary unit (the rankers stride through the range)
  __device__ __forceinline__ void seek_to(int u) {
    if (u >= slot_u1 || slot < 0) {
      slot = find(u);
      slot_u0 = pprefix[slot] * tiles;
      slot_u1 = pprefix[slot + 1] * tiles;
      begin = sb[slot];
      end = se[slot];
      chunks = pprefix[slot + 1] - pprefix[slot];
    }
    const int local = u - slot_u0;
    j = local / chunks;
    chunk = local - j * chunks;
    row0 = begin + chunk * kPartRows;
    nrows = min(kPartRows, end - row0);
  }
  __device__ __forceinline__ int unit_index() const { return pprefix[slot] + chunk; }
  __device__ __forceinline__ bool owns_whole_tile(int u0, int u1) const {
    const int t0 = slot_u0 + j * chunks;
    return u0 <= t0 && t0 + chunks <= u1;
  }
};

constexpr int kLabelStage = 1056;                               // 1024 labels + alignment slack
constexpr int kStage2Bytes = kStageBytes + kLabelStage + kMetaWords * 4;  // + unit record (272 B), 16-byte multiple

// `packed`: the interleaved 16-feature tile with two 16-bit counters per word (half the shared memory,
// so eight row stages fit next to it instead of three; flushed before a counter can overflow)
constexpr int kRecRing = 12;                                   // unit records the rankers may be ahead of the consumers
constexpr int kRecRingBytes = kRecRing * kMetaWords * 4 + 256;  // + ring barriers
__host__ __device__ inline HistGeom move_geom(int FT, int B, int C, bool packed = false) {
  HistGeom g;
  g.tile_words = packed ? FT * B * C / 2 : FT * B * C;
  int tile_bytes = g.tile_words * 4;
  int st = (kSmemBudget - tile_bytes - 256 - kRecRingBytes) / kStage2Bytes;
  if (st > kMaxStages) st = kMaxStages;
  g.stages = st;
  g.smem_bytes = tile_bytes + st * kStage2Bytes + 256 + kRecRingBytes;
  return g;
}
constexpr int kPackedFlushUnits = 63;  // 63 x 1,024 rows < 65,536: no 16-bit counter can overflow between flushes

// ring position of a pipeline role, carried from step A to step B
struct Ring {
  int st;
  uint32_t ph;
  __device__ __forceinline__ void next(int stages) {
    if (++st == stages) { st = 0; ph ^= 1; }
  }
};

struct MoveSmem {
  uint32_t* tile;
  uint8_t* stage_base;
  uint64_t *full, *empty, *taken;  // taken: step A, the unit's ranker has read its bytes of the stage
  uint32_t* rec_ring;             // [kRecRing][kMetaWords]: unit records of step A, written by the rankers
  uint64_t *rec_ready, *rec_free;  // [kRecRing]
  int stages;
};

__device__ __forceinline__ MoveSmem move_smem(uint8_t* smem, const HistGeom& geom) {
  MoveSmem m;
  m.tile = reinterpret_cast<uint32_t*>(smem);
  m.stage_base = smem + (size_t)geom.tile_words * 4;
  m.full = reinterpret_cast<uint64_t*>(m.stage_base + (size_t)geom.stages * kStage2Bytes);
  m.empty = m.full + kMaxStages;
  m.taken = m.empty + kMaxStages;
  m.rec_ring = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(m.full) + 256);
  m.rec_ready = reinterpret_cast<uint64_t*>(m.rec_ring + kRecRing * kMetaWords);
  m.rec_free = m.rec_ready + kRecRing;
  m.stages = geom.stages;
  return m;
}

// step 0: the split feature's plane of every parent (ranker warp active); step 1: every other tile
template <int FT, bool IL = false>
__device__ void move_hist_phase(const MtbCtx& P, const MtbCtx& C, uint8_t* smem, int step, Ring& ring) {
  View vp(P);
  const int B = (int)P.B, Cn = (int)P.C;
  // (a 16-bit packed tile - half the shared memory, eight stages instead of three - was measured here:
  // 4.1 instead of 2.9 ms per level on 100M x 64; its three extra ALU instructions per increment cost more
  // than the stages gain, the ALU pipe being the busiest unit of this phase)
  constexpr bool kPacked = false;
  const MoveSmem ms = move_smem(smem, move_geom(FT, B, Cn, kPacked));
  uint32_t* tile = ms.tile;
  const int stages = ms.stages;
  const int n_slots = vp.n_slots();
  const int S = 16 / FT;
  const int GS = (int)(P.G * S);
  const int tiles = step == 0 ? 1 : GS - 1;
  const int total_units = vp.scalars()[4] * tiles;
  const int u0 = (int)((int64_t)total_units * blockIdx.x / gridDim.x);
  const int u1 = (int)((int64_t)total_units * (blockIdx.x + 1) / gridDim.x);
  const int tid = threadIdx.x;

  const uint8_t* codes = vp.codes_cur();
  const uint8_t* lsrc = vp.labels_cur();
  uint32_t* meta = ptr_of<uint32_t>(P.mask_words);
  const int* rec_feature = ptr_of<const int>(P.rec_feature);
  MoveCursor cur;
  cur.init(P, n_slots, tiles);
  // feature tile of (slot, ordinal j): step 0 handles the split plane's first tile, step 1 skips it
  auto tile_of = [&](int slot, int j) -> int {
    const int gsA = (rec_feature[slot] >> 4) * S;
    return step == 0 ? gsA : (j < gsA ? j : j + 1);
  };

  if (tid >= kConsumers + 32) {
    // ===== ranker warps (step 0): "goes left" bits, prefixes and destination bases of each unit.
    // Ranker k takes the units u0 + k, u0 + k + kRankers, ... and follows the stage ring on its own.
    if (step != 0) return;
    const int lane = tid & 31, rk = (tid - kConsumers - 32) >> 5;
    const int* rec_bin = ptr_of<const int>(P.rec_bin);
    uint32_t* cursors = ptr_of<uint32_t>(P.cursors);
    // A ranker takes its bytes out of the staged rows and lets go of the stage at once: the consumers do
    // not need the unit record until they have counted the unit's rows (they form their own rows' "goes
    // left" bits themselves), so neither the ranker's atomic round trip nor its reads are part of a row
    // stage's cycle.  Records go through a ring of kRecRing entries.  (Measured per plane of the 100M-row
    // fit: 0.88 ms with the record on the stage's critical path, 0.83 ms with the rankers reading global
    // memory ahead of the copies, 0.75 ms as here, 0.67 ms for step B.  Letting the consumers post the bit
    // words of their own rows, so that a ranker only scans and reserves, took 1.20 ms: the record then
    // waits for the slowest consumer warp to reach the unit.)
    Ring mine_ring = ring;
    for (int i = 0; i < rk; ++i) mine_ring.next(stages);
    for (int u = u0 + rk; u < u1; u += kRankers) {
      cur.seek_to(u);
      const int i = u - u0, slot = i % kRecRing;
      const uint32_t use = (uint32_t)(i / kRecRing);
      const int f = rec_feature[cur.slot], bin = rec_bin[cur.slot];
      mbar_wait_sleepy(&ms.full[mine_ring.st], mine_ring.ph, 1000u);
      const uint32_t col = smem_u32(ms.stage_base) + (uint32_t)mine_ring.st * kStage2Bytes + (uint32_t)(f & 15);
      uint32_t code[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int row = j * 32 + lane;
        code[j] = row < cur.nrows ? lds_u8(col + 16u * (uint32_t)row) : 0xffffu;  // (0xffff: past the unit's end)
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ms.taken[mine_ring.st]);
      uint32_t mine = 0, nl = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const uint32_t ml = __ballot_sync(0xffffffffu, (int)code[j] <= bin);
        if (lane == j) mine = ml;
        nl += __popc(ml);
      }
      uint32_t base_l = 0, base_r = 0;
      if (lane == 0) {
        // one 64-bit add reserves both ranges (left count low, right count high)
        const unsigned long long got = atomicAdd(reinterpret_cast<unsigned long long*>(cursors) + cur.slot,
                                                 (unsigned long long)nl | ((unsigned long long)((uint32_t)cur.nrows - nl) << 32));
        base_l = (uint32_t)got;
        base_r = (uint32_t)(got >> 32);
      }
      const uint32_t pc = __popc(mine);
      const uint32_t before = warp_incl_scan(pc) - pc;  // lefts before word `lane`: ranks need no scan later
      mbar_wait_sleepy(&ms.rec_free[slot], (use & 1u) ^ 1u, 1000u);  // the consumers are done with the slot's last record
      uint32_t* rec = ms.rec_ring + slot * kMetaWords;
      rec[lane] = mine;
      rec[32 + lane] = before;
      if (lane == 0) { rec[64] = base_l; rec[65] = base_r; }
      if (GS > 1) {  // step B fetches the record by TMA
        uint32_t* grec = meta + (int64_t)cur.unit_index() * kMetaWords;
        grec[lane] = mine;
        grec[32 + lane] = before;
        if (lane == 0) { grec[64] = base_l; grec[65] = base_r; }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ms.rec_ready[slot]);
#pragma unroll
      for (int k = 0; k < kRankers; ++k) mine_ring.next(stages);
    }
    return;
  }
  if (tid >= kConsumers) {
    // ===== producer: rows of one plane, their labels (and in step B the unit record), all by TMA =====
    if (tid == kConsumers) {
      for (int u = u0; u < u1; ++u) {
        cur.seek(u);
        const int g = tile_of(cur.slot, cur.j) / S;
        mbar_wait_sleepy(&ms.empty[ring.st], ring.ph ^ 1, 1000u);
        if (step == 0) mbar_wait_sleepy(&ms.taken[ring.st], ring.ph ^ 1, 1000u);
        uint8_t* sbuf = ms.stage_base + (size_t)ring.st * kStage2Bytes;
        const uint32_t cbytes = (uint32_t)cur.nrows * 16u;
        const uint32_t loff = (uint32_t)cur.row0 & 15u;
        const uint32_t lbytes = (loff + (uint32_t)cur.nrows + 15u) & ~15u;
        mbar_arrive_expect_tx(&ms.full[ring.st], cbytes + lbytes + (step ? kMetaWords * 4 : 0));
        tma_load_1d(sbuf, codes + ((int64_t)g * P.n_cap + cur.row0) * 16, cbytes, &ms.full[ring.st]);
        tma_load_1d(sbuf + kStageBytes, lsrc + (cur.row0 - (int)loff), lbytes, &ms.full[ring.st]);
        if (step)
          tma_load_1d(sbuf + kStageBytes + kLabelStage, meta + (int64_t)cur.unit_index() * kMetaWords, kMetaWords * 4,
                      &ms.full[ring.st]);
        ring.next(stages);
      }
    }
    return;
  }
  // ===== consumer warps: warp w owns rows [64 w, 64 w + 64) of every unit =====
  const int lane = lane_id(), w = warp_id();
  uint8_t* ldst = vp.labels_nxt();
  uint4* dst = reinterpret_cast<uint4*>(vp.codes_nxt());
  const int* part_flags = ptr_of<const int>(P.part_flags);
  const int* part_child = ptr_of<const int>(P.part_child);
  const int* cmode = ptr_of<const int>(C.slot_mode);
  const int64_t plane = P.n_cap;
  uint64_t* avail = ms.full;
  int cur_slot = -1, cur_j = -1, cur_gs = -1, build_child = -1, build_side = -1, flags = 0;
  int split_k = 0, split_bin = 0;  // step A: the consumers test their own rows' bytes of the split feature
  const int* rec_bin_c = ptr_of<const int>(P.rec_bin);
  int units_since_flush = 0;
  bool cur_excl = false;
  for (int u = u0; u < u1; ++u) {
    cur.seek(u);
    if (cur.slot != cur_slot || cur.j != cur_j || (kPacked && units_since_flush >= kPackedFlushUnits)) {
      if (build_child >= 0) {
        named_bar_sync(1, kConsumers);
        if constexpr (kPacked) flush_tile_il16(C, tile, cur_gs / S, tid, build_child);
        else flush_tile<FT, IL>(C, tile, build_child, cur_gs, tid, cur_excl);
        named_bar_sync(1, kConsumers);
      }
      units_since_flush = 0;
      cur_slot = cur.slot;
      cur_j = cur.j;
      cur_gs = tile_of(cur.slot, cur.j);
      cur_excl = cur.owns_whole_tile(u0, u1);
      flags = part_flags[cur_slot];
      split_k = rec_feature[cur_slot] & 15;
      split_bin = rec_bin_c[cur_slot];
      const int a = part_child[cur_slot];
      build_child = -1;
      build_side = -1;
      if (a >= 0) {
        if (cmode[a] == MTB_MODE_BUILD) { build_child = a; build_side = 0; }
        else if (cmode[a + 1] == MTB_MODE_BUILD) { build_child = a + 1; build_side = 1; }
      }
    }
    const int g = cur_gs / S, sub = (cur_gs % S) * FT;
    const bool mover = (cur_gs % S) == 0;  // one feature tile per plane moves the plane's bytes
    const int ra = w * 64 + lane, rb = ra + 32;
    const bool va_ok = ra < cur.nrows, vb_ok = rb < cur.nrows;
    mbar_wait_sleepy(&avail[ring.st], ring.ph, 500u);
    // (explicit shared-space loads: see lds_u4)
    const uint32_t sbuf = smem_u32(ms.stage_base) + (uint32_t)ring.st * kStage2Bytes;
    const uint32_t slab = sbuf + kStageBytes + ((uint32_t)cur.row0 & 15u);
    // unit record: step B reads it from the stage (fetched by TMA); step A takes it from the rankers' ring
    // after the unit's rows are counted
    const int rslot = (u - u0) % kRecRing;
    uint32_t w0 = 0, w1 = 0, before0 = 0, base_l = 0, base_r = 0;
    auto read_record = [&](uint32_t rec) {
      w0 = lds_u32(rec + 8 * w);
      w1 = lds_u32(rec + 8 * w + 4);
      before0 = lds_u32(rec + 4 * (32 + 2 * w));  // lefts before word 2w
      base_l = lds_u32(rec + 4 * 64);
      base_r = lds_u32(rec + 4 * 65);
    };
    if (step != 0) read_record(sbuf + kStageBytes + kLabelStage);
    uint4 xa = make_uint4(0, 0, 0, 0), xb = xa;
    uint32_t lab_a = 0, lab_b = 0;
    if (va_ok) { xa = lds_u4(sbuf + 16 * ra); lab_a = lds_u8(slab + ra); }
    if (vb_ok) { xb = lds_u4(sbuf + 16 * rb); lab_b = lds_u8(slab + rb); }
    __syncwarp();
    if (lane == 0) mbar_arrive(&ms.empty[ring.st]);
    bool la, lb;
    if (step == 0) {
      la = (int)code_byte(xa, split_k) <= split_bin;
      lb = (int)code_byte(xb, split_k) <= split_bin;
    } else {
      la = (w0 >> lane) & 1u;
      lb = (w1 >> lane) & 1u;
    }
    auto move_rows = [&]() {
      if (!mover) return;
      const uint32_t below = (1u << lane) - 1u;
      const uint32_t lefts_a = before0 + __popc(w0 & below);
      const uint32_t lefts_b = before0 + __popc(w0) + __popc(w1 & below);
      const bool keep_l = flags & 1, keep_r = flags & 2;
      if (va_ok && (la ? keep_l : keep_r)) {
        const int64_t d = la ? (int64_t)cur.begin + base_l + lefts_a
                             : (int64_t)cur.end - 1 - (base_r + ((uint32_t)ra - lefts_a));
        st_na_u4(dst + (int64_t)g * plane + d, xa);
        if (g == 0) ldst[d] = (uint8_t)lab_a;
      }
      if (vb_ok && (lb ? keep_l : keep_r)) {
        const int64_t d = lb ? (int64_t)cur.begin + base_l + lefts_b
                             : (int64_t)cur.end - 1 - (base_r + ((uint32_t)rb - lefts_b));
        st_na_u4(dst + (int64_t)g * plane + d, xb);
        if (g == 0) ldst[d] = (uint8_t)lab_b;
      }
    };
    if (step != 0) move_rows();
    // (Tried: compacting the rows to count - the lanes list their rows of the built child in shared memory
    // and lane i reads the i-th listed row from the stage again, so that a warp issues one full pass of
    // increments instead of two half-empty ones.  2.67 instead of 2.50 ms per level: with near-even splits
    // a second pass is still needed for almost half of the units, and full passes conflict more.  Also
    // tried: first pass = a lane's first row that counts, second pass = lanes with two; 2.57 ms.  The
    // cost of the shared-memory increments follows the number of active lanes, not of instructions.)
    if (build_child >= 0) {
      ++units_since_flush;
      if constexpr (IL && kPacked) {
        if (va_ok && (int)(!la) == build_side) tile_count_il16(tile, xa, lab_a, B, lane);
        if (vb_ok && (int)(!lb) == build_side) tile_count_il16(tile, xb, lab_b, B, lane);
      } else if constexpr (IL) {
        if (va_ok && (int)(!la) == build_side) tile_count_il(tile, xa, lab_a, B, lane);
        if (vb_ok && (int)(!lb) == build_side) tile_count_il(tile, xb, lab_b, B, lane);
      } else {
        if (va_ok && (int)(!la) == build_side) {
#pragma unroll
          for (int k = 0; k < FT; ++k) atomicAdd(&tile[(k * Cn + (int)lab_a) * B + (int)code_byte(xa, sub + k)], 1u);
        }
        if (vb_ok && (int)(!lb) == build_side) {
#pragma unroll
          for (int k = 0; k < FT; ++k) atomicAdd(&tile[(k * Cn + (int)lab_b) * B + (int)code_byte(xb, sub + k)], 1u);
        }
      }
    }
    if (step == 0) {
      mbar_wait_sleepy(&ms.rec_ready[rslot], (uint32_t)((u - u0) / kRecRing) & 1u, 500u);
      read_record(smem_u32(ms.rec_ring) + (uint32_t)rslot * kMetaWords * 4);
      __syncwarp();
      if (lane == 0) mbar_arrive(&ms.rec_free[rslot]);
      move_rows();
    }
    ring.next(stages);
  }
  if (build_child >= 0) {
    named_bar_sync(1, kConsumers);
    if constexpr (kPacked) flush_tile_il16(C, tile, cur_gs / S, tid, build_child);
    else flush_tile<FT, IL>(C, tile, build_child, cur_gs, tid, cur_excl);
  }
}

// ----------------------------------------------------------------- fused level
// plan -> hist -> [push partial slabs to owners -> peer barrier] -> gain on owned features
// (records go to every rank's mailbox) -> [peer barrier] -> finalize.  One cooperative
// launch; with world > 1 no NCCL call and no separate reduction or gain kernel runs.
__device__ __forceinline__ void stamp(const MtbCtx& ctx, int i) {
  // optional phase timestamps (ns, %globaltimer) for profiling: ctx.timers = u64[16] or 0
  if (ctx.timers && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    ptr_of<unsigned long long>(ctx.timers)[i] = t;
  }
}

// Grid-wide zero fill of `n_slabs` histogram slabs starting at `base` (128-bit stores).
__device__ void zero_slabs(uint32_t* base, int64_t slab_words, int n_slabs) {
  uint4* p = reinterpret_cast<uint4*>(base);
  const int64_t n16 = slab_words * n_slabs / 4;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
    st_na_u4(p + i, z);
}

// After a level's search: the slabs the next level accumulates into must start at zero.  They are
// the other ping-pong side (dead once this level's DERIVE pairs have read their parents); with
// peers the one side is reused, dead once every rank has passed the second barrier.  The next
// frontier has at most two slots per current slot.  Runs while CTA 0 finalizes.
__device__ void zero_next_level(const MtbCtx& ctx, int n_slots) {
  if (ctx.max_depth >= 0 && ctx.depth + 1 >= ctx.max_depth) return;
  View v(ctx);
  int64_t n = 2 * (int64_t)n_slots;
  if (n > ctx.h_cap) n = ctx.h_cap;
  zero_slabs(ptr_of<uint32_t>(ctx.hist[ctx.hcur ^ 1]), v.slab_words(), (int)n);
}

// ------------------------------------------------------ root level from row-major bytes (cluster)
// The first touch of a uint8 data set: rows arrive row-major ([n][16 G] bytes, the user's layout) and
// have to become feature-group planes, labels have to be copied, and the root histogram has to be
// counted.  As three passes (retile kernel, label copy, root histogram) that is 6.4 + 6.4 + 6.4 GB of
// traffic for 100M x 64; here it is one pass, 6.4 GB in and 6.4 GB out, run by thread-block clusters
// of G CTAs:
//   * a unit is 256 whole rows (256 x 16 G bytes, contiguous).  It is loaded ONCE per cluster, by TMA
//     bulk copies that are multicast into the same stage of all G CTAs (the source is this GPU's own
//     memory, see tma_load_1d_multicast);
//   * CTA g of the cluster owns feature group g: its consumers take bytes [16 g, 16 g + 16) of every
//     row from the shared tile, store them into plane g (one 128-bit store per row, coalesced), and
//     count them into the CTA's [C*B][16] histogram tile; CTA 0 also writes the labels;
//   * two consumer groups of eight warps alternate units, so stores and shared-memory atomics of
//     consecutive units overlap;
//   * a stage holds the same bytes in all G CTAs, so the bytes a cluster has in flight are bounded by
//     ONE CTA's spare shared memory: with the 32-bit tile (160 KB at 256 bins x 10 classes) only four
//     16 KB stages fit and the kernel was latency-bound at 11 ms for 100M x 64.  The tile therefore
//     packs two 16-bit counters per word (80 KB, nine stages) and is merged into the slab every 240
//     units, before any counter can overflow;
//   * stage hand-shake: a CTA's producer thread waits until its own consumers have released the
//     stage, arms its own `full` barrier with the unit's byte count and arrives on the leader's `go`
//     barrier; the leader issues the multicast copy once all G CTAs have done so.  The copy is the
//     tiled (tensor-map) form: the untiled cp.async.bulk with .multicast::cluster was measured at
//     ~0.34 us + 0.043 us/KB per operation, serialised across the cluster (23 GB/s per cluster).
// The tiles are merged into the root slab with red.global.add at the end (the entry point zeroes the
// slab first).  mtb_level_fused then runs with ctx.root_hist_done set and skips its histogram phase.
constexpr int kRootRows = 256;
constexpr int kRootGroupWarps = 8;

struct RootArgs {
  const uint8_t* X;   // [n][ldx] row-major bytes, ldx == 16 * G, 16-byte aligned, this GPU's memory
  const uint8_t* y;   // [n] labels (class codes)
  int64_t n;
  int multicast;      // 1: the leader loads each unit once for the whole cluster; 0: every CTA loads it itself (experiment)
};

struct RootSmem {
  uint32_t* tile;
  uint8_t* stage_base;
  uint64_t *full, *empty, *go;
  int stages, stage_bytes;
};

__host__ __device__ inline int root_stage_bytes(int G) { return kRootRows * 16 * G; }
constexpr int kRootMaxStages = 12;
constexpr int kRootFlushUnits = 240;  // 240 x 256 rows < 65,536: no 16-bit counter can overflow between flushes
__host__ __device__ inline int root_tile_bytes(int B, int C) { return 16 * B * C * 2; }  // two counters per word
__host__ __device__ inline int root_stages(int B, int C, int G) {
  int st = (kSmemBudget - root_tile_bytes(B, C) - 512) / root_stage_bytes(G);
  return st > kRootMaxStages ? kRootMaxStages : st;
}

__device__ __forceinline__ RootSmem root_smem(uint8_t* smem, int B, int C, int G) {
  RootSmem m;
  m.tile = reinterpret_cast<uint32_t*>(smem);
  m.stage_bytes = root_stage_bytes(G);
  m.stages = root_stages(B, C, G);
  m.stage_base = smem + (size_t)root_tile_bytes(B, C);
  m.full = reinterpret_cast<uint64_t*>(m.stage_base + (size_t)m.stages * m.stage_bytes);
  m.empty = m.full + kRootMaxStages;
  m.go = m.empty + kRootMaxStages;
  return m;
}

__global__ void __launch_bounds__(kHistThreads, 1) k_root_cluster(const __grid_constant__ MtbCtx ctx, const RootArgs a,
                                                                  const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int B = (int)ctx.B, C = (int)ctx.C, G = (int)ctx.G;
  const RootSmem ms = root_smem(smem, B, C, G);
  const int tid = threadIdx.x;
  const uint32_t crank = cluster_ctarank();  // == feature group of this CTA
  const int n_clusters = gridDim.x / G, cluster_id = blockIdx.x / G;
  const int64_t units = (a.n + kRootRows - 1) / kRootRows;
  const int64_t u0 = units * cluster_id / n_clusters, u1 = units * (cluster_id + 1) / n_clusters;
  if (tid == 0) {
    for (int i = 0; i < ms.stages; ++i) {
      mbar_init(&ms.full[i], 1);
      mbar_init(&ms.empty[i], kRootGroupWarps);
      mbar_init(&ms.go[i], (uint32_t)G);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < 8 * B * C; i += blockDim.x) ms.tile[i] = 0;
  __syncthreads();
  cluster_sync_all();  // every CTA's barriers exist before anybody arrives on them remotely

  if (tid >= kConsumers) {
    // ===== producer thread of this CTA =====
    if (tid == kConsumers) {
      int st = 0;
      uint32_t ph = 0;
      for (int64_t u = u0; u < u1; ++u) {
        const int64_t row0 = u * kRootRows;
        uint32_t bytes = (uint32_t)min((int64_t)kRootRows, a.n - row0) * 16u * (uint32_t)G;
        if (a.multicast) bytes = (uint32_t)ms.stage_bytes;  // the tiled copy always delivers the whole box
        mbar_wait(&ms.empty[st], ph ^ 1);                // own consumers are done with the stage
        mbar_arrive_expect_tx(&ms.full[st], bytes);      // armed for the unit's bytes
        if (!a.multicast) {
          tma_load_1d(ms.stage_base + (size_t)st * ms.stage_bytes, a.X + row0 * 16 * G, bytes, &ms.full[st]);
          if (++st == ms.stages) { st = 0; ph ^= 1; }
          continue;
        }
        mbar_arrive_cluster(&ms.go[st], 0);              // tell the leader
        if (crank == 0) {
          mbar_wait_cluster(&ms.go[st], ph);             // all G CTAs are ready to receive
          // one tiled copy of the unit (a 16 G x 256 box of the [n][16 G] byte matrix), multicast to the
          // cluster; rows past the end arrive as zeros, so the byte count is always the full box
          tma_load_2d_multicast(ms.stage_base + (size_t)st * ms.stage_bytes, &tmap, 0, (int)row0, &ms.full[st],
                                (uint16_t)((1u << G) - 1u));
        }
        if (++st == ms.stages) { st = 0; ph ^= 1; }
      }
    }
  } else {
    // ===== consumers: group `grp` takes the units u0 + grp, u0 + grp + 2, ... =====
    const int lane = tid & 31, w = tid >> 5;
    const int grp = w / kRootGroupWarps, r = (w % kRootGroupWarps) * 32 + lane;  // row of the unit
    uint4* plane = reinterpret_cast<uint4*>(ptr_of<uint8_t>(ctx.codes[ctx.cur])) + (int64_t)crank * ctx.n_cap;
    uint8_t* lab_out = ptr_of<uint8_t>(ctx.labels[ctx.cur]);
    int st = grp % ms.stages;
    uint32_t ph = (uint32_t)(grp / ms.stages) & 1u;
    int64_t next_flush = u0 + kRootFlushUnits;
    // labels come straight from global memory, one unit ahead (a unit is short: the load's latency
    // would otherwise be paid once per unit)
    uint32_t lab_next = 0;
    if (u0 + grp < u1 && (u0 + grp) * kRootRows + r < a.n) lab_next = a.y[(u0 + grp) * kRootRows + r];
    for (int64_t u = u0 + grp; u < u1; u += 2) {
      if (u >= next_flush) {  // both groups pass every flush point, each before its first unit beyond it
        named_bar_sync(1, kConsumers);
        flush_tile_il16(ctx, ms.tile, (int)crank, tid);
        named_bar_sync(1, kConsumers);
        next_flush += kRootFlushUnits;
      }
      const int64_t row = u * kRootRows + r;
      const bool ok = row < a.n;
      const uint32_t lab = lab_next;
      if (u + 2 < u1 && row + 2 * kRootRows < a.n) lab_next = a.y[row + 2 * kRootRows];
      mbar_wait(&ms.full[st], ph);
      uint4 x = make_uint4(0, 0, 0, 0);
      if (ok) x = lds_u4(smem_u32(ms.stage_base) + (uint32_t)(st * ms.stage_bytes + r * 16 * G + 16 * (int)crank));
      __syncwarp();
      if (lane == 0) mbar_arrive(&ms.empty[st]);  // the tile's bytes are in registers: release the stage
      if (ok) {
        st_na_u4(plane + row, x);
        if (crank == 0) lab_out[row] = (uint8_t)lab;
        tile_count_il16(ms.tile, x, lab, B, lane);
      }
      // two units further on: the ring position advances by two stages
      st += 2;
      if (st >= ms.stages) { st -= ms.stages; ph ^= 1; }
    }
    // flush points this group did not reach because its units ran out (the other group may have)
    for (; next_flush < u1; next_flush += kRootFlushUnits) {
      named_bar_sync(1, kConsumers);
      flush_tile_il16(ctx, ms.tile, (int)crank, tid);
      named_bar_sync(1, kConsumers);
    }
    named_bar_sync(1, kConsumers);
    flush_tile_il16(ctx, ms.tile, (int)crank, tid);
  }
  cluster_sync_all();  // nobody leaves while a peer may still multicast into it or arrive on its barriers
}

// Root level of row-major uint8 input without staging (k_root_direct): CTA b works on feature group
// b mod G of the rows of row range b / G.  A lane loads the 16 bytes of its row's group straight from
// the caller's matrix (128-bit loads, four rows per lane in flight), stores them into the plane,
// and counts them into the CTA's 32-bit [C*B][16] tile; the CTAs of group 0 also write the labels.
// The G CTAs of a row range read the same lines at about the same time, so DRAM sees the input once
// (the others hit in L2); each CTA uses half of every 32-byte sector it receives.  Against the
// separate passes (retile: read + write; root histogram: read the planes again) one read of the
// data set is saved.  The tiles are merged into the root slab with red.global.add (the entry point
// zeroes the slab first); mtb_level_fused then runs with ctx.root_hist_done set.
constexpr int kRootDirectThreads = 1024;
constexpr int kRootDirectUnroll = 4;

__global__ void __launch_bounds__(kRootDirectThreads, 1) k_root_direct(const __grid_constant__ MtbCtx ctx, const RootArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint32_t* tile = reinterpret_cast<uint32_t*>(smem);
  const int B = (int)ctx.B, C = (int)ctx.C, G = (int)ctx.G;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, nw = kRootDirectThreads / 32;
  const int g = (int)(blockIdx.x % (unsigned)G), range = (int)(blockIdx.x / (unsigned)G), n_ranges = (int)(gridDim.x / (unsigned)G);
  for (int i = tid; i < 16 * B * C; i += kRootDirectThreads) tile[i] = 0;
  __syncthreads();
  const int64_t blocks = (a.n + 31) / 32;  // 32 rows: one per lane
  const int64_t b0 = blocks * range / n_ranges, b1 = blocks * (range + 1) / n_ranges;
  const uint4* src = reinterpret_cast<const uint4*>(a.X) + g;  // row r of the group: src[r * G]
  uint4* plane = reinterpret_cast<uint4*>(ptr_of<uint8_t>(ctx.codes[ctx.cur])) + (int64_t)g * ctx.n_cap;
  uint8_t* lab_out = ptr_of<uint8_t>(ctx.labels[ctx.cur]);
  for (int64_t b = b0 + w; b < b1; b += (int64_t)nw * kRootDirectUnroll) {
    uint4 x[kRootDirectUnroll];
    uint32_t lab[kRootDirectUnroll];
#pragma unroll
    for (int j = 0; j < kRootDirectUnroll; ++j) {
      const int64_t row = (b + (int64_t)j * nw) * 32 + lane;
      x[j] = make_uint4(0, 0, 0, 0);
      lab[j] = 0;
      if (b + (int64_t)j * nw < b1 && row < a.n) {
        x[j] = ld_cg_u4(src + row * G);
        lab[j] = __ldg(a.y + row);
      }
    }
#pragma unroll
    for (int j = 0; j < kRootDirectUnroll; ++j) {
      const int64_t row = (b + (int64_t)j * nw) * 32 + lane;
      if (b + (int64_t)j * nw < b1 && row < a.n) {
        st_na_u4(plane + row, x[j]);
        if (g == 0) lab_out[row] = (uint8_t)lab[j];
        tile_count_il(tile, x[j], lab[j], B, lane);
      }
    }
  }
  __syncthreads();
  if (tid < kConsumers) flush_tile<16, true>(ctx, tile, 0, g, tid, false);
}

// Everything after a level's histograms are complete on this rank (the caller has synchronised the
// grid): [push partial slabs to the owners, peer barrier,] gain on the owned features (records go to
// every rank's mailbox), [peer barrier | grid sync,] finalize on CTA 0 while the others zero the
// next level's slabs.
template <bool NV>
__device__ __forceinline__ void search_tail(const MtbCtx& ctx, cg::grid_group& grid, uint8_t* smem, int smem_bytes,
                                            int* smem_warp, int n_slots, bool zero_next, bool rows_from_cursors) {
  stamp(ctx, 2);
  if (ctx.world > 1) {
    if constexpr (!NV) push_phase(ctx, n_slots, rows_from_cursors);
    peer_barrier(ctx, grid, (uint32_t)ctx.epoch + 1u);
    stamp(ctx, 3);
    gain_phase<NV>(ctx, smem, kFusedGainWarps, smem_bytes, -1, n_slots);
    stamp(ctx, 4);
    peer_barrier(ctx, grid, (uint32_t)ctx.epoch + 2u);
    stamp(ctx, 5);
  } else {
    stamp(ctx, 3);
    gain_phase(ctx, smem, kFusedGainWarps, smem_bytes, -1, n_slots);
    grid.sync();
    stamp(ctx, 4);
    stamp(ctx, 5);
  }
  if (blockIdx.x == 0) finalize_phase(ctx, smem_warp, reinterpret_cast<const double*>(smem));
  stamp(ctx, 6);
  if (zero_next) zero_next_level(ctx, n_slots);
}

template <int FT, bool IL = false, bool NV = false>
__global__ void __launch_bounds__(kFusedThreads, 1) k_level_fused(const __grid_constant__ MtbCtx ctx, int smem_bytes) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int smem_warp[33];
  cg::grid_group grid = cg::this_grid();
  // empty frontier (the tree finished at an earlier level): uniform early exit on every CTA of
  // every rank, before any grid or peer barrier.  (Depth 0 starts a fit: CTA 0 writes the root slot.)
  if (ctx.depth > 0 && ptr_of<const int>(ctx.scalars)[0] == 0) return;
  stamp(ctx, 0);
  if (blockIdx.x == 0) plan_phase(ctx, smem_warp);
  const bool have_hist = ctx.depth == 0 && ctx.root_hist_done != 0;  // counted by k_root_cluster
  if (!have_hist) {  // the slabs this level accumulates into start at zero (shared tiles are merged with red.add)
    View v(ctx);
    zero_slabs(v.hist_cur(), v.slab_words(), ctx.depth == 0 ? 1 : v.n_slots());
  }
  grid.sync();
  stamp(ctx, 1);
  if (!have_hist) {
    hist_phase<FT, IL>(ctx, smem);
    grid.sync();
  }
  search_tail<NV>(ctx, grid, smem, smem_bytes, smem_warp, ptr_of<const int>(ctx.scalars)[0], ctx.zero_next != 0, false);
}

template <int FT, bool IL = false, bool NV = false>
__global__ void __launch_bounds__(kFusedThreads, 1) k_level_fused2(const __grid_constant__ MtbCtx P,
                                                                   const __grid_constant__ MtbCtx C, int smem_bytes) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int smem_warp[33];
  cg::grid_group grid = cg::this_grid();
  // nothing to partition and nothing to search: the tree finished at an earlier level
  const int n_children = ptr_of<const int>(P.scalars)[1];  // the children's slot count (finalize overwrites it at the end)
  if (n_children == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) ptr_of<int>(P.scalars)[0] = 0;
    return;
  }
  stamp(C, 0);
  {
    const HistGeom geom = move_geom(FT, (int)P.B, (int)P.C, false);
    const MoveSmem ms = move_smem(smem, geom);
    if (threadIdx.x == 0) {
      for (int i = 0; i < geom.stages; ++i) {
        mbar_init(&ms.full[i], 1);
        mbar_init(&ms.empty[i], kConsumerWarps);
        mbar_init(&ms.taken[i], 1);
      }
      for (int i = 0; i < kRecRing; ++i) {
        mbar_init(&ms.rec_ready[i], 1);
        mbar_init(&ms.rec_free[i], kConsumerWarps);
      }
      fence_barrier_init();
    }
    for (int i = threadIdx.x; i < geom.tile_words; i += blockDim.x) ms.tile[i] = 0;
    __syncthreads();
  }
  Ring ring{0, 0};
  move_hist_phase<FT, IL>(P, C, smem, 0, ring);
  if (P.G * (16 / FT) > 1) {
    fence_proxy_async();  // the ranker's generic-proxy writes precede step B's TMA writes to the stages
    grid.sync();          // unit records (global) and destination bases are complete
    stamp(C, 8);
    move_hist_phase<FT, IL>(P, C, smem, 1, ring);
  }
  grid.sync();  // children's histograms and rows are complete on this rank
  stamp(C, 1);
  // CTA 0 turns the partition cursors into the children's row segments while the others go on: only
  // its own finalize (and the next launch) read them.  The search below takes the slot count as an
  // argument; scalars[0] is brought up to date for finalize and the zeroing at the end.
  if (blockIdx.x == 0) {
    commit_phase(P, threadIdx.x, blockDim.x);
    if (threadIdx.x == 0) ptr_of<int>(P.scalars)[0] = n_children;
    __syncthreads();
  }
  search_tail<NV>(C, grid, smem, smem_bytes, smem_warp, n_children, true, true);
}

int g_sm_count = 0;
int sm_count() {
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
  }
  return g_sm_count;
}

template <typename K>
int set_smem(K kernel, int bytes) {
  MTB_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

// unfused gain kernel: as many warps as fit in shared memory with their slabs
int gain_smem(const MtbCtx* ctx) {
  GainGeom g = gain_geom((int)ctx->B, (int)ctx->C, kGainWarps, 227 * 1024);
  return g.warps < 1 ? -1 : kGainTabBytes + g.warps * g.slab_words * 4;
}

}  // namespace

// ctx->tile_layout == 1 selects the interleaved shared-memory tile (FT == 16 only).
static bool interleaved(const MtbCtx* ctx) { return ctx->tile_layout == 1 && ctx->FT == 16; }
// ctx->mc_hist set (world > 1): the histogram slabs are multicast-mapped and the owners reduce in the switch.
static bool in_switch(const MtbCtx* ctx) { return ctx->world > 1 && ctx->mc_hist[ctx->hcur] != 0; }

template <int FT>
static int launch_hist(const MtbCtx* ctx, const HistGeom& g, cudaStream_t stream) {
  if constexpr (FT == 16) {
    if (interleaved(ctx)) {
      if (set_smem(k_hist<16, true>, g.smem_bytes)) return -1;
      k_hist<16, true><<<sm_count(), kHistThreads, g.smem_bytes, stream>>>(*ctx);
      MTB_CHECK(cudaGetLastError());
      return 0;
    }
  }
  if (set_smem(k_hist<FT>, g.smem_bytes)) return -1;
  k_hist<FT><<<sm_count(), kHistThreads, g.smem_bytes, stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

template <int FT>
static int launch_fused(const MtbCtx* ctx, int smem, cudaStream_t stream) {
  MtbCtx local = *ctx;
  void* args[] = {&local, &smem};
  void* fn = (void*)k_level_fused<FT>;
  if constexpr (FT == 16) {
    const bool il = interleaved(ctx), nv = in_switch(ctx);
    if (il || nv)
      fn = il ? (nv ? (void*)k_level_fused<16, true, true> : (void*)k_level_fused<16, true, false>)
              : (void*)k_level_fused<16, false, true>;
  } else if (in_switch(ctx)) {
    mtb_set_error("fused", "the in-switch reduction is only built for 16-feature tiles", __FILE__, __LINE__);
    return -2;
  }
  MTB_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  MTB_CHECK(cudaLaunchCooperativeKernel(fn, dim3(sm_count()), dim3(kFusedThreads), args, smem, stream));
  return 0;
}

template <int FT>
static int launch_fused2(const MtbCtx* P, const MtbCtx* C, int smem, cudaStream_t stream) {
  MtbCtx p = *P, c = *C;
  void* args[] = {&p, &c, &smem};
  void* fn = (void*)k_level_fused2<FT>;
  if constexpr (FT == 16) {
    const bool il = interleaved(P), nv = in_switch(C);
    if (il || nv)
      fn = il ? (nv ? (void*)k_level_fused2<16, true, true> : (void*)k_level_fused2<16, true, false>)
              : (void*)k_level_fused2<16, false, true>;
  } else if (in_switch(C)) {
    mtb_set_error("fused2", "the in-switch reduction is only built for 16-feature tiles", __FILE__, __LINE__);
    return -2;
  }
  MTB_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  MTB_CHECK(cudaLaunchCooperativeKernel(fn, dim3(sm_count()), dim3(kFusedThreads), args, smem, stream));
  return 0;
}

#define FT_SWITCH(ft, FN, ...)                                                                         \
  switch (ft) {                                                                                        \
    case 16: return FN<16>(__VA_ARGS__);                                                               \
    case 8: return FN<8>(__VA_ARGS__);                                                                 \
    case 4: return FN<4>(__VA_ARGS__);                                                                 \
    case 2: return FN<2>(__VA_ARGS__);                                                                 \
    case 1: return FN<1>(__VA_ARGS__);                                                                 \
    default:                                                                                           \
      mtb_set_error("FT", "features per tile must be 1, 2, 4, 8 or 16", __FILE__, __LINE__);           \
      return -2;                                                                                       \
  }

// Shared memory of the fused kernels: the larger of the histogram / move pipelines, grown if needed so
// that at least one gain warp's slab fits.  -1: the shape does not fit (no FT helps: the gain slab is B x C).
static int fused_smem(const MtbCtx* ctx, bool with_move) {
  HistGeom g = hist_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C);
  int smem = g.smem_bytes;
  if (with_move) {
    HistGeom mg = move_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C, false);
    if (mg.smem_bytes > smem) smem = mg.smem_bytes;
  }
  GainGeom gg = gain_geom((int)ctx->B, (int)ctx->C, kFusedGainWarps, smem);
  if (gg.warps < 1) smem = kGainTabBytes + gg.slab_words * 4;
  return smem <= 227 * 1024 ? smem : -1;
}

extern "C" int mtb_hist_config(const MtbCtx* ctx, int* threads, int* smem_bytes, int* stages, int* rows_per_stage) {
  HistGeom g = hist_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C);
  if (threads) *threads = kFusedThreads;
  if (smem_bytes) *smem_bytes = g.smem_bytes;
  if (stages) *stages = g.stages;
  if (rows_per_stage) *rows_per_stage = kRows;
  if (g.stages < 2 || move_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C).stages < 2) return -2;
  return fused_smem(ctx, true) > 0 && gain_smem(ctx) > 0 ? 0 : -2;
}

extern "C" int mtb_level_plan(const MtbCtx* ctx, void* stream) {
  k_plan<<<1, 1024, 0, (cudaStream_t)stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_hist(const MtbCtx* ctx, void* stream) {
  HistGeom g = hist_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C);
  if (g.stages < 2) {
    mtb_set_error("hist", "histogram tile does not fit in shared memory", __FILE__, __LINE__);
    return -2;
  }
  FT_SWITCH((int)ctx->FT, launch_hist, ctx, g, (cudaStream_t)stream);
}

extern "C" int mtb_level_gain(const MtbCtx* ctx, void* stream) {
  const int smem = gain_smem(ctx);
  if (smem < 0) {
    mtb_set_error("gain", "bins x classes too large for a shared-memory gain slab", __FILE__, __LINE__);
    return -2;
  }
  if (set_smem(k_gain, smem)) return -1;
  const GainGeom gg = gain_geom((int)ctx->B, (int)ctx->C, kGainWarps, smem);
  int64_t pairs = ctx->n_slots_host * (ctx->world > 1 ? ctx->fpo : ctx->F);
  int blocks = (int)((pairs + gg.warps - 1) / gg.warps);
  int maxb = sm_count() * 2;
  if (blocks > maxb) blocks = maxb;
  if (blocks < 1) blocks = 1;
  k_gain<<<blocks, kGainWarps * 32, smem, (cudaStream_t)stream>>>(*ctx, -1, smem);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_push(const MtbCtx* ctx, void* stream) {
  if (ctx->world <= 1) return 0;
  k_push<<<sm_count() * 2, 512, 0, (cudaStream_t)stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_finalize(const MtbCtx* ctx, void* stream) {
  k_finalize<<<1, 1024, 0, (cudaStream_t)stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_partition(const MtbCtx* ctx, void* stream) {
  k_partition<<<sm_count() * 4, kPartThreads, 0, (cudaStream_t)stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_commit(const MtbCtx* ctx, void* stream) {
  k_commit<<<64, 256, 0, (cudaStream_t)stream>>>(*ctx);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int mtb_level_fused(const MtbCtx* ctx, void* stream) {
  if (hist_geom((int)ctx->FT, (int)ctx->B, (int)ctx->C).stages < 2) {
    mtb_set_error("hist", "histogram tile does not fit in shared memory", __FILE__, __LINE__);
    return -2;
  }
  const int smem = fused_smem(ctx, false);
  if (smem < 0) {
    mtb_set_error("fused", "gain slab does not fit in shared memory", __FILE__, __LINE__);
    return -2;
  }
  FT_SWITCH((int)ctx->FT, launch_fused, ctx, smem, (cudaStream_t)stream);
}

extern "C" int mtb_level_fused2(const MtbCtx* parents, const MtbCtx* children, void* stream) {
  if (hist_geom((int)parents->FT, (int)parents->B, (int)parents->C).stages < 2) {
    mtb_set_error("hist", "histogram tile does not fit in shared memory", __FILE__, __LINE__);
    return -2;
  }
  if (move_geom((int)parents->FT, (int)parents->B, (int)parents->C).stages < 2) {
    mtb_set_error("fused2", "histogram tile leaves no room for two row stages", __FILE__, __LINE__);
    return -3;
  }
  const int smem = fused_smem(parents, true);
  if (smem < 0) {
    mtb_set_error("fused", "gain slab does not fit in shared memory", __FILE__, __LINE__);
    return -2;
  }
  FT_SWITCH((int)parents->FT, launch_fused2, parents, children, smem, (cudaStream_t)stream);
}

/* Root level of a uint8 data set in one pass: row-major bytes -> feature-group planes + labels + root
 * histogram (k_root_cluster).  Returns 1 (nothing launched) when the shape or the pointers do not qualify:
 * the caller then encodes and counts the usual way.                                              */
extern "C" int mtb_root_retile_hist(const MtbCtx* ctx, const void* X, int64_t ldx, int64_t n, const void* y, void* stream) {
  const int G = (int)ctx->G, B = (int)ctx->B, C = (int)ctx->C;
  if (!(G == 2 || G == 4 || G == 8) || ctx->F != 16 * G || ldx != 16 * G || ctx->FT != 16 || ctx->tile_layout != 1) return 1;
  if ((reinterpret_cast<uintptr_t>(X) & 15u) != 0 || n <= 0 || n > ctx->n_cap) return 1;
  if (root_stages(B, C, G) < 2) return 1;
  // the multicast copy must read this GPU's own memory (see tma_load_1d_multicast)
  int dev = -1;
  cudaPointerAttributes attr;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaPointerGetAttributes(&attr, X) != cudaSuccess ||
      attr.type != cudaMemoryTypeDevice || attr.device != dev) {
    cudaGetLastError();
    return 1;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int smem = root_tile_bytes(B, C) + root_stages(B, C, G) * root_stage_bytes(G) + 512;
  MTB_CHECK(cudaFuncSetAttribute(k_root_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  MTB_CHECK(cudaMemsetAsync(ptr_of<void>(ctx->hist[ctx->hcur]), 0, (size_t)ctx->G * 16 * B * C * 4, st));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((sm_count() / G) * G));
  cfg.blockDim = dim3(kHistThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)G;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  // one wave: as many clusters as can be resident at once (a cluster needs G SMs of one GPC)
  static int max_clusters[9] = {0};
  if (!max_clusters[G]) {
    int nc = 0;
    if (cudaOccupancyMaxActiveClusters(&nc, k_root_cluster, &cfg) != cudaSuccess || nc < 1) {
      cudaGetLastError();
      nc = sm_count() / G;
    }
    max_clusters[G] = nc;
  }
  if ((unsigned)(max_clusters[G] * G) < cfg.gridDim.x) cfg.gridDim = dim3((unsigned)(max_clusters[G] * G));
  // tensor map of the row-major byte matrix: [n][16 G] uint8, box = 16 G x 256 rows
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
      cudaGetLastError();
      return 1;
    }
    encode = (EncodeFn)fn;
  }
  CUtensorMap tmap;
  {
    const cuuint64_t gdim[2] = {(cuuint64_t)(16 * G), (cuuint64_t)n};
    const cuuint64_t gstride[1] = {(cuuint64_t)(16 * G)};
    const cuuint32_t box[2] = {(cuuint32_t)(16 * G), (cuuint32_t)kRootRows};
    const cuuint32_t estride[2] = {1, 1};
    if (encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(X), gdim, gstride, box, estride,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 1;
  }
  static int multicast = -1;
  if (multicast < 0) {
    const char* e = getenv("MTB_ROOT_MULTICAST");
    multicast = (e && e[0] == '0') ? 0 : 1;
  }
  RootArgs a{reinterpret_cast<const uint8_t*>(X), reinterpret_cast<const uint8_t*>(y), n, multicast};
  MTB_CHECK(cudaLaunchKernelEx(&cfg, k_root_cluster, *ctx, a, tmap));
  return 0;
}

/* The same pass without staging and without clusters (k_root_direct): any whole number of 16-feature groups.  */
extern "C" int mtb_root_direct_hist(const MtbCtx* ctx, const void* X, int64_t ldx, int64_t n, const void* y, void* stream) {
  const int G = (int)ctx->G, B = (int)ctx->B, C = (int)ctx->C;
  if (G < 1 || G > sm_count() || ctx->F != 16 * G || ldx != 16 * G || ctx->FT != 16 || ctx->tile_layout != 1) return 1;
  if ((reinterpret_cast<uintptr_t>(X) & 15u) != 0 || n <= 0 || n > ctx->n_cap) return 1;
  const int smem = 16 * B * C * 4;
  if (smem > kSmemBudget) return 1;
  cudaStream_t st = (cudaStream_t)stream;
  MTB_CHECK(cudaFuncSetAttribute(k_root_direct, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  MTB_CHECK(cudaMemsetAsync(ptr_of<void>(ctx->hist[ctx->hcur]), 0, (size_t)G * 16 * B * C * 4, st));
  RootArgs a{reinterpret_cast<const uint8_t*>(X), reinterpret_cast<const uint8_t*>(y), n, 0};
  k_root_direct<<<(unsigned)((sm_count() / G) * G), kRootDirectThreads, smem, st>>>(*ctx, a);
  MTB_CHECK(cudaGetLastError());
  return 0;
}

/* Test hook: out[i] = g(k[i]) = k log2 k evaluated by the device function the gain phase uses
 * (one thread per value; `table` = the klogk constants).                                    */
namespace {
__global__ void k_klogk_eval(const uint32_t* k, double* out, int64_t n, const double* table) {
  KlogK g;
  g.bind(table, table);
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = g(k[i]);
}
}  // namespace
extern "C" int mtb_klogk_eval(const uint32_t* k, double* out, int64_t n, const double* table, void* stream) {
  if (n <= 0) return 0;
  k_klogk_eval<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(k, out, n, table);
  MTB_CHECK(cudaGetLastError());
  return 0;
}
