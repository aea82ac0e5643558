/*
 * mpitree_b200 native library — C ABI (libmpitree_b200.so, sm_100a only).
 *
 * Every entry point is `extern "C"`, takes raw device pointers and a CUDA stream
 * (as void*), launches asynchronously on that stream and returns 0 on success or
 * a negative error code (text via mtb_last_error()).  Nothing here allocates device
 * memory unless its comment says so: buffers are owned by the caller (the Python
 * layer allocates them as torch tensors).
 *
 * Reference hot paths these kernels replace (paths relative to the reference repo):
 *   K1 encode      np.unique candidate discovery      mpitree/tree/decision_tree.py:73
 *   K2 split search feature loop :133-138, threshold loop :76-86, entropy :49-51,
 *                   argmin/argmax :88-91 / :140-141
 *   K3 partition    boolean-mask copies               :150-164
 *   K4 predict      walk + apply_along_axis + argmax  :208-227, :248
 *   K5 all-reduce   (no counterpart; replaces the role of mpi4py collectives, :456)
 *   K6 hand-off     parity split + subtree allgather  :447-466
 */
#ifndef MPITREE_B200_H
#define MPITREE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTB_GROUP 16        /* features per 128-bit code chunk              */
#define MTB_MAX_BINS 256    /* bin codes are one byte                       */
#define MTB_MAX_CLASSES 128 /* CUDA engine limit (smem tile = B*C*4 bytes)  */
#define MTB_MAX_WORLD 16

/* slot modes */
#define MTB_MODE_NONE 0
#define MTB_MODE_BUILD 1
#define MTB_MODE_DERIVE 2
#define MTB_MODE_HANDOFF 3 /* rows are kept and listed for the subtree phase; no histogram */

/*
 * Build context.  Plain struct of 64-bit fields so that ctypes / any FFI can fill
 * it without worrying about padding.  Pointers are device pointers unless noted.
 */
typedef struct MtbCtx {
  /* ---- problem shape ---- */
  int64_t n_cap;    /* rows per feature-group plane (allocation stride)        */
  int64_t n_local;  /* rows of this rank's shard                               */
  int64_t F;        /* real features                                           */
  int64_t G;        /* feature groups = ceil(F/16)                             */
  int64_t B;        /* bin stride (power of two, 32..256)                      */
  int64_t C;        /* classes                                                 */
  int64_t FT;       /* features per shared-memory histogram tile (1,2,4,8,16)  */
  int64_t max_depth;          /* -1 = unlimited                                */
  int64_t min_samples_split;
  int64_t criterion;          /* 0 entropy, 1 gini                             */
  int64_t dot_mode;           /* host np.dot rounding model, see cpu_engine    */
  int64_t world, rank;
  int64_t m_cap;              /* frontier slot capacity                        */
  int64_t node_cap;           /* tree node capacity                            */
  /* ---- row data, ping-pong: [G][n_cap][16] codes, [n_cap] labels ---- */
  int64_t codes[2];
  int64_t labels[2];
  int64_t cur;                /* which ping-pong side holds the current level  */
  /* ---- histograms, ping-pong by level: [m_cap][G*16][C][B] u32 (class-major) ---- */
  int64_t hist[2];
  int64_t hcur;               /* side holding the current level's histograms   */
  /* ---- frontier (current level, SoA length m_cap) ---- */
  int64_t slot_node, slot_begin, slot_end, slot_mode, slot_parent, slot_sibling;
  /* ---- frontier being produced for the next level ---- */
  int64_t nslot_node, nslot_begin, nslot_end, nslot_mode, nslot_parent, nslot_sibling;
  int64_t nslot_src;          /* i32: parent slot * 2 + side (0 left, 1 right)   */
  int64_t n_slots_host;       /* host's copy of the current slot count (launch sizing) */
  /* ---- per (slot, feature) results of the gain phase ---- */
  int64_t fb_cost;   /* f64 [m_cap][F]   min cost (inf if no candidate)        */
  int64_t fb_bin;    /* i32 [m_cap][F]                                         */
  int64_t fb_left;   /* u32 [m_cap][F][C] left class counts at the best bin    */
  int64_t node_cnt;  /* u32 [m_cap][C]   class counts of the slot's node       */
  /* ---- per slot decisions ---- */
  int64_t rec_feature, rec_bin;  /* i32 [m_cap]                                 */
  int64_t part_flags;            /* i32 [m_cap] bit0: keep left, bit1: keep right */
  int64_t cursors;               /* u32 [m_cap][2] rows written left / right    */
  /* ---- tree (SoA length node_cap) ---- */
  int64_t t_feature, t_bin, t_left, t_parent, t_depth; /* i32 */
  int64_t t_gain;                /* f64 */
  int64_t t_counts;              /* u32 [node_cap][C] */
  /* ---- device scalars: i32[8] = {n_slots, n_next, n_nodes, units_hist, units_part, handed-off nodes, 0,
   *      peer barrier gave up (a rank did not arrive)} ---- */
  int64_t scalars;
  int64_t unit_prefix;           /* i32 [m_cap+1] scratch for work decomposition */
  int64_t depth;                 /* depth of the current level                  */
  /* ---- multi-GPU (filled by the comm runtime; zero when world == 1) ---- */
  /* Feature f is owned by rank f / fpo.  Rank o's staging area is u32 [world][h_cap][fpo*B*C]:
   * source rank p pushes its partial counts of o's features for slot s to [p][s][...].
   * Rank q's record mailbox holds fb_cost | fb_bin | fb_left | node_cnt back to back (the local
   * fb_* / node_cnt pointers above point into the own mailbox), written by the owners.        */
  int64_t peer_hist[MTB_MAX_WORLD];   /* peers' staging areas (mapped)          */
  int64_t peer_rec[MTB_MAX_WORLD];    /* peers' record mailboxes (mapped)       */
  int64_t peer_flags[MTB_MAX_WORLD];  /* peers' barrier flags u32[world] (mapped) */
  int64_t epoch;                      /* barrier epoch base for this level (host advances by 2) */
  /* ---- host-computed (k/n)*log2(k/n) table for n <= plogp_n: entry (n-1)*n/2 + (k-1), f64.
   * Small nodes are where rounding-level ties happen; using the host's own log2 values there
   * makes the device arithmetic bit-identical to numpy's.  0 = no table. ---- */
  int64_t plogp;
  int64_t plogp_n;
  int64_t h_cap;                      /* histogram slabs per side == staging slots per source */
  int64_t fpo;                        /* features per owner rank (F when world == 1)          */
  int64_t timers;                     /* u64[16] phase timestamps (%globaltimer) or 0         */
  /* Nodes with more than plogp_n rows use a cheaper, still deterministic evaluation:
   * cost = (g(nL) + g(nR) - sum_c g(kL_c) - sum_c g(kR_c)) / n with g(k) = k*log2(k), evaluated
   * arithmetically (csrc/impurity.cuh).  klogk: f64 [128 + 32*128 + 6] constants built by the host
   * (1/m0 per interval, e + log2 m0 per (exponent, interval), then the six coefficients of
   * log2(1+r)); klogk_n = 128.                                                              */
  int64_t klogk;
  int64_t klogk_n;
  /* world > 1: the owner's reduced slabs for its own features, ping-pong by level:
   * u32 [h_cap][fpo*B*C] each.  BUILD pairs write (own partial + staged partials) here and
   * DERIVE pairs write (parent - sibling) here, so neither touches the local partial slabs and
   * both kinds of pair can run in one pass.                                                  */
  int64_t rstore[2];
  /* ---- hand-off (phase 2): a non-terminal child with <= small_rows global rows leaves the
   * level-wise build; after its rows are partitioned into place the commit phase appends
   * (node id, row segment, ping-pong side) to this list; scalars[5] counts the entries. ---- */
  int64_t small_rows;                 /* 0 = never hand off                                   */
  int64_t ho_cap;
  int64_t ho_node, ho_begin, ho_end, ho_side; /* i32 [ho_cap]                                 */
  /* ---- fused partition + next-level histogram (mtb_level_fused2) ----
   * part_child: i32 [m_cap] first child slot of a split slot that keeps rows, else -1
   * mask_words: u32 [max_units][68] per 1024-row partition unit: 32 "row goes left" bit words, 32
   *             exclusive left-count prefixes, the rows already placed left / right before the
   *             unit, 2 pad words (272 B; written by the ranker warp while the split feature's
   *             plane is moved, fetched by TMA next to the rows of the other planes)
   * zero_next : mtb_level_fused only: also zero the slabs the next level accumulates into (set
   *             when the next launch is mtb_level_fused2, which always does this for its successor) */
  int64_t part_child, mask_words, zero_next;
  /* shared-memory histogram tile: 0 = [f][B][C], 1 = [B*C][16] interleaved by feature (FT == 16
   * only; the default there).  Slabs in global memory are [f][B][C] either way.                */
  int64_t tile_layout;
  /* world > 1, opt-in: multicast (NVSwitch) mappings of hist[0] / hist[1].  When set, the fused level
   * kernels skip the peer push and the owners read cross-rank sums with multimem.ld_reduce.  The
   * slabs must then live in a VMM allocation bound to the multicast object (parallel/nvls.py). 0 = off. */
  int64_t mc_hist[2];
  /* world > 1: how long (ns, %globaltimer) a rank waits for a peer in an in-kernel barrier before it
   * gives up, sets scalars[7] and raises the abort word (u32 index MTB_MAX_WORLD of every rank's flag
   * block), after which all barriers of the fit fall through.  0 = wait for ever.                */
  int64_t barrier_timeout_ns;
  /* depth 0 only: the root slab already holds the root histogram (mtb_root_retile_hist ran), so
   * mtb_level_fused skips its zeroing and histogram phase.                                        */
  int64_t root_hist_done;
} MtbCtx;

const char* mtb_last_error(void);
int mtb_version(void);
int mtb_sm_count(int device);
/* Make `device` current for this library's (statically linked) CUDA runtime. */
int mtb_set_device(int device);

/* ---- K1: encode raw rows into tiled bin codes ------------------------------------ */
/* X: [n][ldx] row-major (device or mapped pinned host); edges: f64 [F][256] padded with
 * +inf; writes codes[g][row0 + i][k] and (optionally) counts values that match no edge
 * exactly into *mismatch (u64, may be NULL).  dtype: 0 = f32, 1 = f64.                */
int mtb_encode_float(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx, const double* edges,
                     const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0,
                     unsigned long long* mismatch, void* stream);
/* float32 rows with F % 4 == 0, ldx % 4 == 0 and a 16-byte aligned base: same codes as
 * mtb_encode_float(dtype 0), one lane per (row, 4 features), 128-bit loads (the default
 * float32 path: 1.55 vs 4.45 ms on 10M x 32; MTB_ENCODE=scalar selects the other).     */
int mtb_encode_f32_quads(const float* X, int64_t n, int64_t F, int64_t ldx, const double* edges,
                         const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0, void* stream);
/* Same contract and codes as mtb_encode_f32_quads; the search starts from a 512-slot table over each
 * feature's value range (a guess + a short probe instead of eight bisection steps).  The default
 * float32 path (MTB_ENCODE=quads / scalar select the others).                              */
int mtb_encode_f32_slots(const float* X, int64_t n, int64_t F, int64_t ldx, const double* edges,
                         const int32_t* n_bins, uint8_t* codes, int64_t n_cap, int64_t row0, void* stream);
/* uint8 input; lut: u8 [F][256] value->code, or NULL for identity.                    */
int mtb_encode_u8(const uint8_t* X, int64_t n, int64_t F, int64_t ldx, const uint8_t* lut, uint8_t* codes,
                  int64_t n_cap, int64_t row0, void* stream);

/* ---- K2/K3: level-wise tree build (separate phases) ------------------------------- */
int mtb_level_plan(const MtbCtx* ctx, void* stream);      /* work decomposition + zero cursors       */
int mtb_level_hist(const MtbCtx* ctx, void* stream);      /* TMA-fed shared-memory histograms        */
int mtb_level_gain(const MtbCtx* ctx, void* stream);      /* derive siblings, prefix, gain, arg-min  */
int mtb_level_finalize(const MtbCtx* ctx, void* stream);  /* arg-max, children, next frontier        */
int mtb_level_partition(const MtbCtx* ctx, void* stream); /* stream compaction into the other side   */
int mtb_level_commit(const MtbCtx* ctx, void* stream);    /* next frontier's row segments            */
int mtb_level_push(const MtbCtx* ctx, void* stream);      /* world > 1: push partial slabs to their owners */
/* Levels >= 1 in one cooperative launch: partition the parents' rows (view P, the finished
 * previous level) into the children's segments while building the children's histograms
 * (view C, same buffers with the two slot sets / ping-pong sides swapped), then commit, plan,
 * peer push + barriers, gain and finalize for the children.                                  */
int mtb_level_fused2(const MtbCtx* parents, const MtbCtx* children, void* stream);
/* One cooperative launch doing plan + hist + (peer push + barriers) + gain + finalize.  Zeroes the
 * slabs it accumulates into itself (no host-side fill), and the next level's when ctx->zero_next. */
int mtb_level_fused(const MtbCtx* ctx, void* stream);
/* uint8 rows in the user's row-major layout ([n][ldx], ldx == F == 32 / 64 / 128, 16-byte aligned, memory
 * of the current GPU) -> feature-group planes + labels + root histogram in ONE pass, by clusters of F/16
 * CTAs that share each row tile through a multicast TMA copy.  ctx: codes / labels / hist / cur / hcur /
 * n_cap / F / G / B / C / FT == 16 / tile_layout == 1 must be set.  Returns 0 when launched, 1 when the
 * shape or the pointers do not qualify (nothing was launched: encode and count the usual way), < 0 on a
 * CUDA error.  Follow with mtb_level_fused(ctx with root_hist_done = 1).                             */
int mtb_root_retile_hist(const MtbCtx* ctx, const void* X, int64_t ldx, int64_t n, const void* y, void* stream);
/* The same pass without staging or clusters: CTA b takes feature group b mod G of a row range with 128-bit
 * loads from the row-major matrix (ldx == F == 16 G, any G up to the SM count).  Same contract and
 * return values.                                                                                      */
int mtb_root_direct_hist(const MtbCtx* ctx, const void* X, int64_t ldx, int64_t n, const void* y, void* stream);
/* Test hook: out[i] = k[i] * log2(k[i]) by the device function of the gain phase (table = ctx.klogk). */
int mtb_klogk_eval(const uint32_t* k, double* out, int64_t n, const double* table, void* stream);
/* Shared memory / launch geometry the hist phase will use for this ctx (for reports). */
int mtb_hist_config(const MtbCtx* ctx, int* threads, int* smem_bytes, int* stages, int* rows_per_stage);

/* ---- K6: finish small subtrees, one CTA each (struct MtbSubtreeArgs is defined in
 * mpitree_b200/csrc/subtree.cu and mirrored in mpitree_b200/ops/_lib.py) ------------------------ */
struct MtbSubtreeArgs;
int mtb_subtree_config(int64_t B, int64_t C); /* features per tile, 0 = does not fit */
int mtb_subtree_build(const struct MtbSubtreeArgs* args, int grid, void* stream);
/* Copy row segments (all feature-group planes + labels) into another (possibly peer-mapped)
 * data set: item i copies rows [src_begin[i], src_end[i]) of side src_side[i] to
 * dst_begin[i].. of the destination planes (K6 gather half).                                 */
int mtb_gather_rows(const void* src_codes0, const void* src_codes1, const void* src_labels0, const void* src_labels1,
                    int64_t src_n_cap, int64_t G, const int32_t* src_side, const int32_t* src_begin,
                    const int32_t* src_end, const int64_t* dst_codes, const int64_t* dst_labels,
                    const int64_t* dst_n_cap, const int32_t* dst_begin, int64_t n_items, void* stream);

/* ---- K4: batched predict ----------------------------------------------------------- */
/* Pack {feature, left child, leaf label} into 16-byte node records (packed_out: n_nodes*16 B). */
int mtb_pack_nodes(const int32_t* feature, const int32_t* left, const int32_t* leaf_label, int64_t n_nodes,
                   void* packed_out, void* stream);
/* dtype: 0 f32 X / f32 thresholds, 1 u8 X / i32 bin thresholds, 2 f64 X / f64 thresholds.
 * X [n][ldx]; out_leaf / out_label are i32 [n] and either may be NULL.                  */
int mtb_predict(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx, const void* packed_nodes,
                const void* threshold, int64_t n_nodes, int32_t* out_leaf, int32_t* out_label, void* stream);

/* Prediction for callers that cannot allocate device memory (a server that pre-partitioned the
 * GPU and forbids later cudaMalloc; a plugin running inside another application's allocator).
 * Same result as mtb_pack_nodes + mtb_predict, but the node-record scratch is a fixed 1 MiB array
 * inside the library (held by every GPU that loads it), so n_nodes must be <= 65536 or the call
 * fails with an error.  Takes the raw tree arrays (feature/left/leaf_label i32 [n_nodes], threshold
 * typed as for mtb_predict).  Allocates nothing.  The first call on a device resolves the array's
 * address: make it before any stream capture.  Calls share the array: overlapping calls race.
 * This entry point has no counterpart in the reference.                                          */
int mtb_predict_noalloc(const void* X, int dtype, int64_t n, int64_t F, int64_t ldx, const int32_t* feature,
                        const void* threshold, const int32_t* left, const int32_t* leaf_label, int64_t n_nodes,
                        int32_t* out_leaf, int32_t* out_label, void* stream);

/* ---- K5 / comm runtime -------------------------------------------------------------- */
int mtb_ipc_handle_size(void);
int mtb_ipc_get_handle(void* dev_ptr, void* handle_out);          /* cudaIpcGetMemHandle   */
int mtb_ipc_open_handle(const void* handle, void** dev_ptr_out);  /* cudaIpcOpenMemHandle  */
int mtb_ipc_close(void* dev_ptr);
int mtb_enable_peer_access(int peer_device);
int mtb_can_access_peer(int device, int peer);
/* The only allocating entry points: raw cudaMalloc/cudaFree for peer-visible (IPC) regions. */
int mtb_device_malloc(int64_t bytes, void** out);
int mtb_device_free(void* ptr);
int mtb_memset_async(void* ptr, int value, int64_t bytes, void* stream);
/* device-to-device copy; either address may lie in a peer's region opened with mtb_ipc_open_handle */
int mtb_memcpy_async(void* dst, const void* src, int64_t bytes, void* stream);
/* u32 sum all-reduce over peer-mapped buffers (K5).  bufs[q] = rank q's symmetric area as mapped
 * in this process: [in: cap u32][res: cap u32]; flags[q] = rank q's flag words u32[512][16].  Every
 * rank writes its `count` inputs at the start of its own area, then all ranks call this with the
 * same arguments and a common `epoch` (advance by 2 per call).  algo: 0 one-shot (pull + sum),
 * 1 two-shot (reduce-scatter + push; result in the own result area bufs[rank] + cap and, when `out`
 * is non-null and different, copied there).  One-shot writes to `out`.  Plain launches, CTAs of
 * equal index synchronise through the flags.  Spins on peers: one rank per GPU.                   */
int mtb_allreduce_u32(int algo, int rank, int world, void* const* bufs, void* const* flags, int64_t count,
                      int64_t cap, uint32_t epoch, uint32_t* out, void* stream);
/* Same contract, reduced inside the NVSwitch: the symmetric areas must come from a VMM allocation
 * bound to a multicast object (parallel/nvls.py) and `mc_buf` is that object's mapping (the same
 * layout seen through one address).  Each rank sums its 1/world slice with multimem.ld_reduce and
 * broadcasts it with multimem.st.                                                                 */
int mtb_allreduce_u32_nvls(int rank, int world, void* const* bufs, void* const* flags, void* mc_buf, int64_t count,
                           int64_t cap, uint32_t epoch, uint32_t* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MPITREE_B200_H */
