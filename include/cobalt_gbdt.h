/*
 * cobalt_gbdt.h -- C ABI of libcobalt_gbdt.so, the sm_100a histogram-GBDT hot path.
 *
 * The reference has no FFI for this path: it calls the third-party xgboost package
 * through its Python estimator.  Each entry point below therefore cites
 * the reference call site whose work it takes over and the xgboost stage it replaces
 * (K-numbers of SURVEY.md 2.1).  Callers: cobalt_smart_lender_ai_b200/_lib.py (ctypes);
 * INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - every function returns 0 on success, a negative cb_status otherwise; the message is
 *     available from cb_last_error() (thread local); nothing throws across the boundary;
 *   - all data pointers are DEVICE pointers owned by the caller unless a name says host;
 *     the library never frees or reallocates caller memory;
 *   - scratch memory is caller provided: ask the matching *_workspace_bytes() first;
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued, never synchronised,
 *     unless the function name ends in _sync;
 *   - integers are exact: gradients are fixed point (R4), histograms int64 (R5).
 */
#ifndef COBALT_GBDT_H_
#define COBALT_GBDT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CB_VERSION 200
#define CB_MISSING_BIN 255 /* u8 code of a missing value (R2) */
#define CB_MAX_BINS 255    /* real bins per feature (R2) */
#define CB_CUT_STRIDE 256  /* cuts are stored [n_features][256], padded with +inf */
#define CB_QBITS 17        /* |q_g| <= 2^17 per row (R4); equals oracle.QBITS */

typedef enum {
  CB_OK = 0,
  CB_ERR_INVALID = -1,   /* bad argument */
  CB_ERR_WORKSPACE = -2, /* workspace too small */
  CB_ERR_CUDA = -3,      /* a CUDA call failed; see cb_last_error() */
  CB_ERR_UNSUPPORTED = -4
} cb_status;

int cb_version(void);
const char* cb_last_error(void);
/* SM count and opt-in shared memory per block of the current device. */
int cb_device_info(int* sm_count, int* smem_optin_bytes);

/* ---- K1 quantile sketch: per-feature cut arrays (R1).
 * Replaces the implicit sketch inside XGBClassifier.fit
 * (src/model_train_test/model_tree_train_test.py:118,159; max_bin 256 in the shipped model).
 * X: f32 row major [n][ldx].  cuts: f32 [f][256]; n_cuts: i32 [f]; min_vals: f32 [f]. */
int cb_sketch_workspace_bytes(int64_t n_rows, int n_features, size_t* bytes);
int cb_sketch_cuts(const float* X, int64_t n_rows, int n_features, int64_t ldx, int max_bin, float* cuts,
                   int32_t* n_cuts, float* min_vals, void* workspace, size_t workspace_bytes, void* stream);

/* ---- one-hot detection on the data (one-off, before the placement is planned).
 * The reference's feature engineering emits one-hot dummies with pd.get_dummies(..., drop_first=True)
 * (src/data_preprocessing/feature_engineering.py:142-147): 2-bin features of which at most one per categorical is
 * set.  For up to 64 candidate features cand[i] (2-bin features, chosen by the host) this ADDS to
 * pair[i * n_cand + j] (i < j) the number of rows where both are in bin 1 and to miss[i] the number of rows where
 * candidate i is missing (uint64 device arrays, zero them first).  Candidates that never co-occur and are never
 * missing can be stored as one code byte (a "group").  When rows are sharded the caller sums the counts over ranks. */
int cb_onehot_pair_counts(const float* X, int64_t n_rows, int n_features, int64_t ldx, const float* cuts,
                          const int32_t* cand, int n_cand, uint64_t* pair, uint64_t* miss, void* stream);

/* ---- column placement of the binned matrix (HOST arrays, no device work).
 * A binned row holds one byte per COLUMN: a plain feature, or a group of mutually exclusive 2-bin features packed
 * as one code (0 = no member in bin 1, m + 1 = member m in bin 1).  Groups: grp_ptr_host [n_groups + 1],
 * grp_feat_host [grp_ptr[n_groups]] (members, group after group); n_groups = 0 and NULLs for none.
 * pos_feature_host[p] = column stored in byte p of a binned row: >= 0 a plain feature, -1 padding, <= -2 group
 * -2 - value.  Columns are dealt to u32 words by descending bin count so that the histogram kernel can give every
 * lane its own shared-memory bank; *conflict_free = 0 means the layout did not fit and the natural order (tiled
 * kernel) is returned -- with groups that case is refused (CB_ERR_UNSUPPORTED): plan again without them.
 * ldb: bytes per binned row, a multiple of 4, >= number of columns.
 * No reference counterpart (layout is internal to xgboost). */
int cb_plan_layout(const int32_t* n_cuts_host, int n_features, const int32_t* grp_ptr_host, const int32_t* grp_feat_host,
                   int n_groups, int64_t ldb, int32_t* pos_feature_host, int32_t* conflict_free);

/* ---- K2 bin-index packing (R2): bins u8, n * ldb bytes, ldb % 4 == 0, pad = 255.
 * pos_feature: i32 [ldb] (device) from cb_plan_layout, or NULL for the natural order; grp_ptr / grp_feat: the
 * groups as device arrays (NULL when pos_feature names no group).
 * blocked = 0: row major [n][ldb].  blocked = 1: column-block major -- bytes [128k, 128k+128) of
 * every row form their own [n][width_k] array at byte offset n * 128 * k (width 128, last block
 * ldb - 128k), so each block streams as whole aligned lines (what the conflict-free histogram reads).
 * Replaces xgboost's GHistIndexMatrix build inside fit (same call sites as K1). */
int cb_bin_pack_u8(const float* X, int64_t n_rows, int n_features, int64_t ldx, const float* cuts,
                   const int32_t* n_cuts, const int32_t* pos_feature, const int32_t* grp_ptr, const int32_t* grp_feat,
                   uint8_t* bins, int64_t ldb, int blocked, void* stream);

/* position-major copy of a binned matrix: bins_cm[p * n + row] = byte p of `row`.  Optional input of the
 * trainer: the row partition (K6) reads one byte per row of the split feature, and rows of a node are
 * ascending, so this copy turns a DRAM burst per row into a near-sequential read.  No reference counterpart. */
int cb_bin_transpose_u8(const uint8_t* bins, int64_t n_rows, int64_t ldb, int blocked, uint8_t* bins_cm, void* stream);

/* ---- K3 logistic gradient/hessian in fixed point (R4).
 * Replaces the binary:logistic objective with scale_pos_weight
 * (src/model_train_test/model_tree_train_test.py:103-116).
 * gpair: int32 [n][2].  totals: int64 [2], ADDED to (zero it first).  weight may be NULL.
 * keep_thresh: 0 = keep every row, else a row is kept iff hash(seed, tree, row_offset+i) < keep_thresh. */
int cb_grad_logistic_q(const float* margin, const uint8_t* y, const float* weight, int64_t n_rows, float spw,
                       int sg, int sh, uint32_t keep_thresh, uint32_t seed, uint32_t tree, int64_t row_offset,
                       int32_t* gpair, int64_t* totals, void* stream);
/* p = sigmoid(margin) with the specified operation sequence (R4/R11). */
int cb_sigmoid(const float* margin, int64_t n, float* out, void* stream);

/* ---- K4 per-node histogram build (R5).  Replaces xgboost BuildHist inside fit.
 * row_idx: u32 [count] or NULL (rows 0..count-1).  bin_ptr: i32 [f+1] (device).
 * hist: int64 [total_bins][2], overwritten.  Missing rows are skipped. */
int cb_hist_workspace_bytes(int64_t max_rows, int total_bins, size_t* bytes);
int cb_hist_build(const uint8_t* bins, int64_t ldb, int n_features, const int32_t* gpair, const uint32_t* row_idx,
                  int64_t count, const int32_t* bin_ptr, int total_bins, int64_t* hist, void* workspace,
                  size_t workspace_bytes, void* stream);

/* ---- K4 through the conflict-free kernel (the one the trainer launches at every level), on one node, standalone:
 * bins in the placement of cb_plan_layout / cb_bin_pack_u8 (pos_feature i32 [ldb], groups, blocked as there; all device
 * arrays), row_idx u32 [count] or NULL.  hist: int64 [total_bins][2] in natural feature order, overwritten; members of a
 * one-hot group get bin 0 = (sum of the rows' pairs) - bin 1, as the trainer's split scan derives it.  Builds the
 * shared-memory tables on the host from device copies of the placement and checks the launch's status, hence _sync.
 * CB_ERR_UNSUPPORTED when the placement has no conflict-free image (then cb_hist_build is the path). */
int cb_hist_build_cf_workspace_bytes(int64_t max_rows, int n_features, size_t* bytes);
int cb_hist_build_cf_sync(const uint8_t* bins, int64_t n_rows, int64_t ldb, int blocked, const int32_t* pos_feature,
                          const int32_t* grp_ptr, const int32_t* grp_feat, int n_groups, int n_features,
                          const int32_t* gpair, const uint32_t* row_idx, int64_t count, const int32_t* bin_ptr,
                          int total_bins, int64_t* hist, void* workspace, size_t workspace_bytes, void* stream);

/* ---- K9 sibling = parent - built (exact). */
int cb_hist_subtract(const int64_t* parent, const int64_t* built, int64_t* sibling, int64_t n_values, void* stream);

/* ---- K5 split-gain scan (R6).  Replaces HistEvaluator::EnumerateSplit. */
typedef struct {
  float eta, reg_lambda, reg_alpha, gamma, min_child_weight;
  int32_t max_depth, sg, sh;
} cb_tree_params;

typedef struct {
  float loss_chg;
  int32_t feature; /* -1: no candidate */
  int32_t split_bin;
  int32_t default_left;
  float threshold;
  int32_t pad_;
  int64_t GL, HL, GR, HR;
} cb_split;

/* totals: int64 [2] (G, H of the node); feature_mask: u8 [f] or NULL; out: one cb_split (device). */
int cb_split_scan(const int64_t* hist, const int32_t* bin_ptr, int n_features, const float* cuts,
                  const float* min_vals, const int64_t* totals, const cb_tree_params* params,
                  const uint8_t* feature_mask, cb_split* out, void* workspace, size_t workspace_bytes,
                  void* stream);
int cb_split_scan_workspace_bytes(int n_features, size_t* bytes);

/* ---- K5 for hosts that hand over inputs and outputs only.
 * No counterpart in the reference: this entry point exists for callers that are not able to give the
 * library a workspace -- a serving process whose device memory is carved up once at start-up and
 * where a later cudaMalloc is forbidden (it stalls every stream of the device and can fail when the
 * pool is full), or a foreign-language binding that can only pass the buffers it owns.  Same result
 * as cb_split_scan, bit for bit.  The per-feature candidate records live in an array inside the
 * library's device image (128 KB per GPU for the life of the process), which bounds the size:
 * n_features <= 4096, larger inputs are refused with CB_ERR_UNSUPPORTED.  It allocates nothing.
 *   - The first call on each device looks up the array's device address; make that call before
 *     any stream capture begins on that device.
 *   - All calls share the one array.  Each call overwrites the records it reads, so earlier contents
 *     never leak into a result, but two calls whose kernels overlap on the GPU (different streams,
 *     or different host threads on one device) race on it: keep them on one stream or serialise them. */
int cb_split_scan_noalloc(const int64_t* hist, const int32_t* bin_ptr, int n_features, const float* cuts,
                          const float* min_vals, const int64_t* totals, const cb_tree_params* params,
                          const uint8_t* feature_mask, cb_split* out, void* stream);

/* ---- K6 stable row partition by compaction (R8).  Replaces RowPartitioner.
 * seg_in/seg_out: u32 [count]; n_left: int64 [1] (device). */
int cb_partition_workspace_bytes(int64_t count, size_t* bytes);
int cb_partition_stable(const uint8_t* bins, int64_t ldb, const uint32_t* seg_in, int64_t count, int feature,
                        int split_bin, int default_left, uint32_t* seg_out, int64_t* n_left, void* workspace,
                        size_t workspace_bytes, void* stream);

/* ---- K7 leaf-value / prediction-cache update (R10), standalone form: margin[row_idx[i]] += leaf_value for the
 * `count` rows of one leaf's segment (row_idx NULL: rows 0..count-1).  Replaces xgboost's UpdatePredictionCache
 * inside fit (src/model_train_test/model_tree_train_test.py:118); the trainer does the same inside its partition
 * and final-level kernels. */
int cb_leaf_update(float* margin, const uint32_t* row_idx, int64_t count, float leaf_value, void* stream);

/* ---- K8 batched ensemble traversal (R11).  Replaces Booster.predict behind
 * predict_proba (src/api/cobalt_fast_api.py:91, src/model_train_test/model_tree_train_test.py:171-172).
 * nodes: packed 8-byte nodes {f32 cond; u32 meta} (meta: bits 0-14 feature, bit 15 default_left,
 * bits 16-31 left child relative to the tree start, 0 = leaf; right = left + 1).
 * tree_off: i32 [n_trees + 1].  out_margin and out_proba may each be NULL. */
int cb_predict(const float* X, int64_t n_rows, int n_features, int64_t ldx, const uint64_t* nodes,
               const int32_t* tree_off, int n_trees, float base_margin, float* out_margin, float* out_proba,
               void* stream);

/* ---- K8, complete-tree form (the fast path for forests of bounded depth): every tree padded to a
 * complete binary tree of `depth` levels in heap order (node i -> children 2i+1, 2i+2).
 * cb_predict_heap_plan gives, for the current device, the trees per shared-memory chunk and the rows
 * per CTA tile.  nodes: u64 [Tp][2^depth - 1] = threshold bits | (feature * tile_rows * 4 | default_left << 31) << 32;
 * leaf: f32 [Tp][2^depth]; Tp = n_trees rounded up to whole chunks; both 16-byte aligned (chunks are
 * streamed into shared memory with bulk copies).  Same call sites as cb_predict; bit-identical results. */
int cb_predict_heap_plan(int n_features, int depth, int* chunk_trees, int* tile_rows);
int cb_predict_heap(const float* X, int64_t n_rows, int n_features, int64_t ldx, const uint64_t* nodes,
                    const float* leaf, int n_trees, int depth, float base_margin, float* out_margin,
                    float* out_proba, void* stream);

/* ---- K8 on quantised inputs (the fastest form; same call sites, bit-identical results).  A split "x < thr" only
 * compares x with one of the forest's own thresholds, so a row is first reduced to one byte per feature
 * (cb_bin_rows_u8 against the forest's threshold table: per feature its distinct split_cond values ascending, then
 * +inf; code = number of thresholds <= x, 255 = missing; at most 254 thresholds per feature), and "x < t_k"
 * becomes "code <= k".
 * cb_predict_q8_plan: trees per shared-memory chunk and plane_base, the shared-window address the kernel's code
 * planes start at on this device (it goes into every node word and back into cb_predict_q8, which refuses a value
 * that is not this device's with CB_ERR_INVALID).  The first call on a device runs a one-thread probe kernel and
 * synchronises; it allocates nothing.
 * codes: u8 [n_rows][ldc] row major, ldc = 4 * odd >= n_features, 16-byte aligned.
 * forest: u32 [Tp][2 << depth] per tree in 1-based heap order: word 0 unused, words 1 .. 2^depth - 1 the nodes
 * = span << 16 | plane_base + (default_left * n_features + feature) * 128 with span = k (+1 when default_left;
 * pass-through and padding nodes: plane_base), words 2^depth .. 2^(depth+1) - 1 the leaf values (f32 bits).
 * Tp = n_trees rounded up to whole chunks, 16-byte aligned.  depth <= 10; plane_base + 256 * n_features <= 65536
 * (about 250 features), else cb_predict_q8_plan reports CB_ERR_UNSUPPORTED and cb_predict_heap is the path.
 * cb_bin_rows_u8: K2 with the table of a 32-feature slab held in shared memory; natural column order, row major,
 * same codes as cb_bin_pack_u8 (cuts f32 [n_features][256], n_cuts i32 [n_features]). */
int cb_predict_q8_plan(int n_features, int depth, int64_t ldc, int* chunk_trees, uint32_t* plane_base);
int cb_bin_rows_u8(const float* X, int64_t n_rows, int n_features, int64_t ldx, const float* cuts,
                   const int32_t* n_cuts, uint8_t* codes, int64_t ldc, void* stream);
int cb_predict_q8(const uint8_t* codes, int64_t n_rows, int n_features, int64_t ldc, const uint32_t* forest,
                  int n_trees, int depth, float base_margin, uint32_t plane_base, float* out_margin,
                  float* out_proba, void* stream);

/* ---- K10 path-dependent TreeSHAP (Lundberg, Erion & Lee 2018, Algorithm 2).  Replaces
 * shap.TreeExplainer(xgb_model).shap_values(row) / .expected_value (src/api/cobalt_fast_api.py:46, :100-105).
 * Forest in the wire format: arrays concatenated over trees, tree_off i64 [n_trees + 1], child ids
 * relative to the tree start and larger than the parent's, left == -1 marks a leaf whose value is split_cond;
 * sum_hess is the node cover.  max_depth: depth of the deepest tree (<= 12, computed by the host).
 * out: f32 [n_rows][n_features + 1]; the last column is the expected value (cover-weighted mean of the leaves of
 * every tree, plus base_margin); every row of out sums to the row's margin.  f32 arithmetic in a fixed order:
 * bit-identical to the oracle's R13. */
int cb_tree_shap_workspace_bytes(int64_t n_rows, int n_features, int n_trees, int64_t n_nodes, size_t* bytes);
int cb_tree_shap(const float* X, int64_t n_rows, int n_features, int64_t ldx, const int64_t* tree_off, int n_trees,
                 int64_t n_nodes, const int32_t* left, const int32_t* right, const int32_t* split_index,
                 const float* split_cond, const uint8_t* default_left, const float* sum_hess, float base_margin,
                 int max_depth, float* out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- fused trainer: one boosting round = K3 + per level (K4, K9, K5, K6) + K7, with every
 * split decision kept on the device (no host synchronisation inside a tree).
 * Replaces the body of XGBClassifier.fit's boosting loop. */
typedef struct {
  int64_t n_rows;     /* rows on this rank */
  int64_t ldb;
  int64_t row_offset; /* global index of local row 0 (row sampling is a function of the global row) */
  int32_t n_features;
  int32_t total_bins;
  int32_t max_trees;  /* forest capacity (rows of the forest arrays) */
  int32_t world_size; /* >1: the caller sums the level buffer over ranks between level_hist and level_split */
  cb_tree_params tree;
  float scale_pos_weight;
  uint32_t keep_thresh;
  uint32_t seed;
  int32_t blocked;    /* storage of `bins`: 0 row major, 1 column-block major (see cb_bin_pack_u8) */
  int32_t n_groups;   /* one-hot groups packed as one column each (cb_plan_layout); 0 = none */
} cb_train_config;

typedef struct {
  const uint8_t* bins;     /* n * ldb bytes, layout per cb_train_config.blocked */
  const uint8_t* y;        /* [n] */
  const float* weight;     /* [n] or NULL */
  float* margin;           /* [n] in/out */
  const int32_t* bin_ptr;  /* [f+1] */
  const float* cuts;       /* [f][256] */
  const float* min_vals;   /* [f] */
  const int32_t* pos_feature; /* [ldb] column placement used when the rows were binned, or NULL */
  const uint8_t* bins_cm;  /* [ldb][n] position-major copy (cb_bin_transpose_u8), or NULL */
  /* forest out, each [max_trees][max_nodes], max_nodes = 2^(max_depth+1)-1 (wire format a-9) */
  int32_t* left; int32_t* right; int32_t* parent; int32_t* split_index; int32_t* split_bin;
  float* split_cond; uint8_t* default_left; float* base_weight; float* loss_chg; float* sum_hess;
  int32_t* n_nodes;        /* [max_trees] */
  const int32_t* grp_ptr;  /* [n_groups + 1] or NULL */
  const int32_t* grp_feat; /* [grp_ptr[n_groups]] member features, group after group, or NULL */
} cb_train_buffers;

typedef struct cb_trainer_s cb_trainer;

int cb_trainer_workspace_bytes(const cb_train_config* cfg, size_t* bytes);
int cb_trainer_create(const cb_train_config* cfg, const cb_train_buffers* buf, void* workspace,
                      size_t workspace_bytes, cb_trainer** out);
void cb_trainer_destroy(cb_trainer* tr);
/* whole round on one GPU.  feature_masks: u8 [max_trees][f] (device; row t masks tree t) or NULL
 * (colsample_bytree).  tree_index >= 0 selects the forest row; tree_index < 0 continues from the
 * device-side counter that every round_end advances, so one captured CUDA graph of a round can be
 * replayed round after round without any host value changing. */
int cb_trainer_boost_round(cb_trainer* tr, int tree_index, const uint8_t* feature_masks, void* stream);
/* the same round in steps, for row-sharded training:
 *   round_begin; for level in 0..max_depth-1: level_hist, <allreduce level buffer>, level_split; round_end */
int cb_trainer_round_begin(cb_trainer* tr, int tree_index, const uint8_t* feature_masks, void* stream);
int cb_trainer_level_hist(cb_trainer* tr, int level, void* stream);
int cb_trainer_level_split(cb_trainer* tr, int level, void* stream);
int cb_trainer_round_end(cb_trainer* tr, void* stream);
/* ---- row-sharded training over NVLink peer memory (the fused form of the path's one collective).
 * Every rank allocates one "symmetric" region (cb_sym_alloc; size from cb_trainer_sym_bytes), exchanges the
 * 64-byte IPC handles with its peers by any means (torch.distributed here), maps theirs (cb_sym_open) and hands
 * all `world` addresses, own included at index `rank`, to cb_trainer_set_peers.  After that the level calls need
 * no all-reduce in between: the scan kernel of each rank owns a slice of the features and adds that slice of every
 * rank's local histogram straight out of peer memory, then publishes one candidate per node to all ranks; ranks
 * synchronise through epoch flags in the regions.  cb_trainer_boost_round then runs a whole sharded round.
 * No reference counterpart (the reference has no distributed code); replaces the NCCL all-reduce of the level buffer. */
int cb_sym_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int cb_sym_open(const unsigned char* handle64, void** ptr);
int cb_sym_close(void* ptr);
int cb_sym_free(void* ptr);
int cb_trainer_sym_bytes(const cb_train_config* cfg, size_t* bytes);
int cb_trainer_set_peers(cb_trainer* tr, int rank, int world, void* const* regions);
/* *timed_out != 0: a wait on a peer flag gave up (a peer died); results are then invalid.  Synchronises. */
int cb_trainer_peer_status(cb_trainer* tr, int* timed_out);
/* int64 buffer to sum over ranks after level_hist(level): pointer and element count */
int cb_trainer_level_buffer(cb_trainer* tr, int level, int64_t** ptr, int64_t* n_values);
/* introspection for parity tests: current row order u32 [n], histogram pool slot of a node */
int cb_trainer_row_order(cb_trainer* tr, const uint32_t** ptr);
int cb_trainer_node_hist(cb_trainer* tr, int node_id, const int64_t** ptr);
/* measurement hooks (bench.py): CUDA events around the histogram kernels of level_hist.
 * last_hist_timing_sync blocks until the last level_hist finished and also returns the rows it scanned. */
int cb_trainer_enable_timing(cb_trainer* tr, int on);
int cb_trainer_last_hist_timing_sync(cb_trainer* tr, float* ms_accum, float* ms_reduce, int64_t* rows, int* n_build);
/* per-launch timeline of one round (warm caches, back to back): profile_begin, then one round through the
 * calls above, then profile_end_sync returns the elapsed ms between consecutive launches and their names */
int cb_trainer_profile_begin(cb_trainer* tr);
int cb_trainer_profile_end_sync(cb_trainer* tr, float* ms, const char** names, int capacity, int* n);
/* which histogram kernel this trainer runs: conflict-free layout or the tiled fallback, shared-memory rows of 32 banks
 * per array (R), dynamic shared memory per CTA */
int cb_trainer_layout_info(cb_trainer* tr, int* conflict_free, int* smem_rows, int* smem_bytes);
/* kernels launched by this trainer so far (bench.py's gpu_launches) */
int64_t cb_trainer_launch_count(cb_trainer* tr);

#ifdef __cplusplus
}
#endif
#endif /* COBALT_GBDT_H_ */
