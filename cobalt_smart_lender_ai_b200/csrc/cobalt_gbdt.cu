// libcobalt_gbdt.so -- C ABI over the sm_100a kernels (see include/cobalt_gbdt.h).
#include <cstdlib>
#include <mutex>
#include <new>

#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler is attached

#include "sketch.cuh"
#include "hist_cf.cuh"
#include "trainer.cuh"
#include "shap.cuh"

namespace cb {
thread_local char g_err[512] = "";

int device_info(DeviceInfo* out) {
  int dev = 0;
  CB_CUDA(cudaGetDevice(&dev));
  static thread_local DeviceInfo cache;
  if (cache.device != dev) {
    CB_CUDA(cudaDeviceGetAttribute(&cache.sm_count, cudaDevAttrMultiProcessorCount, dev));
    CB_CUDA(cudaDeviceGetAttribute(&cache.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    cache.device = dev;
  }
  *out = cache;
  return CB_OK;
}

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// NVTX range around the launches of one stage of the path (sketch / bin / grad / hist / scan+partition / leaf /
// predict): Nsight Systems then shows the stages -- and the all-reduce between "hist" and "split" -- by name.
struct StageRange {
  explicit StageRange(const char* name) { nvtxRangePushA(name); }
  ~StageRange() { nvtxRangePop(); }
};

// Launch with programmatic stream serialisation (common.cuh pdl_enter): the kernel may be scheduled while its
// predecessor in the stream is still draining; it blocks in griddepcontrol.wait until that one has completed.
// COBALT_PDL=1 (read once) switches it on; the default launches plainly.
static bool pdl_enabled() {     // measured on B200: 5 % slower with the overlap than without at 2.2M x 150, so opt-in
  static const bool on = [] { const char* e = getenv("COBALT_PDL"); return e && e[0] == '1'; }();
  return on;
}
template <typename... P, typename... A>
static inline cudaError_t launch_dep(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, A&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<P>(std::forward<A>(args))...);
}

// shared-memory budget of hist_accum in histogram slots (8 bytes per slot)
static inline int hist_budget_slots(const DeviceInfo& di) { return (di.smem_optin - 2048) / 8; }
static inline int hist_tiles_bound(int S, int budget) {
  int eff = budget - 4 * CB_MAX_BINS;
  int t = (S + eff - 1) / eff;
  return t < 1 ? 1 : t;
}
static inline int64_t max_partials(int64_t rows, int grid) { return (rows + kChunkRows - 1) / kChunkRows + 2 * (int64_t)grid + kMaxBuild; }

static int launch_hist_accum(const uint8_t* bins, int64_t ldb, const int32_t* gpair, const uint32_t* row_idx,
                             int identity, const HistPlan* plan, const int32_t* bin_ptr, int f, int S, int grid,
                             int n_tiles_bound, int smem_bytes, int32_t* partials, cudaStream_t s) {
  static thread_local int attr_dev = -1, attr_smem = 0;
  int dev = 0;
  CB_CUDA(cudaGetDevice(&dev));
  if (attr_dev != dev || attr_smem < smem_bytes) {
    CB_CUDA(cudaFuncSetAttribute(hist_accum_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    attr_dev = dev; attr_smem = smem_bytes;
  }
  dim3 g(grid, n_tiles_bound);
  hist_accum_kernel<4><<<g, 1024, smem_bytes, s>>>(reinterpret_cast<const uint32_t*>(bins), (int)(ldb / 4),
                                                   reinterpret_cast<const int2*>(gpair), row_idx, identity, plan,
                                                   bin_ptr, f, S, partials);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

}  // namespace cb

using namespace cb;

extern "C" {

int cb_version(void) { return CB_VERSION; }
const char* cb_last_error(void) { return g_err; }

int cb_device_info(int* sm_count, int* smem_optin_bytes) {
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  if (sm_count) *sm_count = di.sm_count;
  if (smem_optin_bytes) *smem_optin_bytes = di.smem_optin;
  return CB_OK;
}

// ---------------------------------------------------------------------------- K2
int cb_onehot_pair_counts(const float* X, int64_t n, int f, int64_t ldx, const float* cuts, const int32_t* cand, int n_cand,
                          uint64_t* pair, uint64_t* miss, void* stream) {
  if (!X || !cuts || !cand || !pair || !miss || n < 0 || f <= 0 || ldx < f || n_cand < 1 || n_cand > 64)
    return fail(CB_ERR_INVALID, "cb_onehot_pair_counts: bad arguments (1..64 candidates)");
  if (n == 0) return CB_OK;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  const int grid = (int)std::min<int64_t>(ceil_div(n, 256), (int64_t)di.sm_count * 4);
  onehot_pair_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, n, ldx, cuts, cand, n_cand,
                                                            reinterpret_cast<unsigned long long*>(pair),
                                                            reinterpret_cast<unsigned long long*>(miss));
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_plan_layout(const int32_t* n_cuts_host, int f, const int32_t* grp_ptr_host, const int32_t* grp_feat_host, int n_groups,
                   int64_t ldb, int32_t* pos_feature_host, int32_t* conflict_free) {
  if (!n_cuts_host || !pos_feature_host || f <= 0 || ldb <= 0 || (ldb & 3) || n_groups < 0 ||
      (n_groups > 0 && (!grp_ptr_host || !grp_feat_host)))
    return fail(CB_ERR_INVALID, "cb_plan_layout: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  std::vector<int32_t> bp((size_t)f + 1, 0);
  for (int i = 0; i < f; ++i) {
    if (n_cuts_host[i] < 1 || n_cuts_host[i] > CB_MAX_BINS) return fail(CB_ERR_INVALID, "cb_plan_layout: n_cuts[%d] out of range", i);
    bp[i + 1] = bp[i] + n_cuts_host[i];
  }
  GroupTable gt;
  gt.n_groups = n_groups; gt.ptr = grp_ptr_host; gt.feat = grp_feat_host;
  std::vector<int> group_of;
  if (!check_groups(n_cuts_host, f, gt, &group_of))
    return fail(CB_ERR_INVALID, "cb_plan_layout: a group holds a feature twice, out of range, or with other than 2 bins");
  if (count_columns(f, gt) > ldb) return fail(CB_ERR_INVALID, "cb_plan_layout: ldb %lld < %d columns", (long long)ldb, count_columns(f, gt));
  bool ok = plan_positions(n_cuts_host, f, gt, ldb, pos_feature_host);
  CfHostTables tb;
  ok = ok && build_cf_tables(bp.data(), pos_feature_host, f, gt, ldb, hist_budget_slots(di) * 8, &tb);
  if (!ok) {                       // does not fit one shared-memory image: natural column order, tiled kernel (no groups)
    if (n_groups > 0) return fail(CB_ERR_UNSUPPORTED, "cb_plan_layout: grouped columns need the conflict-free layout, which does not fit; plan again without groups");
    if (ldb < f) return fail(CB_ERR_INVALID, "cb_plan_layout: ldb < n_features");
    for (int64_t p = 0; p < ldb; ++p) pos_feature_host[p] = p < f ? (int32_t)p : -1;
  }
  if (conflict_free) *conflict_free = ok ? 1 : 0;
  return CB_OK;
}

int cb_bin_pack_u8(const float* X, int64_t n, int f, int64_t ldx, const float* cuts, const int32_t* n_cuts,
                   const int32_t* pos_feature, const int32_t* grp_ptr, const int32_t* grp_feat, uint8_t* bins, int64_t ldb,
                   int blocked, void* stream) {
  if (!X || !cuts || !n_cuts || !bins || n < 0 || f <= 0 || ldx < f || ldb <= 0 || (ldb & 3) || (!pos_feature && ldb < f) ||
      ((grp_ptr == nullptr) != (grp_feat == nullptr)))
    return fail(CB_ERR_INVALID, "cb_bin_pack_u8: bad arguments (ldb must be a multiple of 4 and hold every column)");
  if (n == 0) return CB_OK;
  StageRange nv("cobalt:bin_pack");
  int wpr = (int)(ldb / 4);
  int64_t items = n * wpr;
  bin_pack_kernel<<<ceil_div(items, 256), 256, 0, (cudaStream_t)stream>>>(X, n, f, ldx, cuts, n_cuts, pos_feature, grp_ptr, grp_feat,
                                                                         bins, wpr, make_view(bins, n, ldb, blocked));
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_bin_transpose_u8(const uint8_t* bins, int64_t n, int64_t ldb, int blocked, uint8_t* bins_cm, void* stream) {
  if (!bins || !bins_cm || n < 0 || ldb <= 0 || (ldb & 3)) return fail(CB_ERR_INVALID, "cb_bin_transpose_u8: bad arguments");
  if (n == 0) return CB_OK;
  bin_transpose_kernel<<<dim3((unsigned)((n + 31) / 32), (unsigned)((ldb + 31) / 32)), 256, 0, (cudaStream_t)stream>>>(
      make_view(bins, n, ldb, blocked), n, bins_cm);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K3
int cb_grad_logistic_q(const float* margin, const uint8_t* y, const float* weight, int64_t n, float spw, int sg,
                       int sh, uint32_t keep_thresh, uint32_t seed, uint32_t tree, int64_t row_offset,
                       int32_t* gpair, int64_t* totals, void* stream) {
  if (!margin || !y || !gpair || !totals || n < 0) return fail(CB_ERR_INVALID, "cb_grad_logistic_q: bad arguments");
  if (n == 0) return CB_OK;
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  int grid = (int)std::min<int64_t>(ceil_div(n, 256), (int64_t)di.sm_count * 8);
  grad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(margin, y, weight, n, spw, ldexpf(1.0f, sg), ldexpf(1.0f, sh),
                                                      keep_thresh, seed, tree, row_offset,
                                                      reinterpret_cast<int2*>(gpair),
                                                      reinterpret_cast<unsigned long long*>(totals), nullptr);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_sigmoid(const float* margin, int64_t n, float* out, void* stream) {
  if (!margin || !out || n < 0) return fail(CB_ERR_INVALID, "cb_sigmoid: bad arguments");
  if (n == 0) return CB_OK;
  sigmoid_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(margin, n, out);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K4
int cb_hist_workspace_bytes(int64_t max_rows, int total_bins, size_t* bytes) {
  if (!bytes || max_rows < 0 || total_bins <= 0) return fail(CB_ERR_INVALID, "cb_hist_workspace_bytes: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  Bump b(nullptr);
  b.take<HistPlan>(1);
  b.take<int32_t>((size_t)max_partials(max_rows, di.sm_count) * 2 * total_bins);
  *bytes = b.used();
  return CB_OK;
}

int cb_hist_build(const uint8_t* bins, int64_t ldb, int f, const int32_t* gpair, const uint32_t* row_idx,
                  int64_t count, const int32_t* bin_ptr, int total_bins, int64_t* hist, void* ws, size_t ws_bytes,
                  void* stream) {
  if (!bins || !gpair || !bin_ptr || !hist || !ws || f <= 0 || total_bins <= 0 || count < 0 || (ldb & 3) ||
      count > 2147483647LL)
    return fail(CB_ERR_INVALID, "cb_hist_build: bad arguments");
  size_t need = 0;
  int rc = cb_hist_workspace_bytes(count, total_bins, &need);
  if (rc) return rc;
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_hist_build: workspace %zu < %zu bytes", ws_bytes, need);
  DeviceInfo di;
  device_info(&di);
  cudaStream_t s = (cudaStream_t)stream;
  Bump b(ws);
  HistPlan* plan = b.take<HistPlan>(1);
  int32_t* partials = b.take<int32_t>((size_t)max_partials(count, di.sm_count) * 2 * total_bins);
  const int budget = hist_budget_slots(di);
  const int tiles = hist_tiles_bound(total_bins, budget);
  if (tiles > kMaxTiles) return fail(CB_ERR_UNSUPPORTED, "cb_hist_build: %d bins need more than %d tiles", total_bins, kMaxTiles);
  plan_tiles_kernel<<<1, 1, 0, s>>>(bin_ptr, f, budget, plan);
  plan_single_kernel<<<1, 1, 0, s>>>(plan, (int)count, di.sm_count);
  CB_LAUNCH_CHECK();
  const int smem = std::min(budget, total_bins) * 8;
  rc = launch_hist_accum(bins, ldb, gpair, row_idx, row_idx == nullptr, plan, bin_ptr, f, total_bins, di.sm_count,
                         tiles, smem, partials, s);
  if (rc) return rc;
  hist_reduce_kernel<<<dim3(ceil_div(total_bins, 256), 1), 256, 0, s>>>(partials, plan, total_bins, hist);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---- K4 through the conflict-free kernel, standalone (the kernel the trainer runs at every level)
}  // extern "C"
namespace cb {
__global__ void __launch_bounds__(256) sum_gpair_kernel(const int2* __restrict__ gpair, const uint32_t* __restrict__ rows,
                                                        int64_t count, unsigned long long* __restrict__ totals) {
  long long g = 0, h = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const int2 q = gpair[rows ? (int64_t)rows[i] : i];
    g += q.x; h += q.y;
  }
  g = warp_sum_i64(g); h = warp_sum_i64(h);
  if ((threadIdx.x & 31) == 0) { atomicAdd(totals, (unsigned long long)g); atomicAdd(totals + 1, (unsigned long long)h); }
}
// members of a one-hot group column: bin 0 = node totals - bin 1 (what the trainer's split scan derives)
__global__ void group_bin0_kernel(int64_t* __restrict__ hist, const int32_t* __restrict__ bin_ptr,
                                  const int32_t* __restrict__ feat_code, int f, const long long* __restrict__ totals) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= f || feat_code[j] == 0) return;
  const int b0 = bin_ptr[j];
  hist[2 * (size_t)b0] = totals[0] - hist[2 * (size_t)(b0 + 1)];
  hist[2 * (size_t)b0 + 1] = totals[1] - hist[2 * (size_t)(b0 + 1) + 1];
}
static size_t hist_cf_layout(int64_t max_rows, int f, const DeviceInfo& di, void* ws, HistPlan** plan, CfLayout** lay,
                             int32_t** slot_of, int32_t** fsrc, int32_t** fdst, int32_t** feat_code, long long** totals,
                             unsigned int** flag, int32_t** partials) {
  Bump b(ws);
  const int words = hist_budget_slots(di);              // upper bound of R * 32
  HistPlan* p = b.take<HistPlan>(1);
  CfLayout* l = b.take<CfLayout>(1);
  int32_t* so = b.take<int32_t>((size_t)words);
  int32_t* a = b.take<int32_t>((size_t)words);
  int32_t* c = b.take<int32_t>((size_t)words);
  int32_t* fc = b.take<int32_t>((size_t)f);
  long long* t = b.take<long long>(2);
  unsigned int* fl = b.take<unsigned int>(1);
  int32_t* pa = b.take<int32_t>((size_t)max_partials(max_rows, di.sm_count) * 2 * words);
  if (plan) { *plan = p; *lay = l; *slot_of = so; *fsrc = a; *fdst = c; *feat_code = fc; *totals = t; *flag = fl; *partials = pa; }
  return b.used();
}
}  // namespace cb
extern "C" {

int cb_hist_build_cf_workspace_bytes(int64_t max_rows, int n_features, size_t* bytes) {
  if (!bytes || max_rows < 0 || n_features <= 0) return fail(CB_ERR_INVALID, "cb_hist_build_cf_workspace_bytes: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  *bytes = hist_cf_layout(max_rows, n_features, di, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
  return CB_OK;
}

int cb_hist_build_cf_sync(const uint8_t* bins, int64_t n_rows, int64_t ldb, int blocked, const int32_t* pos_feature,
                          const int32_t* grp_ptr, const int32_t* grp_feat, int n_groups, int f, const int32_t* gpair,
                          const uint32_t* row_idx, int64_t count, const int32_t* bin_ptr, int total_bins, int64_t* hist,
                          void* ws, size_t ws_bytes, void* stream) {
  if (!bins || !pos_feature || !gpair || !bin_ptr || !hist || !ws || f <= 0 || total_bins <= 0 || count < 0 || n_rows <= 0 ||
      ldb <= 0 || (ldb & 3) || count > 2147483647LL || n_groups < 0 || (n_groups > 0 && (!grp_ptr || !grp_feat)))
    return fail(CB_ERR_INVALID, "cb_hist_build_cf_sync: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  size_t need = 0;
  cb_hist_build_cf_workspace_bytes(count, f, &need);
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_hist_build_cf_sync: workspace %zu < %zu bytes", ws_bytes, need);
  cudaStream_t s = (cudaStream_t)stream;
  CB_CUDA(cudaStreamSynchronize(s));
  std::vector<int32_t> bp((size_t)f + 1), pf((size_t)ldb), gptr((size_t)n_groups + 1, 0), gfeat, fc((size_t)f, 0);
  CB_CUDA(cudaMemcpy(bp.data(), bin_ptr, sizeof(int32_t) * (f + 1), cudaMemcpyDeviceToHost));
  CB_CUDA(cudaMemcpy(pf.data(), pos_feature, sizeof(int32_t) * ldb, cudaMemcpyDeviceToHost));
  if (n_groups > 0) {
    CB_CUDA(cudaMemcpy(gptr.data(), grp_ptr, sizeof(int32_t) * (n_groups + 1), cudaMemcpyDeviceToHost));
    if (gptr[0] != 0 || gptr[n_groups] < n_groups || gptr[n_groups] > f) return fail(CB_ERR_INVALID, "cb_hist_build_cf_sync: grp_ptr is malformed");
    gfeat.resize((size_t)gptr[n_groups]);
    CB_CUDA(cudaMemcpy(gfeat.data(), grp_feat, sizeof(int32_t) * gfeat.size(), cudaMemcpyDeviceToHost));
  }
  if (bp[f] != total_bins) return fail(CB_ERR_INVALID, "cb_hist_build_cf_sync: total_bins does not match bin_ptr");
  GroupTable gt;
  gt.n_groups = n_groups; gt.ptr = gptr.data(); gt.feat = gfeat.data();
  CfHostTables tb;
  if (!build_cf_tables(bp.data(), pf.data(), f, gt, ldb, hist_budget_slots(di) * 8, &tb))
    return fail(CB_ERR_UNSUPPORTED, "cb_hist_build_cf_sync: the placement has no conflict-free shared-memory image; use cb_hist_build");
  const BinsView bv = make_view(bins, n_rows, ldb, blocked);
  if (tb.lay.n_blocks > 1 && !bv.blocked) return fail(CB_ERR_UNSUPPORTED, "cb_hist_build_cf_sync: rows wider than 128 bytes must be stored block-major");
  for (int g = 0; g < n_groups; ++g)
    for (int k = gptr[g]; k < gptr[g + 1]; ++k) fc[gfeat[k]] = k - gptr[g] + 1;
  HistPlan* plan; CfLayout* lay; int32_t *slot_of, *fsrc, *fdst, *feat_code, *partials; long long* totals; unsigned int* flag;
  hist_cf_layout(count, f, di, ws, &plan, &lay, &slot_of, &fsrc, &fdst, &feat_code, &totals, &flag, &partials);
  CB_CUDA(cudaMemsetAsync(plan, 0, sizeof(HistPlan), s));
  CB_CUDA(cudaMemsetAsync(totals, 0, 2 * sizeof(long long), s));
  CB_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned int), s));
  CB_CUDA(cudaMemcpyAsync(lay, &tb.lay, sizeof(CfLayout), cudaMemcpyHostToDevice, s));
  CB_CUDA(cudaMemcpyAsync(slot_of, tb.slot_of.data(), sizeof(int32_t) * tb.slot_of.size(), cudaMemcpyHostToDevice, s));
  if (tb.lay.n_fold) {
    CB_CUDA(cudaMemcpyAsync(fsrc, tb.fold_src.data(), sizeof(int32_t) * tb.fold_src.size(), cudaMemcpyHostToDevice, s));
    CB_CUDA(cudaMemcpyAsync(fdst, tb.fold_dst.data(), sizeof(int32_t) * tb.fold_dst.size(), cudaMemcpyHostToDevice, s));
  }
  CB_CUDA(cudaMemcpyAsync(feat_code, fc.data(), sizeof(int32_t) * f, cudaMemcpyHostToDevice, s));
  CB_CUDA(cudaStreamSynchronize(s));                     // the host vectors go out of scope below
  const int grid = std::min(di.sm_count, kMaxGrid);
  plan_single_kernel<<<1, 1, 0, s>>>(plan, (int)count, grid);
  CB_CUDA(cudaFuncSetAttribute(hist_accum_cf_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCfSmemBytes));
  CB_CUDA(cudaFuncSetAttribute(hist_accum_cf_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCfSmemBytes));
  auto kern = row_idx ? hist_accum_cf_kernel<4, false> : hist_accum_cf_kernel<4, true>;
  kern<<<grid, 1024, kCfSmemBytes, s>>>(bins, bv.blocked ? n_rows : 0, bv.blocked ? bv.last_w : (int)ldb,
                                        reinterpret_cast<const int2*>(gpair), row_idx, plan, lay, fsrc, fdst, partials, slot_of,
                                        total_bins, hist, plan->done, flag, nullptr, nullptr);
  if (n_groups > 0) {
    if (count > 0)
      sum_gpair_kernel<<<std::min(ceil_div(count, 256), di.sm_count * 4), 256, 0, s>>>(
          reinterpret_cast<const int2*>(gpair), row_idx, count, reinterpret_cast<unsigned long long*>(totals));
    group_bin0_kernel<<<ceil_div(f, 128), 128, 0, s>>>(hist, bin_ptr, feat_code, f, totals);
  }
  CB_LAUNCH_CHECK();
  unsigned int timed_out = 0;
  CB_CUDA(cudaMemcpyAsync(&timed_out, flag, sizeof timed_out, cudaMemcpyDeviceToHost, s));
  CB_CUDA(cudaStreamSynchronize(s));
  if (timed_out) return fail(CB_ERR_CUDA, "cb_hist_build_cf_sync: the CTAs of the launch were not co-resident (another kernel holds SMs)");
  return CB_OK;
}

int cb_hist_subtract(const int64_t* parent, const int64_t* built, int64_t* sibling, int64_t n, void* stream) {
  if (!parent || !built || !sibling || n < 0) return fail(CB_ERR_INVALID, "cb_hist_subtract: bad arguments");
  if (n == 0) return CB_OK;
  subtract_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(parent, built, sibling, n);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K5
}  // extern "C"

namespace cb {
__global__ void __launch_bounds__(128) scan_single_kernel(const int64_t* __restrict__ hist,
                                                          const int32_t* __restrict__ bin_ptr, int f,
                                                          const int64_t* __restrict__ totals, cb_tree_params p,
                                                          const uint8_t* __restrict__ feature_mask,
                                                          FeatBest* __restrict__ feat_best) {
  const int feat = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (feat >= f) return;
  FeatBest fb;
  fb.chg = 0.0f; fb.split_bin = -1; fb.default_left = 0; fb.pad = 0; fb.GL = 0; fb.HL = 0;
  if (!feature_mask || feature_mask[feat]) {
    const int b0 = __ldg(bin_ptr + feat), nb = __ldg(bin_ptr + feat + 1) - b0;
    fb = scan_feature(hist + 2 * (size_t)b0, nb, totals[0], totals[1], p);
  }
  if ((threadIdx.x & 31) == 0) feat_best[feat] = fb;
}

__global__ void pick_single_kernel(const FeatBest* __restrict__ feat_best, int f, const float* __restrict__ cuts,
                                   const float* __restrict__ min_vals, const int64_t* __restrict__ totals,
                                   cb_split* out) {
  if (threadIdx.x || blockIdx.x) return;
  float bc = 0.0f; int bf = -1;
  for (int j = 0; j < f; ++j)
    if (feat_best[j].chg > bc) { bc = feat_best[j].chg; bf = j; }
  cb_split s;
  memset(&s, 0, sizeof s);
  s.feature = bf; s.loss_chg = bc;
  if (bf >= 0) {
    const FeatBest fb = feat_best[bf];
    s.split_bin = fb.split_bin; s.default_left = fb.default_left;
    s.threshold = fb.split_bin >= 0 ? cuts[(size_t)bf * CB_CUT_STRIDE + fb.split_bin] : min_vals[bf];
    s.GL = fb.GL; s.HL = fb.HL; s.GR = totals[0] - fb.GL; s.HR = totals[1] - fb.HL;
  }
  *out = s;
}
}  // namespace cb

extern "C" {

int cb_split_scan_workspace_bytes(int f, size_t* bytes) {
  if (!bytes || f <= 0) return fail(CB_ERR_INVALID, "cb_split_scan_workspace_bytes: bad arguments");
  Bump b(nullptr);
  b.take<FeatBest>((size_t)f);
  *bytes = b.used();
  return CB_OK;
}

int cb_split_scan(const int64_t* hist, const int32_t* bin_ptr, int f, const float* cuts, const float* min_vals,
                  const int64_t* totals, const cb_tree_params* params, const uint8_t* feature_mask, cb_split* out,
                  void* ws, size_t ws_bytes, void* stream) {
  if (!hist || !bin_ptr || !cuts || !min_vals || !totals || !params || !out || !ws || f <= 0)
    return fail(CB_ERR_INVALID, "cb_split_scan: bad arguments");
  size_t need = 0;
  cb_split_scan_workspace_bytes(f, &need);
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_split_scan: workspace %zu < %zu bytes", ws_bytes, need);
  cudaStream_t s = (cudaStream_t)stream;
  FeatBest* fbuf = Bump(ws).take<FeatBest>((size_t)f);
  scan_single_kernel<<<ceil_div(f, 4), 128, 0, s>>>(hist, bin_ptr, f, totals, *params, feature_mask, fbuf);
  pick_single_kernel<<<1, 1, 0, s>>>(fbuf, f, cuts, min_vals, totals, out);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K5 without a workspace
}  // extern "C"

// Scratch of cb_split_scan_noalloc: one candidate record per feature.  It lives in the module's
// device image (128 KB on every GPU that loads the library, until the process exits).
#define CB_NOALLOC_MAX_FEATURES 4096
static __device__ __align__(256) cb::FeatBest g_noalloc_feat_best[CB_NOALLOC_MAX_FEATURES];

namespace cb {
// Device address of g_noalloc_feat_best on the current device.  In host code the array's name only
// stands for the symbol, so the address has to be asked for, once per device, with that device current.
static int noalloc_scratch(FeatBest** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static FeatBest* ptrs[kMaxDev] = {nullptr};
  int dev = 0;
  CB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(CB_ERR_UNSUPPORTED, "cb_split_scan_noalloc: device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev]) {
    // a cudaDeviceReset() since the set-up leaves a stale address behind: ask the driver whether it still is device
    // memory of this GPU (a query, not an allocation) and look the symbol up again if not
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptrs[dev]) != cudaSuccess || at.type != cudaMemoryTypeDevice || at.device != dev) {
      cudaGetLastError();
      ptrs[dev] = nullptr;
    }
  }
  if (!ptrs[dev]) {
    FeatBest* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_feat_best);
    if (e != cudaSuccess || !p)
      return fail(CB_ERR_CUDA, "cb_split_scan_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return CB_OK;
}
}  // namespace cb

extern "C" {

int cb_split_scan_noalloc(const int64_t* hist, const int32_t* bin_ptr, int f, const float* cuts, const float* min_vals,
                          const int64_t* totals, const cb_tree_params* params, const uint8_t* feature_mask,
                          cb_split* out, void* stream) {
  if (!hist || !bin_ptr || !cuts || !min_vals || !totals || !params || !out || f <= 0)
    return fail(CB_ERR_INVALID, "cb_split_scan_noalloc: bad arguments");
  if (f > CB_NOALLOC_MAX_FEATURES)
    return fail(CB_ERR_UNSUPPORTED, "cb_split_scan_noalloc: %d features exceed the built-in scratch of %d; use cb_split_scan with a workspace",
                f, CB_NOALLOC_MAX_FEATURES);
  FeatBest* fbuf = nullptr;
  int rc = noalloc_scratch(&fbuf);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // every record in [0, f) is written by the first kernel before the second reads it, so nothing
  // left over from an earlier call (possibly with more features) can be picked up
  scan_single_kernel<<<ceil_div(f, 4), 128, 0, s>>>(hist, bin_ptr, f, totals, *params, feature_mask, fbuf);
  pick_single_kernel<<<1, 1, 0, s>>>(fbuf, f, cuts, min_vals, totals, out);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K6
int cb_partition_workspace_bytes(int64_t count, size_t* bytes) {
  if (!bytes || count < 0) return fail(CB_ERR_INVALID, "cb_partition_workspace_bytes: bad arguments");
  Bump b(nullptr);
  b.take<uint32_t>((size_t)(count / 32 + 2));
  b.take<int>((size_t)(count / kPartTile + 2));
  *bytes = b.used();
  return CB_OK;
}

int cb_partition_stable(const uint8_t* bins, int64_t ldb, const uint32_t* seg_in, int64_t count, int feature,
                        int split_bin, int default_left, uint32_t* seg_out, int64_t* n_left, void* ws,
                        size_t ws_bytes, void* stream) {
  if (!bins || !n_left || !ws || count < 0 || feature < 0 || feature >= ldb || count > 2147483647LL ||
      (count > 0 && (!seg_in || !seg_out)))
    return fail(CB_ERR_INVALID, "cb_partition_stable: bad arguments");
  size_t need = 0;
  cb_partition_workspace_bytes(count, &need);
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_partition_stable: workspace %zu < %zu bytes", ws_bytes, need);
  cudaStream_t s = (cudaStream_t)stream;
  if (count == 0) {
    CB_CUDA(cudaMemsetAsync(n_left, 0, sizeof(int64_t), s));
    return CB_OK;
  }
  Bump b(ws);
  uint32_t* flag_bits = b.take<uint32_t>((size_t)(count / 32 + 2));
  int* tile_cnt = b.take<int>((size_t)(count / kPartTile + 2));
  const int nt = ceil_div(count, kPartTile);
  part1_flags_kernel<<<nt, 256, 0, s>>>(make_view(bins, 0, ldb, 0), seg_in, count, feature, split_bin, default_left, flag_bits, tile_cnt);
  part2_scan_kernel<<<1, 1024, 0, s>>>(tile_cnt, nt, n_left);
  part3_scatter_kernel<<<nt, 256, 0, s>>>(seg_in, count, flag_bits, tile_cnt, n_left, seg_out);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K7
int cb_leaf_update(float* margin, const uint32_t* row_idx, int64_t count, float leaf_value, void* stream) {
  if (!margin || count < 0) return fail(CB_ERR_INVALID, "cb_leaf_update: bad arguments");
  if (count == 0) return CB_OK;
  leaf_update_kernel<<<ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(margin, row_idx, count, leaf_value);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- K8
int cb_predict(const float* X, int64_t n, int f, int64_t ldx, const uint64_t* nodes, const int32_t* tree_off,
               int n_trees, float base_margin, float* out_margin, float* out_proba, void* stream) {
  if (n == 0) return CB_OK;
  if (!X || !nodes || !tree_off || n < 0 || f <= 0 || f > 32767 || ldx < f || n_trees < 0 || (!out_margin && !out_proba))
    return fail(CB_ERR_INVALID, "cb_predict: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  const int smem = f * kPredRows * (int)sizeof(float);
  if (smem > di.smem_optin)
    return fail(CB_ERR_UNSUPPORTED, "cb_predict: %d features need %d bytes of shared memory (> %d)", f, smem, di.smem_optin);
  static thread_local int attr_dev = -1, attr_smem = 0;
  if (attr_dev != di.device || attr_smem < smem) {
    CB_CUDA(cudaFuncSetAttribute(predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_dev = di.device; attr_smem = smem;
  }
  predict_kernel<<<ceil_div(n, kPredRows), kPredRows, smem, (cudaStream_t)stream>>>(
      X, n, f, ldx, reinterpret_cast<const uint2*>(nodes), tree_off, n_trees, base_margin, out_margin, out_proba);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// trees per shared-memory chunk for the complete-tree scorer: two buffers next to the row tile
static int heap_chunk_trees(const DeviceInfo& di, int f, int depth, int* rows_out) {
  const int NI = (1 << depth) - 1, NL = 1 << depth;
  const int per_tree = NI * 8 + NL * 4;
  // two CTAs per SM (one stages its row tile while the other walks trees) when 128-row tiles fit
  static const int env_rows = [] { const char* e = getenv("COBALT_PRED_ROWS"); return e ? atoi(e) : 128; }();   // read once
  int rows = env_rows;
  if (rows != 256 && rows != 128 && rows != 64 && rows != 32) rows = 128;
  for (;;) {
    const int64_t tile = (((int64_t)f * rows * 4 + 127) & ~(int64_t)127);
    const int per_sm = rows <= 128 ? 2 : 1;
    const int64_t avail = (int64_t)(di.smem_optin + 1024) / per_sm - tile - 1280 - (per_sm > 1 ? 1024 : 0);
    int chunk = avail > 0 ? (int)(avail / 2 / per_tree) & ~7 : 0;
    if (chunk >= 8 || rows <= 32) {
      *rows_out = rows;
      return chunk > 512 ? 512 : chunk;
    }
    if (rows == 128) { rows = 256; continue; }      // 128 did not fit twice: one CTA per SM with 256 rows
    if (rows == 256) { rows = 64; continue; }
    rows >>= 1;
  }
}

int cb_predict_heap_plan(int n_features, int depth, int* chunk_trees, int* tile_rows) {
  if (!chunk_trees || !tile_rows || n_features <= 0 || depth < 1 || depth > 12) return fail(CB_ERR_INVALID, "cb_predict_heap_plan: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  int rows = 0;
  *chunk_trees = heap_chunk_trees(di, n_features, depth, &rows);
  *tile_rows = rows;
  if (*chunk_trees < 8) return fail(CB_ERR_UNSUPPORTED, "cb_predict_heap: %d features at depth %d do not fit shared memory", n_features, depth);
  return CB_OK;
}

int cb_predict_heap(const float* X, int64_t n, int f, int64_t ldx, const uint64_t* nodes, const float* leaf,
                    int n_trees, int depth, float base_margin, float* out_margin, float* out_proba, void* stream) {
  if (n == 0) return CB_OK;
  if (!X || !nodes || !leaf || n < 0 || f <= 0 || f > 32767 || ldx < f || n_trees < 0 || depth < 1 || depth > 12 ||
      (!out_margin && !out_proba))
    return fail(CB_ERR_INVALID, "cb_predict_heap: bad arguments");
  if ((reinterpret_cast<uintptr_t>(nodes) | reinterpret_cast<uintptr_t>(leaf)) & 15)
    return fail(CB_ERR_INVALID, "cb_predict_heap: forest arrays must be 16-byte aligned");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  int rows = 0;
  const int chunk = heap_chunk_trees(di, f, depth, &rows);
  if (chunk < 8) return fail(CB_ERR_UNSUPPORTED, "cb_predict_heap: %d features at depth %d do not fit shared memory", f, depth);
  const int NI = (1 << depth) - 1, NL = 1 << depth;
  const int64_t tile = (((int64_t)f * rows * 4 + 127) & ~(int64_t)127);
  const int smem = (int)(tile + 2 * (int64_t)chunk * (NI * 8 + NL * 4) + 128);
  static thread_local int attr_dev = -1;
  if (attr_dev != di.device) {
    CB_CUDA(cudaFuncSetAttribute(predict_heap_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, di.smem_optin - 1024));
    CB_CUDA(cudaFuncSetAttribute(predict_heap_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, di.smem_optin - 1024));
    attr_dev = di.device;
  }
  static const int env_ilp = [] { const char* e = getenv("COBALT_PRED_ILP"); return e ? atoi(e) : 8; }();          // read once
  if (env_ilp == 4)
    predict_heap_kernel<4><<<ceil_div(n, rows), rows, smem, (cudaStream_t)stream>>>(
        X, n, f, ldx, reinterpret_cast<const uint2*>(nodes), leaf, n_trees, depth, chunk, base_margin, out_margin, out_proba);
  else
    predict_heap_kernel<8><<<ceil_div(n, rows), rows, smem, (cudaStream_t)stream>>>(
        X, n, f, ldx, reinterpret_cast<const uint2*>(nodes), leaf, n_trees, depth, chunk, base_margin, out_margin, out_proba);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// quantised-input scoring: CTAs of 128 rows; as many per SM as shared memory allows (tile + staging are fixed,
// the two tree-chunk buffers take what is left, >= 8 trees each)
constexpr int kQ8Ilp = 16;                 // trees in flight per thread
static int q8_chunk_trees(const DeviceInfo& di, int f, int depth, int64_t ldc, int* smem_bytes) {
  (void)ldc;
  const int64_t TB = (int64_t)8 << depth;
  const int64_t planes = ((int64_t)2 * f * 128 + 127) & ~(int64_t)127;
  const int64_t per_sm = (int64_t)di.smem_optin + 1024;          // opt-in per-CTA limit + its 1 KB reserve = the SM's carve-out
  for (int k = 5; k >= 1; --k) {                                 // CTAs per SM: the walk is latency bound, warps count
    const int64_t avail = per_sm / k - 1024 - 256 - planes;
    int chunk = (int)std::min<int64_t>(avail / (kQ8Stages * TB), 64) & ~(kQ8Ilp - 1);
    if (chunk >= kQ8Ilp) {
      if (smem_bytes) *smem_bytes = (int)(planes + kQ8Stages * chunk * TB + 8 * kQ8Stages);
      return chunk;
    }
  }
  return 0;
}

static int q8_plane_base(const DeviceInfo& di, uint32_t* base) {
  static std::mutex mu;
  static uint32_t cached[64];
  static bool have[64];
  std::lock_guard<std::mutex> lock(mu);
  const int d = di.device & 63;
  if (!have[d]) {
    smem_base_probe_kernel<<<1, 1, 1024>>>();
    cudaError_t e = cudaMemcpyFromSymbol(&cached[d], g_smem_base_probe, sizeof(uint32_t));   // synchronises; no allocation
    if (e != cudaSuccess) return fail(CB_ERR_CUDA, "cb_predict_q8_plan: probe failed: %s", cudaGetErrorString(e));
    have[d] = true;
  }
  *base = cached[d];
  return CB_OK;
}

int cb_predict_q8_plan(int n_features, int depth, int64_t ldc, int* chunk_trees, uint32_t* plane_base) {
  if (!chunk_trees || !plane_base || n_features <= 0 || n_features > 255 || depth < 1 || depth > 10 || ldc < n_features ||
      (ldc & 3) || !((ldc >> 2) & 1))
    return fail(CB_ERR_INVALID, "cb_predict_q8_plan: need 1..255 features, depth 1..10 and ldc = 4 * odd >= n_features");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  if ((rc = q8_plane_base(di, plane_base))) return rc;
  *chunk_trees = q8_chunk_trees(di, n_features, depth, ldc, nullptr);
  if (*chunk_trees < kQ8Ilp || (*plane_base & 127u) || *plane_base + 2u * (uint32_t)n_features * 128u > 0x10000u)
    return fail(CB_ERR_UNSUPPORTED, "cb_predict_q8: %d features at depth %d do not fit shared memory", n_features, depth);
  return CB_OK;
}

int cb_bin_rows_u8(const float* X, int64_t n, int f, int64_t ldx, const float* cuts, const int32_t* n_cuts, uint8_t* codes,
                   int64_t ldc, void* stream) {
  if (n == 0) return CB_OK;
  if (!X || !cuts || !n_cuts || !codes || n < 0 || f <= 0 || ldx < f || ldc < f)
    return fail(CB_ERR_INVALID, "cb_bin_rows_u8: bad arguments");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  const int slabs = ceil_div(f, 32);
  // about 4 CTAs per SM and slab, at least 256 rows each: the table load is amortised over the CTA's rows
  int64_t rows_per_cta = std::max<int64_t>(256, ceil_div(n, (int64_t)di.sm_count * 4));
  rows_per_cta = (rows_per_cta + 31) & ~(int64_t)31;
  bin_rows_q8_kernel<<<dim3((unsigned)ceil_div(n, rows_per_cta), slabs), 256, 0, (cudaStream_t)stream>>>(
      X, n, f, ldx, cuts, n_cuts, codes, ldc, (int)rows_per_cta);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_predict_q8(const uint8_t* codes, int64_t n, int f, int64_t ldc, const uint32_t* forest, int n_trees, int depth,
                  float base_margin, uint32_t plane_base, float* out_margin, float* out_proba, void* stream) {
  if (n == 0) return CB_OK;
  if (!codes || !forest || n < 0 || f <= 0 || f > 255 || ldc < f || (ldc & 3) || !((ldc >> 2) & 1) || n_trees < 0 || depth < 1 ||
      depth > 10 || (!out_margin && !out_proba))
    return fail(CB_ERR_INVALID, "cb_predict_q8: bad arguments");
  if ((reinterpret_cast<uintptr_t>(codes) | reinterpret_cast<uintptr_t>(forest)) & 15)
    return fail(CB_ERR_INVALID, "cb_predict_q8: codes and forest must be 16-byte aligned");
  DeviceInfo di;
  int rc = device_info(&di);
  if (rc) return rc;
  int smem = 0;
  const int chunk = q8_chunk_trees(di, f, depth, ldc, &smem);
  if (chunk < 8) return fail(CB_ERR_UNSUPPORTED, "cb_predict_q8: %d features at depth %d do not fit shared memory", f, depth);
  uint32_t want_base = 0;
  if ((rc = q8_plane_base(di, &want_base))) return rc;
  if (want_base != plane_base)
    return fail(CB_ERR_INVALID, "cb_predict_q8: plane_base %u is not this device's (%u): lay the forest out with cb_predict_q8_plan on this device",
                plane_base, want_base);
  auto launch = [&](auto kernel) -> int {
    CB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, di.smem_optin - 1024));
    kernel<<<ceil_div(n, 128), 128, smem, (cudaStream_t)stream>>>(codes, n, f, ldc, forest, n_trees, chunk, base_margin,
                                                                  plane_base, out_margin, out_proba);
    CB_LAUNCH_CHECK();
    return CB_OK;
  };
  switch (depth) {
    case 1: return launch(predict_q8_kernel<kQ8Ilp, 1>);
    case 2: return launch(predict_q8_kernel<kQ8Ilp, 2>);
    case 3: return launch(predict_q8_kernel<kQ8Ilp, 3>);
    case 4: return launch(predict_q8_kernel<kQ8Ilp, 4>);
    case 5: return launch(predict_q8_kernel<kQ8Ilp, 5>);
    case 6: return launch(predict_q8_kernel<kQ8Ilp, 6>);
    case 7: return launch(predict_q8_kernel<kQ8Ilp, 7>);
    case 8: return launch(predict_q8_kernel<kQ8Ilp, 8>);
    case 9: return launch(predict_q8_kernel<kQ8Ilp, 9>);
    default: return launch(predict_q8_kernel<kQ8Ilp, 10>);
  }
}

// ---------------------------------------------------------------------------- K10 TreeSHAP
int cb_tree_shap_workspace_bytes(int64_t n_rows, int n_features, int n_trees, int64_t n_nodes, size_t* bytes) {
  if (!bytes || n_rows < 0 || n_features <= 0 || n_trees < 0 || n_nodes < 0) return fail(CB_ERR_INVALID, "cb_tree_shap_workspace_bytes: bad arguments");
  Bump b(nullptr);
  b.take<float>((size_t)n_rows * n_trees * (n_features + 1));
  b.take<float>((size_t)n_nodes);
  b.take<float>((size_t)n_trees);
  *bytes = b.used();
  return CB_OK;
}

int cb_tree_shap(const float* X, int64_t n, int f, int64_t ldx, const int64_t* tree_off, int n_trees, int64_t n_nodes,
                 const int32_t* left, const int32_t* right, const int32_t* split_index, const float* split_cond,
                 const uint8_t* default_left, const float* sum_hess, float base_margin, int max_depth, float* out,
                 void* workspace, size_t workspace_bytes, void* stream) {
  if (n == 0) return CB_OK;
  if (!X || !tree_off || !left || !right || !split_index || !split_cond || !default_left || !sum_hess || !out || !workspace ||
      n < 0 || f <= 0 || ldx < f || n_trees < 0 || n_nodes < 0)
    return fail(CB_ERR_INVALID, "cb_tree_shap: bad arguments");
  if (max_depth > kShapMaxDepth) return fail(CB_ERR_UNSUPPORTED, "cb_tree_shap: trees deeper than %d levels", kShapMaxDepth);
  size_t need = 0;
  cb_tree_shap_workspace_bytes(n, f, n_trees, n_nodes, &need);
  if (workspace_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_tree_shap: workspace %zu < %zu bytes", workspace_bytes, need);
  cudaStream_t s = (cudaStream_t)stream;
  Bump b(workspace);
  float* ws = b.take<float>((size_t)n * n_trees * (f + 1));
  float* node_mean = b.take<float>((size_t)n_nodes);
  float* tree_mean = b.take<float>((size_t)n_trees);
  ShapForest fo{tree_off, left, right, split_index, split_cond, default_left, sum_hess};
  CB_CUDA(cudaMemsetAsync(ws, 0, sizeof(float) * (size_t)n * n_trees * (f + 1), s));
  if (n_trees > 0) {
    shap_means_kernel<<<ceil_div(n_trees, 128), 128, 0, s>>>(fo, n_trees, node_mean, tree_mean);
    shap_trees_kernel<<<(unsigned)ceil_div(n * n_trees, (int64_t)128), 128, 0, s>>>(X, n, f, ldx, fo, n_trees, ws);
  }
  shap_sum_kernel<<<(unsigned)ceil_div(n * (f + 1), (int64_t)256), 256, 0, s>>>(ws, tree_mean, n, f, n_trees, base_margin, out);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

// ---------------------------------------------------------------------------- trainer
struct cb_trainer_s {
  cb_train_config cfg;
  cb_train_buffers buf;
  DeviceInfo di;
  int max_nodes, S, S2, grid, tiles, smem, n_part_tiles, pool_slots, lb_slots;
  int res_flags, res_scatter, res_final;   // CTAs of the level-tail kernels that are resident at once (whole GPU)
  int2* gpair;
  uint32_t* rows[2];
  int cur;
  bool identity;
  uint32_t* flag_bits;
  int* tile_cnt;
  int* group_cnt;
  TreeState* st;
  LevelSegs* segs;
  HistPlan* plan;
  NodeArrays na;
  int64_t* pool;
  int64_t* level_buf;
  int32_t* partials;
  FeatBest* feat_best;
  int tree_index;
  const uint8_t* fmask;
  int64_t launches;
  bool timing;
  cudaEvent_t ev[3];
  // per-launch timeline of one eager round (cb_trainer_profile_*): an event after every launch
  bool profiling;
  int n_marks;
  cudaEvent_t marks[96];
  const char* mark_names[96];
  // conflict-free histogram path
  bool cf;
  int cf_R, cf_smem;
  CfLayout* cf_lay;
  int32_t* cf_slot_of;
  int32_t* cf_fold_src;
  int32_t* cf_fold_dst;
  int32_t* feat_pos;      // [f] byte position of a feature's column inside a binned row
  int32_t* feat_code;     // [f] 0 = plain feature, m + 1 = member m of a group column (bin 1 <=> column code == m + 1)
  bool grouped;
  int part_words;         // int32 words per partial histogram
  // peer mode (row-sharded training over NVLink peer memory)
  PeerCtx pc;
  SymHeader** d_peers;
  int64_t** d_peer_lb;
};

static int trainer_layout(const cb_train_config* cfg, const DeviceInfo& di, void* ws, cb_trainer_s* t, size_t* bytes) {
  const int D = cfg->tree.max_depth;
  const int64_t n = cfg->n_rows;
  const int S = cfg->total_bins;
  const int max_nodes = (1 << (D + 1)) - 1;
  const int pool_slots = (1 << D) - 1;
  const int lb_slots = D >= 2 ? (1 << (D - 2)) : 1;
  const int lvl_nodes = 1 << (D - 1);
  Bump b(ws);
  int2* gpair = b.take<int2>((size_t)n);
  uint32_t* r0 = b.take<uint32_t>((size_t)n);
  uint32_t* r1 = b.take<uint32_t>((size_t)n);
  uint32_t* flag_bits = b.take<uint32_t>((size_t)(n / 32 + 2));
  int* tile_cnt = b.take<int>((size_t)(n / kPartTile + 2));
  int* group_cnt = b.take<int>((size_t)(n / kPartTile / kTileGroup + 2));
  TreeState* st = b.take<TreeState>(1);
  LevelSegs* segs = b.take<LevelSegs>(1);
  HistPlan* plan = b.take<HistPlan>(1);
  NodeArrays na;
  na.begin = b.take<int>((size_t)max_nodes);
  na.count = b.take<int>((size_t)max_nodes);
  na.G = b.take<long long>((size_t)max_nodes);
  na.H = b.take<long long>((size_t)max_nodes);
  int64_t* pool = b.take<int64_t>((size_t)pool_slots * 2 * S);
  int64_t* level_buf = b.take<int64_t>((size_t)lb_slots * 2 * S + 2);
  const int cf_words = hist_budget_slots(di);                 // upper bound of R * 32
  const int part_words = 2 * std::max(S, cf_words);
  int32_t* partials = b.take<int32_t>((size_t)max_partials(n, di.sm_count) * part_words);
  FeatBest* feat_best = b.take<FeatBest>((size_t)lvl_nodes * cfg->n_features);
  CfLayout* cf_lay = b.take<CfLayout>(1);
  int32_t* cf_slot_of = b.take<int32_t>((size_t)cf_words);
  int32_t* cf_fold_src = b.take<int32_t>((size_t)cf_words);
  int32_t* cf_fold_dst = b.take<int32_t>((size_t)cf_words);
  int32_t* feat_pos = b.take<int32_t>((size_t)cfg->n_features);
  int32_t* feat_code = b.take<int32_t>((size_t)cfg->n_features);
  SymHeader** d_peers = b.take<SymHeader*>(kMaxRanks);
  int64_t** d_peer_lb = b.take<int64_t*>(kMaxRanks);
  if (bytes) *bytes = b.used();
  if (t) {
    t->max_nodes = max_nodes; t->S = S; t->S2 = 2 * S; t->pool_slots = pool_slots; t->lb_slots = lb_slots;
    t->gpair = gpair; t->rows[0] = r0; t->rows[1] = r1; t->flag_bits = flag_bits; t->tile_cnt = tile_cnt;
    t->group_cnt = group_cnt;
    t->st = st; t->segs = segs; t->plan = plan; t->na = na; t->pool = pool; t->level_buf = level_buf;
    t->partials = partials; t->feat_best = feat_best;
    t->cf_lay = cf_lay; t->cf_slot_of = cf_slot_of; t->cf_fold_src = cf_fold_src; t->cf_fold_dst = cf_fold_dst;
    t->feat_pos = feat_pos; t->feat_code = feat_code; t->part_words = part_words;
    t->d_peers = d_peers; t->d_peer_lb = d_peer_lb;
  }
  return CB_OK;
}

static int trainer_check_cfg(const cb_train_config* cfg) {
  if (!cfg) return fail(CB_ERR_INVALID, "trainer: null config");
  if (cfg->n_rows <= 0 || cfg->n_rows > 2147483647LL) return fail(CB_ERR_INVALID, "trainer: n_rows out of range");
  if (cfg->n_features <= 0 || cfg->ldb <= 0 || (cfg->ldb & 3) || cfg->n_groups < 0 || (cfg->n_groups == 0 && cfg->ldb < cfg->n_features))
    return fail(CB_ERR_INVALID, "trainer: ldb must be a multiple of 4 and hold every column");
  if (cfg->tree.max_depth < 1 || cfg->tree.max_depth > kMaxDepth)
    return fail(CB_ERR_UNSUPPORTED, "trainer: max_depth must be in [1, %d]", kMaxDepth);
  if (cfg->total_bins <= 0 || cfg->total_bins > cfg->n_features * CB_MAX_BINS)
    return fail(CB_ERR_INVALID, "trainer: total_bins out of range");
  if (cfg->max_trees <= 0) return fail(CB_ERR_INVALID, "trainer: max_trees must be positive");
  return CB_OK;
}

int cb_trainer_workspace_bytes(const cb_train_config* cfg, size_t* bytes) {
  int rc = trainer_check_cfg(cfg);
  if (rc) return rc;
  if (!bytes) return fail(CB_ERR_INVALID, "cb_trainer_workspace_bytes: null output");
  DeviceInfo di;
  rc = device_info(&di);
  if (rc) return rc;
  return trainer_layout(cfg, di, nullptr, nullptr, bytes);
}

int cb_trainer_create(const cb_train_config* cfg, const cb_train_buffers* buf, void* ws, size_t ws_bytes,
                      cb_trainer** out) {
  int rc = trainer_check_cfg(cfg);
  if (rc) return rc;
  if (!buf || !ws || !out) return fail(CB_ERR_INVALID, "cb_trainer_create: null argument");
  if (!buf->bins || !buf->y || !buf->margin || !buf->bin_ptr || !buf->cuts || !buf->min_vals || !buf->left ||
      !buf->right || !buf->parent || !buf->split_index || !buf->split_bin || !buf->split_cond || !buf->default_left ||
      !buf->base_weight || !buf->loss_chg || !buf->sum_hess || !buf->n_nodes)
    return fail(CB_ERR_INVALID, "cb_trainer_create: a required buffer is null");
  DeviceInfo di;
  rc = device_info(&di);
  if (rc) return rc;
  size_t need = 0;
  trainer_layout(cfg, di, nullptr, nullptr, &need);
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_trainer_create: workspace %zu < %zu bytes", ws_bytes, need);
  cb_trainer_s* t = new (std::nothrow) cb_trainer_s();
  if (!t) return fail(CB_ERR_INVALID, "cb_trainer_create: out of host memory");
  t->cfg = *cfg; t->buf = *buf; t->di = di;
  trainer_layout(cfg, di, ws, t, nullptr);
  const int budget = hist_budget_slots(di);
  t->tiles = hist_tiles_bound(t->S, budget);
  if (t->tiles > kMaxTiles) { delete t; return fail(CB_ERR_UNSUPPORTED, "trainer: too many bins for %d tiles", kMaxTiles); }
  t->smem = std::min(budget, t->S) * 8;
  t->grid = std::min(di.sm_count, kMaxGrid);
  t->n_part_tiles = ceil_div(cfg->n_rows, kPartTile);
  {
    // persistent level-tail grids = what the GPU holds at once (the serial extra CTA included): one CTA more and a
    // tile CTA waits for the extra CTA before it even starts (tools/level_trace.py: + 4 us per kernel)
    int a = 0, b = 0, c = 0;
    CB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, part_level_flags_kernel, 256, 0));
    CB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, part_level_scatter_kernel, 256, 0));
    CB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c, final_apply_kernel, 256, 0));
    t->res_flags = std::max(2, a * t->di.sm_count);
    t->res_scatter = std::max(2, b * t->di.sm_count);
    t->res_final = std::max(2, c * t->di.sm_count);
  }
  t->cur = 0; t->identity = true; t->tree_index = 0; t->fmask = nullptr; t->launches = 0; t->timing = false; t->profiling = false; t->n_marks = 0;
  memset(&t->pc, 0, sizeof t->pc); t->pc.world = 1;
  plan_tiles_kernel<<<1, 1>>>(buf->bin_ptr, cfg->n_features, budget, t->plan);
  if (const char* mr = getenv("COBALT_HIST_MIN_ROWS")) plan_set_min_rows_kernel<<<1, 1>>>(t->plan, atoi(mr));
  init_state_kernel<<<1, 1>>>(t->st);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { delete t; return fail(CB_ERR_CUDA, "cb_trainer_create: %s", cudaGetErrorString(e)); }
  // column placement -> conflict-free shared-memory tables (host side, one-off)
  {
    const int f = cfg->n_features;
    std::vector<int32_t> bp((size_t)f + 1), pf((size_t)cfg->ldb), fp((size_t)f, -1), fc((size_t)f, 0);
    e = cudaMemcpy(bp.data(), buf->bin_ptr, sizeof(int32_t) * (f + 1), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && buf->pos_feature)
      e = cudaMemcpy(pf.data(), buf->pos_feature, sizeof(int32_t) * cfg->ldb, cudaMemcpyDeviceToHost);
    const int ng = cfg->n_groups;
    std::vector<int32_t> gptr((size_t)ng + 1, 0), gfeat;
    if (ng < 0 || (ng > 0 && (!buf->grp_ptr || !buf->grp_feat || !buf->pos_feature))) {
      delete t; return fail(CB_ERR_INVALID, "trainer: n_groups > 0 needs grp_ptr, grp_feat and pos_feature");
    }
    if (e == cudaSuccess && ng > 0) {
      e = cudaMemcpy(gptr.data(), buf->grp_ptr, sizeof(int32_t) * (ng + 1), cudaMemcpyDeviceToHost);
      if (e == cudaSuccess) {
        if (gptr[0] != 0 || gptr[ng] < ng || gptr[ng] > f) { delete t; return fail(CB_ERR_INVALID, "trainer: grp_ptr is malformed"); }
        gfeat.resize((size_t)gptr[ng]);
        e = cudaMemcpy(gfeat.data(), buf->grp_feat, sizeof(int32_t) * gfeat.size(), cudaMemcpyDeviceToHost);
      }
    }
    if (e != cudaSuccess) { delete t; return fail(CB_ERR_CUDA, "cb_trainer_create: %s", cudaGetErrorString(e)); }
    if (!buf->pos_feature)
      for (int64_t p = 0; p < cfg->ldb; ++p) pf[p] = p < f ? (int32_t)p : -1;
    if (bp[f] != cfg->total_bins) { delete t; return fail(CB_ERR_INVALID, "trainer: total_bins does not match bin_ptr"); }
    GroupTable gt;
    gt.n_groups = ng; gt.ptr = gptr.data(); gt.feat = gfeat.data();
    {
      std::vector<int32_t> nc((size_t)f);
      for (int i = 0; i < f; ++i) nc[i] = bp[i + 1] - bp[i];
      std::vector<int> group_of;
      if (!check_groups(nc.data(), f, gt, &group_of)) { delete t; return fail(CB_ERR_INVALID, "trainer: malformed feature groups"); }
    }
    for (int64_t p = 0; p < cfg->ldb; ++p) {
      if (pf[p] >= 0 && pf[p] < f) fp[pf[p]] = (int32_t)p;
      else if (is_group_key(pf[p]) && group_of_key(pf[p]) < ng) {
        const int g = group_of_key(pf[p]);
        for (int k = gptr[g]; k < gptr[g + 1]; ++k) { fp[gfeat[k]] = (int32_t)p; fc[gfeat[k]] = k - gptr[g] + 1; }
      }
    }
    for (int i = 0; i < f; ++i)
      if (fp[i] < 0) { delete t; return fail(CB_ERR_INVALID, "trainer: pos_feature does not place feature %d", i); }
    t->grouped = ng > 0;
    CfHostTables tb;
    t->cf = build_cf_tables(bp.data(), pf.data(), f, gt, cfg->ldb, budget * 8, &tb);
    const bool multi = (cfg->ldb + 127) / 128 > 1;
    if (t->cf && multi && !cfg->blocked) t->cf = false;     // row-major rows wider than one block: tiled kernel
    if (!t->cf && multi && cfg->blocked) { delete t; return fail(CB_ERR_UNSUPPORTED, "trainer: block-major bins need a conflict-free layout"); }
    if (!t->cf && ng > 0) { delete t; return fail(CB_ERR_UNSUPPORTED, "trainer: grouped columns need the conflict-free layout"); }
    e = cudaMemcpy(t->feat_code, fc.data(), sizeof(int32_t) * f, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { delete t; return fail(CB_ERR_CUDA, "cb_trainer_create: %s", cudaGetErrorString(e)); }
    e = cudaMemcpy(t->feat_pos, fp.data(), sizeof(int32_t) * f, cudaMemcpyHostToDevice);
    if (t->cf && e == cudaSuccess) {
      t->cf_R = tb.lay.R; t->cf_smem = tb.lay.R * 32 * 8;
      e = cudaMemcpy(t->cf_lay, &tb.lay, sizeof(CfLayout), cudaMemcpyHostToDevice);
      if (e == cudaSuccess) e = cudaMemcpy(t->cf_slot_of, tb.slot_of.data(), sizeof(int32_t) * tb.slot_of.size(), cudaMemcpyHostToDevice);
      if (e == cudaSuccess && tb.lay.n_fold) {
        e = cudaMemcpy(t->cf_fold_src, tb.fold_src.data(), sizeof(int32_t) * tb.fold_src.size(), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(t->cf_fold_dst, tb.fold_dst.data(), sizeof(int32_t) * tb.fold_dst.size(), cudaMemcpyHostToDevice);
      }
      if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_accum_cf_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCfSmemBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(hist_accum_cf_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCfSmemBytes);
    }
    if (e != cudaSuccess) { delete t; return fail(CB_ERR_CUDA, "cb_trainer_create: %s", cudaGetErrorString(e)); }
  }
  *out = t;
  return CB_OK;
}

void cb_trainer_destroy(cb_trainer* tr) {
  if (tr && tr->timing)
    for (int i = 0; i < 3; ++i) cudaEventDestroy(tr->ev[i]);
  delete tr;
}

/* profile one eager round: begin -> run a round with the per-step calls -> end returns names and ms */
int cb_trainer_profile_begin(cb_trainer* t) {
  if (!t) return fail(CB_ERR_INVALID, "trainer: null handle");
  if (!t->profiling) {
    for (int i = 0; i < 96; ++i) CB_CUDA(cudaEventCreate(&t->marks[i]));
    t->profiling = true;
  }
  t->n_marks = 0;
  return CB_OK;
}

int cb_trainer_profile_end_sync(cb_trainer* t, float* ms, const char** names, int cap, int* n) {
  if (!t || !t->profiling || !ms || !names || !n) return fail(CB_ERR_INVALID, "trainer: profiling is not active");
  if (t->n_marks == 0) { *n = 0; return CB_OK; }
  CB_CUDA(cudaEventSynchronize(t->marks[t->n_marks - 1]));
  int k = 0;
  for (int i = 1; i < t->n_marks && k < cap; ++i, ++k) {
    CB_CUDA(cudaEventElapsedTime(&ms[k], t->marks[i - 1], t->marks[i]));
    names[k] = t->mark_names[i];
  }
  *n = k;
  for (int i = 0; i < 96; ++i) cudaEventDestroy(t->marks[i]);
  t->profiling = false;
  return CB_OK;
}

int cb_trainer_enable_timing(cb_trainer* t, int on) {
  if (!t) return fail(CB_ERR_INVALID, "trainer: null handle");
  if (on && !t->timing) {
    for (int i = 0; i < 3; ++i) CB_CUDA(cudaEventCreate(&t->ev[i]));
    t->timing = true;
  } else if (!on && t->timing) {
    for (int i = 0; i < 3; ++i) cudaEventDestroy(t->ev[i]);
    t->timing = false;
  }
  return CB_OK;
}

int cb_trainer_last_hist_timing_sync(cb_trainer* t, float* ms_accum, float* ms_reduce, int64_t* rows, int* n_build) {
  if (!t || !t->timing) return fail(CB_ERR_INVALID, "trainer: timing is not enabled");
  CB_CUDA(cudaEventSynchronize(t->ev[2]));
  if (ms_accum) CB_CUDA(cudaEventElapsedTime(ms_accum, t->ev[0], t->ev[1]));
  if (ms_reduce) CB_CUDA(cudaEventElapsedTime(ms_reduce, t->ev[1], t->ev[2]));
  int h[2] = {0, 0};
  CB_CUDA(cudaMemcpy(&h[0], &t->plan->n_build, sizeof(int), cudaMemcpyDeviceToHost));
  CB_CUDA(cudaMemcpy(&h[1], &t->plan->total_rows, sizeof(int), cudaMemcpyDeviceToHost));
  if (n_build) *n_build = h[0];
  if (rows) *rows = h[1];
  return CB_OK;
}

static inline void mark(cb_trainer_s* t, cudaStream_t s, const char* name) {
  if (!t->profiling || t->n_marks >= 96) return;
  cudaEventRecord(t->marks[t->n_marks], s);
  t->mark_names[t->n_marks++] = name;
}

static TreeOut tree_out(cb_trainer_s* t) {
  const size_t o = 0;   // row 0: kernels add the device-side tree index
  TreeOut to;
  to.left = t->buf.left + o; to.right = t->buf.right + o; to.parent = t->buf.parent + o;
  to.split_index = t->buf.split_index + o; to.split_bin = t->buf.split_bin + o; to.split_cond = t->buf.split_cond + o;
  to.default_left = t->buf.default_left + o; to.base_weight = t->buf.base_weight + o; to.loss_chg = t->buf.loss_chg + o;
  to.sum_hess = t->buf.sum_hess + o;
  return to;
}

int cb_trainer_round_begin(cb_trainer* t, int tree_index, const uint8_t* feature_masks, void* stream) {
  if (!t) return fail(CB_ERR_INVALID, "trainer: null handle");
  if (tree_index >= t->cfg.max_trees) return fail(CB_ERR_INVALID, "trainer: tree_index %d out of range", tree_index);
  cudaStream_t s = (cudaStream_t)stream;
  StageRange nv("cobalt:grad");
  t->fmask = feature_masks; t->identity = true; t->cur = 0;
  const int64_t n = t->cfg.n_rows;
  if (tree_index >= 0) {                 // negative: keep the device-side counter (CUDA-graph replay)
    set_tree_index_kernel<<<1, 1, 0, s>>>(t->st, tree_index);
    t->launches += 1;
  }
  mark(t, s, "start");
  int g = (int)std::min<int64_t>(ceil_div(n, 256), (int64_t)t->di.sm_count * 8);
  // root_tot is zero here: init_state_kernel / the previous round's round_end_kernel cleared it
  CB_CUDA(launch_dep(grad_kernel, dim3(g), dim3(256), 0, s, t->buf.margin, t->buf.y, t->buf.weight, n, t->cfg.scale_pos_weight,
                     ldexpf(1.0f, t->cfg.tree.sg), ldexpf(1.0f, t->cfg.tree.sh), t->cfg.keep_thresh, t->cfg.seed, 0u,
                     t->cfg.row_offset, t->gpair, reinterpret_cast<unsigned long long*>(t->st->root_tot),
                     (const int*)&t->st->tree_index));
  mark(t, s, "grad");
  CB_CUDA(launch_dep(round_init_kernel, dim3(ceil_div(t->max_nodes, 256)), dim3(256), 0, s, t->st, t->na, tree_out(t),
                     t->max_nodes, (int)n, t->plan, t->grid));
  mark(t, s, "round_init");
  CB_LAUNCH_CHECK();
  t->launches += 2;
  return CB_OK;
}

int cb_trainer_level_hist(cb_trainer* t, int level, void* stream) {
  if (!t || level < 0 || level >= t->cfg.tree.max_depth) return fail(CB_ERR_INVALID, "trainer: bad level");
  cudaStream_t s = (cudaStream_t)stream;
  const int nb = level == 0 ? 1 : (1 << (level - 1));
  StageRange nv("cobalt:hist");
  if (t->cf) {
    if (t->timing) CB_CUDA(cudaEventRecord(t->ev[0], s));
    const BinsView bv = make_view(t->buf.bins, t->cfg.n_rows, t->cfg.ldb, t->cfg.blocked);
    auto kern = t->identity ? hist_accum_cf_kernel<4, true> : hist_accum_cf_kernel<4, false>;
    CB_CUDA(launch_dep(kern, dim3(t->grid), dim3(1024), (size_t)kCfSmemBytes, s, t->buf.bins,
                       (int64_t)(bv.blocked ? t->cfg.n_rows : 0), bv.blocked ? bv.last_w : (int)t->cfg.ldb, t->gpair,
                       t->rows[t->cur], t->plan, t->cf_lay, t->cf_fold_src, t->cf_fold_dst, t->partials, t->cf_slot_of,
                       t->S, t->level_buf, t->plan->done, &t->st->peer_timeout, t->st->root_tot,
                       level == 0 ? t->level_buf + t->S2 : nullptr));
    if (t->timing) { CB_CUDA(cudaEventRecord(t->ev[1], s)); CB_CUDA(cudaEventRecord(t->ev[2], s)); }
    mark(t, s, "hist_accum+reduce");
    t->launches += 1;
  } else {
    if (t->timing) CB_CUDA(cudaEventRecord(t->ev[0], s));
    int rc = launch_hist_accum(t->buf.bins, t->cfg.ldb, reinterpret_cast<const int32_t*>(t->gpair), t->rows[t->cur],
                               t->identity ? 1 : 0, t->plan, t->buf.bin_ptr, t->cfg.n_features, t->S, t->grid, t->tiles,
                               t->smem, t->partials, s);
    if (rc) return rc;
    if (t->timing) CB_CUDA(cudaEventRecord(t->ev[1], s));
    hist_reduce_kernel<<<dim3(ceil_div(t->S, 256), nb), 256, 0, s>>>(t->partials, t->plan, t->S, t->level_buf);
    if (t->timing) CB_CUDA(cudaEventRecord(t->ev[2], s));
    t->launches += 2;
    if (level == 0) {
      root_totals_kernel<<<1, 1, 0, s>>>(t->st, t->level_buf + t->S2);
      t->launches += 1;
    }
  }
  if (t->pc.world > 1) {                  // peers may now read this rank's level buffer
    CB_CUDA(launch_dep(peer_signal_kernel, dim3(1), dim3(32), 0, s, t->pc, (const TreeState*)t->st, level, 0));
    t->launches += 1;
    mark(t, s, "signal");
  }
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_trainer_level_buffer(cb_trainer* t, int level, int64_t** ptr, int64_t* n_values) {
  if (!t || level < 0 || level >= t->cfg.tree.max_depth || !ptr || !n_values) return fail(CB_ERR_INVALID, "trainer: bad level");
  const int nb = level == 0 ? 1 : (1 << (level - 1));
  *ptr = t->level_buf;
  *n_values = (int64_t)nb * t->S2 + (level == 0 ? 2 : 0);
  return CB_OK;
}

int cb_trainer_level_split(cb_trainer* t, int level, void* stream) {
  if (!t || level < 0 || level >= t->cfg.tree.max_depth) return fail(CB_ERR_INVALID, "trainer: bad level");
  cudaStream_t s = (cudaStream_t)stream;
  const int D = t->cfg.tree.max_depth;
  const int ln = 1 << level;
  const int n = (int)t->cfg.n_rows;
  const int n_groups = ceil_div(t->n_part_tiles, kTileGroup);
  const int fx = t->pc.world > 1 ? t->pc.f1 - t->pc.f0 : t->cfg.n_features;   // peer mode: this rank's feature slice
  StageRange nv(level + 1 < D ? "cobalt:scan+partition" : "cobalt:scan+leaf");
  {                                         // a rank whose slice is empty still needs the kernel's housekeeping CTA
    CB_CUDA(launch_dep(scan_level_fused_kernel, dim3(fx > 0 ? fx : 1, ln), dim3(128), 0, s,
                       t->level_buf, t->pool, t->plan, t->buf.bin_ptr, t->cfg.n_features, t->S, t->st, level, t->na, t->cfg.tree,
                       t->fmask, t->feat_best, level == 0 ? t->level_buf + t->S2 : nullptr, t->st, t->group_cnt, n_groups, t->pc,
                       t->grouped ? t->feat_code : nullptr));
    t->launches += 1;
  }
  mark(t, s, "scan");
  if (t->pc.world > 1) {
    CB_CUDA(launch_dep(peer_publish_kernel, dim3(1), dim3(256), 0, s, t->pc, (const TreeState*)t->st, level,
                       (const FeatBest*)t->feat_best, t->cfg.n_features));
    t->launches += 1;
    mark(t, s, "publish");
  }
  LevelCtx lc;
  lc.st = t->st; lc.level = level; lc.max_depth = D; lc.f = t->cfg.n_features; lc.feat_best = t->feat_best;
  lc.p = t->cfg.tree; lc.na = t->na; lc.feat_pos = t->feat_pos; lc.feat_code = t->grouped ? t->feat_code : nullptr; lc.pc = t->pc;
  const BinsView view = make_view(t->buf.bins, t->cfg.n_rows, t->cfg.ldb, t->cfg.blocked);
  const uint32_t* rin = t->rows[t->cur];
  const int ident = t->identity ? 1 : 0;
  if (level + 1 < D) {
    uint32_t* rout = t->rows[t->cur ^ 1];
    const int pgrid = std::min(t->n_part_tiles, t->res_flags - 1) + 1;
    const int sgrid = std::min(t->n_part_tiles, t->res_scatter - 1) + 1;
    CB_CUDA(launch_dep(part_level_flags_kernel, dim3(pgrid), dim3(256), 0, s, view, t->buf.bins_cm, rin, ident, n,
                       t->n_part_tiles, lc, t->st, t->buf.cuts, t->buf.min_vals, tree_out(t), t->segs, t->flag_bits,
                       t->tile_cnt, t->group_cnt));
    mark(t, s, "flags+finish");
    CB_CUDA(launch_dep(part_level_scatter_kernel, dim3(sgrid), dim3(256), 0, s, rin, ident, n, t->n_part_tiles,
                       (const LevelSegs*)t->segs, (const uint32_t*)t->flag_bits, (const int*)t->tile_cnt,
                       (const int*)t->group_cnt, rout, t->buf.margin, t->st, level, D, t->na, t->plan, t->grid,
                       t->cfg.world_size > 1 ? 1 : 0));
    mark(t, s, "scatter+plan");
    t->cur ^= 1; t->identity = false;
    t->launches += 2;
  } else {
    const int nblk = ceil_div(n, 1024);
    CB_CUDA(launch_dep(final_apply_kernel, dim3(std::min(nblk, t->res_final - 1) + 1), dim3(256), 0, s, view, t->buf.bins_cm,
                       rin, ident, n, nblk, lc, t->st, t->buf.cuts, t->buf.min_vals, tree_out(t), t->segs, t->buf.margin));
    mark(t, s, "final_apply");
    t->launches += 1;
  }
  CB_LAUNCH_CHECK();
  return CB_OK;
}

int cb_trainer_round_end(cb_trainer* t, void* stream) {
  if (!t) return fail(CB_ERR_INVALID, "trainer: null handle");
  CB_CUDA(launch_dep(round_end_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, t->st, t->buf.n_nodes));
  CB_LAUNCH_CHECK();
  t->launches += 1;
  return CB_OK;
}

int cb_trainer_boost_round(cb_trainer* t, int tree_index, const uint8_t* feature_mask, void* stream) {
  int rc = cb_trainer_round_begin(t, tree_index, feature_mask, stream);
  if (rc) return rc;
  for (int level = 0; level < t->cfg.tree.max_depth; ++level) {
    if ((rc = cb_trainer_level_hist(t, level, stream))) return rc;
    if ((rc = cb_trainer_level_split(t, level, stream))) return rc;
  }
  return cb_trainer_round_end(t, stream);
}

// ---------------------------------------------------------------------------- peer memory (CUDA IPC)
int cb_sym_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  if (!ptr || !handle64 || bytes == 0) return fail(CB_ERR_INVALID, "cb_sym_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  CB_CUDA(cudaMalloc(&p, bytes));
  CB_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { cudaFree(p); return fail(CB_ERR_CUDA, "cb_sym_alloc: cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); }
  memcpy(handle64, &h, 64);
  *ptr = p;
  return CB_OK;
}
int cb_sym_open(const unsigned char* handle64, void** ptr) {
  if (!ptr || !handle64) return fail(CB_ERR_INVALID, "cb_sym_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  CB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return CB_OK;
}
int cb_sym_close(void* ptr) { if (ptr) CB_CUDA(cudaIpcCloseMemHandle(ptr)); return CB_OK; }
int cb_sym_free(void* ptr) { if (ptr) CB_CUDA(cudaFree(ptr)); return CB_OK; }

static size_t sym_lb_offset() { return (sizeof(SymHeader) + 255) & ~size_t(255); }

int cb_trainer_sym_bytes(const cb_train_config* cfg, size_t* bytes) {
  int rc = trainer_check_cfg(cfg);
  if (rc) return rc;
  if (!bytes) return fail(CB_ERR_INVALID, "cb_trainer_sym_bytes: null output");
  const int D = cfg->tree.max_depth;
  const int lb_slots = D >= 2 ? (1 << (D - 2)) : 1;
  *bytes = sym_lb_offset() + sizeof(int64_t) * ((size_t)lb_slots * 2 * cfg->total_bins + 2) + 256;
  return CB_OK;
}

int cb_trainer_set_peers(cb_trainer* t, int rank, int world, void* const* regions) {
  if (!t || !regions || world < 2 || world > kMaxRanks || rank < 0 || rank >= world)
    return fail(CB_ERR_INVALID, "cb_trainer_set_peers: bad arguments (2 <= world <= %d)", kMaxRanks);
  if (world != t->cfg.world_size) return fail(CB_ERR_INVALID, "cb_trainer_set_peers: world differs from cb_train_config.world_size");
  SymHeader* hp[kMaxRanks] = {nullptr};
  int64_t* lp[kMaxRanks] = {nullptr};
  for (int r = 0; r < world; ++r) {
    if (!regions[r]) return fail(CB_ERR_INVALID, "cb_trainer_set_peers: region %d is null", r);
    hp[r] = reinterpret_cast<SymHeader*>(regions[r]);
    lp[r] = reinterpret_cast<int64_t*>(reinterpret_cast<char*>(regions[r]) + sym_lb_offset());
  }
  CB_CUDA(cudaMemcpy(t->d_peers, hp, sizeof hp, cudaMemcpyHostToDevice));
  CB_CUDA(cudaMemcpy(t->d_peer_lb, lp, sizeof lp, cudaMemcpyHostToDevice));
  // feature slices balanced by histogram slots
  const int f = t->cfg.n_features;
  std::vector<int32_t> bp((size_t)f + 1);
  CB_CUDA(cudaMemcpy(bp.data(), t->buf.bin_ptr, sizeof(int32_t) * (f + 1), cudaMemcpyDeviceToHost));
  int bounds[kMaxRanks + 1];
  bounds[0] = 0;
  int feat = 0;
  for (int r = 1; r <= world; ++r) {
    const long long target = (long long)bp[f] * r / world;
    while (feat < f && bp[feat + 1] <= target) ++feat;
    bounds[r] = r == world ? f : feat;
  }
  t->pc.world = world; t->pc.rank = rank; t->pc.f0 = bounds[rank]; t->pc.f1 = bounds[rank + 1];
  t->pc.self = hp[rank]; t->pc.peers = t->d_peers; t->pc.peer_level_buf = t->d_peer_lb;
  t->level_buf = lp[rank];
  return CB_OK;
}

int cb_trainer_peer_status(cb_trainer* t, int* timed_out) {
  if (!t || !timed_out) return fail(CB_ERR_INVALID, "cb_trainer_peer_status: bad arguments");
  unsigned v = 0;
  CB_CUDA(cudaMemcpy(&v, &t->st->peer_timeout, sizeof v, cudaMemcpyDeviceToHost));
  *timed_out = (int)v;
  return CB_OK;
}

int cb_trainer_row_order(cb_trainer* t, const uint32_t** ptr) {
  if (!t || !ptr) return fail(CB_ERR_INVALID, "trainer: null argument");
  *ptr = t->identity ? nullptr : t->rows[t->cur];
  return CB_OK;
}

int cb_trainer_node_hist(cb_trainer* t, int node_id, const int64_t** ptr) {
  if (!t || !ptr || node_id < 0 || node_id >= t->pool_slots) return fail(CB_ERR_INVALID, "trainer: node id out of range");
  *ptr = t->pool + (size_t)node_id * t->S2;
  return CB_OK;
}

int64_t cb_trainer_launch_count(cb_trainer* t) { return t ? t->launches : 0; }

int cb_trainer_layout_info(cb_trainer* t, int* conflict_free, int* smem_rows, int* smem_bytes) {
  if (!t) return fail(CB_ERR_INVALID, "trainer: null handle");
  if (conflict_free) *conflict_free = t->cf ? 1 : 0;
  if (smem_rows) *smem_rows = t->cf ? t->cf_R : 0;
  if (smem_bytes) *smem_bytes = t->cf ? kCfSmemBytes : t->smem;
  return CB_OK;
}

#ifdef CB_LEVEL_TRACE
// debug build only (tools/level_trace.py): [6 kernels][2048 CTAs][start, end] %globaltimer stamps of the last launches
int cb_debug_level_trace(unsigned long long* out_host, int clear) {
  CB_CUDA(cudaDeviceSynchronize());
  CB_CUDA(cudaMemcpyFromSymbol(out_host, g_level_trace, sizeof(unsigned long long) * 6 * 2048 * 2));
  if (clear) {
    void* p = nullptr;
    CB_CUDA(cudaGetSymbolAddress(&p, g_level_trace));
    CB_CUDA(cudaMemset(p, 0, sizeof(unsigned long long) * 6 * 2048 * 2));
  }
  return CB_OK;
}
#endif

}  // extern "C"
