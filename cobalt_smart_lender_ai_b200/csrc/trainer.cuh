// Level-wise tree growth with every decision on the device.  Included by cobalt_gbdt.cu only.
#pragma once
#include "kernels.cuh"

namespace cb {

constexpr int kMaxDepth = 9;              // levels 0..8 are split; level nodes <= 256
constexpr int kMaxLevelNodes = 256;

struct TreeState {                        // device memory
  int n_nodes;
  int lvl_first[kMaxDepth + 2];
  int lvl_count[kMaxDepth + 2];
  int pad_;
  long long root_tot[2];                  // local sum of (q_g, q_h), filled by the gradient kernel
  unsigned int done[2 * (kMaxDepth + 2)]; // per level: CTAs finished in the scan / flags kernels (last one runs the epilogue)
  unsigned long long best_key[kMaxLevelNodes];  // per level node: max over features of (chg bits << 32 | 0xFFFF - feature)
};

// the level's nodes in id order (= ascending position order): what the partition pass needs
struct LevelSegs {
  int n;
  int begin[kMaxLevelNodes], count[kMaxLevelNodes];
  int kind[kMaxLevelNodes];               // 0 leaf now, 1 split + partition, 2 split, children are leaves
  int feat[kMaxLevelNodes], bin[kMaxLevelNodes], dl[kMaxLevelNodes], child[kMaxLevelNodes];
  float leafL[kMaxLevelNodes], leafR[kMaxLevelNodes];
  int Lb[kMaxLevelNodes], nleft[kMaxLevelNodes];
};

struct NodeArrays {                       // device arrays [max_nodes]
  int* begin; int* count; long long* G; long long* H;
};

struct TreeOut {                          // slices of the forest arrays for the tree being grown
  int32_t* left; int32_t* right; int32_t* parent; int32_t* split_index; int32_t* split_bin;
  float* split_cond; uint8_t* default_left; float* base_weight; float* loss_chg; float* sum_hess;
};

__global__ void round_init_kernel(TreeState* st, NodeArrays na, TreeOut to, int max_nodes, int n_rows,
                                  HistPlan* plan, int grid) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < max_nodes) {
    to.left[i] = -1; to.right[i] = -1; to.parent[i] = 2147483647; to.split_index[i] = 0; to.split_bin[i] = 0;
    to.split_cond[i] = 0.0f; to.default_left[i] = 0; to.base_weight[i] = 0.0f; to.loss_chg[i] = 0.0f;
    to.sum_hess[i] = 0.0f;
  }
  if (i == 0) {
    st->n_nodes = 1;
    for (int l = 0; l < kMaxDepth + 2; ++l) { st->lvl_first[l] = 0; st->lvl_count[l] = 0; st->done[2 * l] = 0; st->done[2 * l + 1] = 0; }
    st->lvl_first[0] = 0; st->lvl_count[0] = 1;
    for (int k = 0; k < kMaxLevelNodes; ++k) st->best_key[k] = 0ull;
    na.begin[0] = 0; na.count[0] = n_rows;
    int b0 = 0, c0 = n_rows;
    plan->b_node[0] = 0; plan->b_sibling[0] = -1; plan->b_parent[0] = -1; plan->lvl_src[0] = 0;
    plan_ctas(plan, 1, &b0, &c0, grid);
  }
}

// level buffer tail (level 0 only): the local root totals, summed over ranks with the histogram
__global__ void root_totals_kernel(const TreeState* st, int64_t* level_buf_tail) {
  level_buf_tail[0] = st->root_tot[0];
  level_buf_tail[1] = st->root_tot[1];
}

// pool[node] = built; pool[sibling] = pool[parent] - built   (K9, exact)
__global__ void __launch_bounds__(256) hist_finalize_kernel(const int64_t* __restrict__ level_buf,
                                                            const HistPlan* __restrict__ plan, int S2,
                                                            int64_t* __restrict__ pool, NodeArrays na,
                                                            const int64_t* tail) {
  const int j = blockIdx.y;
  if (j >= plan->n_build) return;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (tail && j == 0 && v == 0) { na.G[0] = tail[0]; na.H[0] = tail[1]; }
  if (v >= S2) return;
  const int64_t built = level_buf[(size_t)j * S2 + v];
  pool[(size_t)plan->b_node[j] * S2 + v] = built;
  const int sib = plan->b_sibling[j];
  if (sib >= 0) pool[(size_t)sib * S2 + v] = pool[(size_t)plan->b_parent[j] * S2 + v] - built;
}

__global__ void __launch_bounds__(128) scan_level_kernel(const int64_t* __restrict__ pool,
                                                         const int32_t* __restrict__ bin_ptr, int f, int S2,
                                                         const TreeState* __restrict__ st, int level, NodeArrays na,
                                                         cb_tree_params p, const uint8_t* __restrict__ feature_mask,
                                                         FeatBest* __restrict__ feat_best) {
  const int i = blockIdx.y;
  if (i >= st->lvl_count[level]) return;
  const int feat = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (feat >= f) return;
  const int nid = st->lvl_first[level] + i;
  FeatBest fb;
  fb.chg = 0.0f; fb.split_bin = -1; fb.default_left = 0; fb.pad = 0; fb.GL = 0; fb.HL = 0;
  if (!feature_mask || feature_mask[feat]) {
    const int b0 = __ldg(bin_ptr + feat), nb = __ldg(bin_ptr + feat + 1) - b0;
    fb = scan_feature(pool + (size_t)nid * S2 + 2 * (size_t)b0, nb, na.G[nid], na.H[nid], p);
  }
  if ((threadIdx.x & 31) == 0) feat_best[(size_t)i * f + feat] = fb;
}

// K9 + K5 fused, one CTA of 128 threads per (level node, feature), 2 bins per thread:
// reads the node's histogram from the level buffer (built child) or as parent - built (sibling),
// stores it in the pool for the next level, and scans it (exact int64 prefix sums, f64 gains).
__device__ __forceinline__ unsigned long long cand_key(float chg, int tiebreak) {
  return chg > 0.0f ? ((unsigned long long)__float_as_uint(chg) << 32) | (unsigned)tiebreak : 0ull;
}
struct FinishArgs {
  int max_depth;
  const float* cuts;
  const float* min_vals;
  TreeOut to;
  LevelSegs* segs;
  const int32_t* feat_pos;
};
struct FinishSmem { float chg[kMaxLevelNodes]; int feat[kMaxLevelNodes]; int child[kMaxLevelNodes]; };

// runs in ONE CTA (any multiple of 32 threads) after every (node, feature) candidate of the level is
// written; reads candidates with L2 loads because other CTAs of the same kernel produced them
__device__ void level_finish_body(TreeState* st, int level, const FeatBest* feat_best, int f, const cb_tree_params& p,
                                  NodeArrays na, const FinishArgs& fa, FinishSmem& sm, long long G0, long long H0) {
  float* s_chg = sm.chg; int* s_feat = sm.feat; int* s_child = sm.child;
  const int max_depth = fa.max_depth;
  const float* cuts = fa.cuts; const float* min_vals = fa.min_vals;
  TreeOut to = fa.to; LevelSegs* segs = fa.segs; const int32_t* feat_pos = fa.feat_pos;
  const int first = st->lvl_first[level], cnt = st->lvl_count[level];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  if (level == 0 && threadIdx.x == 0) { na.G[0] = G0; na.H[0] = H0; }
  __syncthreads();
  const double ig = ldexp(1.0, -p.sg), ih = ldexp(1.0, -p.sh);
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const unsigned long long key = __ldcg(&st->best_key[i]);
    st->best_key[i] = 0ull;                                   // ready for the next level
    const float bc = key ? __uint_as_float((unsigned)(key >> 32)) : 0.0f;
    const int bf = key ? 0xFFFF - (int)(key & 0xFFFFu) : -1;
    s_chg[i] = bc; s_feat[i] = bf;
    bool valid = false;
    if (bf >= 0) {
      const long long HL = __ldcg(&feat_best[(size_t)i * f + bf].HL), HR = na.H[first + i] - HL;
      valid = bc > 1e-6f && !(bc < p.gamma) && level < max_depth && HL != 0 && HR != 0;
    }
    s_child[i] = valid ? 1 : 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (level == 0) {
      to.base_weight[0] = __double2float_rn(calc_weight(__dmul_rn((double)na.G[0], ig), __dmul_rn((double)na.H[0], ih), p));
      to.sum_hess[0] = __double2float_rn(__dmul_rn((double)na.H[0], ih));
    }
    int next = st->n_nodes;
    const int old = next;
    for (int i = 0; i < cnt; ++i) {
      if (s_child[i]) { s_child[i] = next; next += 2; } else s_child[i] = -1;
    }
    st->n_nodes = next;
    st->lvl_first[level + 1] = old;
    st->lvl_count[level + 1] = next - old;
    segs->n = cnt;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int nid = first + i;
    segs->begin[i] = na.begin[nid];
    segs->count[i] = na.count[nid];
    segs->Lb[i] = 0; segs->nleft[i] = 0;
    const int L = s_child[i];
    if (L < 0) {
      const float leaf = __fmul_rn(to.base_weight[nid], p.eta);
      to.split_cond[nid] = leaf;
      to.base_weight[nid] = leaf;
      segs->kind[i] = 0; segs->leafL[i] = leaf; segs->leafR[i] = leaf;
      segs->feat[i] = 0; segs->bin[i] = 0; segs->dl[i] = 0; segs->child[i] = -1;
      continue;
    }
    const int R = L + 1, bf = s_feat[i];
    const FeatBest* fbp = feat_best + (size_t)i * f + bf;
    FeatBest fb;
    fb.split_bin = __ldcg(&fbp->split_bin); fb.default_left = __ldcg(&fbp->default_left);
    fb.GL = __ldcg(&fbp->GL); fb.HL = __ldcg(&fbp->HL);
    const long long GL = fb.GL, HL = fb.HL, GR = na.G[nid] - GL, HR = na.H[nid] - HL;
    to.left[nid] = L; to.right[nid] = R; to.parent[L] = nid; to.parent[R] = nid;
    to.split_index[nid] = bf; to.split_bin[nid] = fb.split_bin;
    to.split_cond[nid] = fb.split_bin >= 0 ? cuts[(size_t)bf * CB_CUT_STRIDE + fb.split_bin] : min_vals[bf];
    to.default_left[nid] = (uint8_t)fb.default_left;
    to.loss_chg[nid] = s_chg[i];
    na.G[L] = GL; na.H[L] = HL; na.G[R] = GR; na.H[R] = HR;
    const float bwL = __double2float_rn(calc_weight(__dmul_rn((double)GL, ig), __dmul_rn((double)HL, ih), p));
    const float bwR = __double2float_rn(calc_weight(__dmul_rn((double)GR, ig), __dmul_rn((double)HR, ih), p));
    to.sum_hess[L] = __double2float_rn(__dmul_rn((double)HL, ih));
    to.sum_hess[R] = __double2float_rn(__dmul_rn((double)HR, ih));
    segs->feat[i] = feat_pos[bf]; segs->bin[i] = fb.split_bin; segs->dl[i] = fb.default_left; segs->child[i] = L;
    if (level + 1 == max_depth) {
      const float lL = __fmul_rn(bwL, p.eta), lR = __fmul_rn(bwR, p.eta);
      to.split_cond[L] = lL; to.base_weight[L] = lL;
      to.split_cond[R] = lR; to.base_weight[R] = lR;
      segs->kind[i] = 2; segs->leafL[i] = lL; segs->leafR[i] = lR;
    } else {
      to.base_weight[L] = bwL; to.base_weight[R] = bwR;
      segs->kind[i] = 1; segs->leafL[i] = 0.0f; segs->leafR[i] = 0.0f;
    }
  }
}

__global__ void __launch_bounds__(128) scan_level_fused_kernel(const int64_t* __restrict__ level_buf,
                                                               int64_t* __restrict__ pool,
                                                               const HistPlan* __restrict__ plan,
                                                               const int32_t* __restrict__ bin_ptr, int f, int S,
                                                               const TreeState* __restrict__ st, int level,
                                                               NodeArrays na, cb_tree_params p,
                                                               const uint8_t* __restrict__ feature_mask,
                                                               FeatBest* __restrict__ feat_best,
                                                               const int64_t* __restrict__ tail, TreeState* stw,
                                                               FinishArgs fa) {
  __shared__ long long sg[4], sh[4];
  __shared__ unsigned long long skf[4], skb[4];
  __shared__ FinishSmem fsm;
  __shared__ int s_last;
  const int i = blockIdx.y;
  const int cnt_level = st->lvl_count[level];
  if (i >= cnt_level) return;
  const int feat = blockIdx.x;
  const int nid = st->lvl_first[level] + i;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  long long Gt, Ht;
  if (level == 0) {
    Gt = tail[0]; Ht = tail[1];
  } else {
    Gt = na.G[nid]; Ht = na.H[nid];
  }
  do {
  if (feature_mask && !feature_mask[feat]) {
    if (tid == 0) {
      FeatBest fb; fb.chg = 0.0f; fb.split_bin = -1; fb.default_left = 0; fb.pad = 0; fb.GL = 0; fb.HL = 0;
      feat_best[(size_t)i * f + feat] = fb;
    }
    break;
  }
  const int b0 = __ldg(bin_ptr + feat), nb = __ldg(bin_ptr + feat + 1) - b0;
  const int src = plan->lvl_src[i], j = src >> 1;
  const longlong2* lb = reinterpret_cast<const longlong2*>(level_buf) + (size_t)j * S + b0;
  longlong2* mine = reinterpret_cast<longlong2*>(pool) + (size_t)nid * S + b0;
  const longlong2* par = (src & 1) ? reinterpret_cast<const longlong2*>(pool) + (size_t)plan->b_parent[j] * S + b0 : nullptr;
  longlong2 v[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int b = 2 * tid + k;
    v[k] = make_longlong2(0, 0);
    if (b < nb) {
      v[k] = lb[b];
      if (par) { const longlong2 q = par[b]; v[k].x = q.x - v[k].x; v[k].y = q.y - v[k].y; }
      mine[b] = v[k];
    }
  }
  // block-wide inclusive scan of the per-thread sums
  long long tg = v[0].x + v[1].x, th = v[0].y + v[1].y;
  long long ig = tg, ih = th;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long a = __shfl_up_sync(0xffffffffu, ig, o), b = __shfl_up_sync(0xffffffffu, ih, o);
    if (lane >= o) { ig += a; ih += b; }
  }
  if (lane == 31) { sg[wid] = ig; sh[wid] = ih; }
  __syncthreads();
  long long wg = 0, wh = 0, Gf = 0, Hf = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    if (w < wid) { wg += sg[w]; wh += sh[w]; }
    Gf += sg[w]; Hf += sh[w];
  }
  const long long exg = wg + ig - tg, exh = wh + ih - th;     // sum of all bins before 2 * tid
  const double dg = ldexp(1.0, -p.sg), dh = ldexp(1.0, -p.sh);
  const float root_gain = __double2float_rn(calc_gain(__dmul_rn((double)Gt, dg), __dmul_rn((double)Ht, dh), p));
  // forward (missing -> right): candidate b has left = prefix_incl(b); scan order ascending
  float fc = 0.0f; int fbin = -1;
  long long pg = exg, ph = exh;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int b = 2 * tid + k;
    pg += v[k].x; ph += v[k].y;
    if (b < nb) {
      const float c = split_chg(pg, ph, Gt - pg, Ht - ph, dg, dh, root_gain, p);
      if (c > fc) { fc = c; fbin = b; }
    }
  }
  // backward (missing -> left): candidate b has right = sum of bins >= b; scan order descending
  float bc = 0.0f; int bbin = -1;
  const bool has_missing = (Gf != Gt) || (Hf != Ht);
  if (has_missing) {
#pragma unroll
    for (int k = 1; k >= 0; --k) {
      const int b = 2 * tid + k;
      if (b < nb) {
        const long long eg = exg + (k ? v[0].x : 0), eh = exh + (k ? v[0].y : 0);   // prefix_excl(b)
        const long long gr = Gf - eg, hr = Hf - eh;
        const float c = split_chg(Gt - gr, Ht - hr, gr, hr, dg, dh, root_gain, p);
        if (c > bc) { bc = c; bbin = b; }
      }
    }
  }
  // block arg-max: forward ties -> lower bin, backward ties -> higher bin (first in scan order)
  unsigned long long kf = cand_key(fc, 0xFFFF - fbin), kb = cand_key(bc, bbin);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, kf, o), b = __shfl_xor_sync(0xffffffffu, kb, o);
    kf = a > kf ? a : kf; kb = b > kb ? b : kb;
  }
  if (lane == 0) { skf[wid] = kf; skb[wid] = kb; }
  __syncthreads();
  kf = max(max(skf[0], skf[1]), max(skf[2], skf[3]));
  kb = max(max(skb[0], skb[1]), max(skb[2], skb[3]));
  const bool use_b = (unsigned)(kb >> 32) > (unsigned)(kf >> 32);       // backward only when strictly better
  const unsigned long long kw = use_b ? kb : kf;
  if (kw == 0ull) {
    if (tid == 0) {
      FeatBest fb; fb.chg = 0.0f; fb.split_bin = -1; fb.default_left = 0; fb.pad = 0; fb.GL = 0; fb.HL = 0;
      feat_best[(size_t)i * f + feat] = fb;
    }
    break;
  }
  const int cbin = use_b ? (int)(kw & 0xFFFFu) : 0xFFFF - (int)(kw & 0xFFFFu);   // candidate bin
  const int sbin = use_b ? cbin - 1 : cbin;                                       // rows with code <= sbin go left
  // the thread that owns bin sbin (or thread 0 when sbin == -1) knows prefix_incl(sbin)
  const int owner = sbin >= 0 ? sbin >> 1 : 0;
  if (tid == owner) {
    long long gl = 0, hl = 0;
    if (sbin >= 0) { gl = exg + v[0].x + ((sbin & 1) ? v[1].x : 0); hl = exh + v[0].y + ((sbin & 1) ? v[1].y : 0); }
    if (use_b) { gl += Gt - Gf; hl += Ht - Hf; }
    FeatBest fb;
    fb.chg = __uint_as_float((unsigned)(kw >> 32)); fb.split_bin = sbin; fb.default_left = use_b ? 1 : 0; fb.pad = 0;
    fb.GL = gl; fb.HL = hl;
    feat_best[(size_t)i * f + feat] = fb;
    __threadfence();
    atomicMax(&stw->best_key[i], ((kw >> 32) << 32) | (unsigned long long)(0xFFFF - feat));
  }
  } while (0);
  // epilogue: the last CTA of the level to finish picks the splits and grows the tree (was a kernel of its own)
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned prev = atomicAdd(&stw->done[2 * level], 1u);
    s_last = prev == (unsigned)(gridDim.x * cnt_level) - 1u;
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    level_finish_body(stw, level, feat_best, f, p, na, fa, fsm, Gt, Ht);
  }
}

// one block: pick each node's best feature, allocate children in node-id order (R9),
// write the tree arrays, the children's totals and the level's segment table

// segment of a position: largest i with begin[i] <= pos, valid when pos < begin[i] + count[i]
__device__ __forceinline__ int find_seg(const int* __restrict__ s_begin, const int* __restrict__ s_count, int n,
                                        int pos) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (s_begin[mid] <= pos) lo = mid + 1; else hi = mid;
  }
  int i = lo - 1;
  if (i < 0 || pos >= s_begin[i] + s_count[i]) return -1;
  return i;
}

struct SegCache {
  int begin[kMaxLevelNodes], count[kMaxLevelNodes];
};

// packed per-segment split description cached in shared memory: feat | dl << 16 | kind << 24, split bin
struct SegSplit { int meta[kMaxLevelNodes]; int bin[kMaxLevelNodes]; };

// the code of column position `pos` of `row`: from the position-major copy when there is one
// (consecutive rows of a node are close together there: a level's gathers touch ~2^level bytes per
// row instead of a DRAM burst per row), else from the row/block-major matrix
__device__ __forceinline__ int code_at(const BinsView& v, const uint8_t* __restrict__ bins_cm, uint32_t row, int pos) {
  if (bins_cm) return __ldg(bins_cm + (size_t)pos * v.n_rows_cm + row);
  return __ldg(v.base + byte_index(v, row, pos));
}

// number of lefts strictly before position p (global prefix over all positions); L2 loads: the flag
// words and tile prefixes were written by other CTAs of the same kernel
__device__ __forceinline__ int lefts_before(const int* tile_ex, const uint32_t* flag_bits, int p, int n_rows) {
  if (p <= 0) return 0;
  int tile = p / kPartTile;
  int w1 = p >> 5, part_bits = p & 31;
  if (tile * kPartTile >= n_rows) {       // p == n_rows on a tile boundary: everything in the last tile
    tile -= 1; w1 = (n_rows + 31) >> 5; part_bits = 0;
  }
  int s = __ldcg(tile_ex + tile);
  for (int w = tile * (kPartTile / 32); w < w1; ++w) s += __popc(__ldcg(flag_bits + w));
  if (part_bits && w1 * 32 < n_rows) s += __popc(__ldcg(flag_bits + w1) & ((1u << part_bits) - 1u));
  return s;
}

// the same, computed by a whole warp (lanes stride over the tile's flag words)
__device__ __forceinline__ int lefts_before_warp(const int* tile_ex, const uint32_t* flag_bits, int p, int n_rows, int lane) {
  if (p <= 0) return 0;
  int tile = p / kPartTile;
  int w1 = p >> 5, part_bits = p & 31;
  if (tile * kPartTile >= n_rows) { tile -= 1; w1 = (n_rows + 31) >> 5; part_bits = 0; }
  int s = 0;
  for (int w = tile * (kPartTile / 32) + lane; w < w1; w += 32) s += __popc(__ldcg(flag_bits + w));
  if (lane == 0) {
    s += __ldcg(tile_ex + tile);
    if (part_bits && w1 * 32 < n_rows) s += __popc(__ldcg(flag_bits + w1) & ((1u << part_bits) - 1u));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}

struct ScanSmem {
  int warp_tot[32];
  int pick[kMaxLevelNodes], cbegin[kMaxLevelNodes], ccount[kMaxLevelNodes];
  int b_begin[kMaxBuild], b_count[kMaxBuild];
};

// P2, run by ONE CTA: exclusive scan of the tile counts, children segments, next level's histogram plan
__device__ void part_scan_body(int* tile_cnt, int n_tiles, int n_rows, const uint32_t* flag_bits, LevelSegs* segs,
                               TreeState* st, int level, int max_depth, NodeArrays na, HistPlan* plan, int grid,
                               int by_hessian, ScanSmem& sm) {
  const int nt = blockDim.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarp = nt >> 5;
  const int per = (n_tiles + nt - 1) / nt;
  const int b = min(n_tiles, (int)threadIdx.x * per), e = min(n_tiles, b + per);
  int s = 0;
  for (int i = b; i < e; ++i) s += __ldcg(tile_cnt + i);
  int inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) sm.warp_tot[wid] = inc;
  __syncthreads();
  int base = 0;
  for (int w = 0; w < wid; ++w) base += sm.warp_tot[w];
  int run = base + inc - s;
  for (int i = b; i < e; ++i) { const int c = __ldcg(tile_cnt + i); tile_cnt[i] = run; run += c; }
  __threadfence();
  __syncthreads();
  const int n = segs->n;
  // one warp per split node: both boundary prefixes
  for (int i = wid; i < n; i += nwarp) {
    if (segs->kind[i] != 1) { if (lane == 0) sm.pick[i] = -1; continue; }
    const int bg = segs->begin[i], ct = segs->count[i];
    const int Lb = lefts_before_warp(tile_cnt, flag_bits, bg, n_rows, lane);
    const int Le = lefts_before_warp(tile_cnt, flag_bits, bg + ct, n_rows, lane);
    if (lane == 0) {
      const int nl = Le - Lb;
      segs->Lb[i] = Lb; segs->nleft[i] = nl;
      const int L = segs->child[i];
      na.begin[L] = bg; na.count[L] = nl;
      na.begin[L + 1] = bg + nl; na.count[L + 1] = ct - nl;
      int pick = -1;
      if (level + 1 < max_depth) {
        const bool left_small = by_hessian ? (na.H[L] <= na.H[L + 1]) : (nl <= ct - nl);
        pick = (L << 1) | (left_small ? 0 : 1);
        sm.cbegin[i] = left_small ? bg : bg + nl;
        sm.ccount[i] = left_small ? nl : ct - nl;
      }
      sm.pick[i] = pick;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int nb = 0;
    const int first = st->lvl_first[level], first_next = st->lvl_first[level + 1];
    for (int i = 0; i < n; ++i) {
      if (sm.pick[i] < 0) continue;
      const int L = sm.pick[i] >> 1, R = L + 1;
      const bool left_small = (sm.pick[i] & 1) == 0;
      plan->b_node[nb] = left_small ? L : R; plan->b_sibling[nb] = left_small ? R : L; plan->b_parent[nb] = first + i;
      plan->lvl_src[L - first_next] = 2 * nb + (left_small ? 0 : 1);
      plan->lvl_src[R - first_next] = 2 * nb + (left_small ? 1 : 0);
      sm.b_begin[nb] = sm.cbegin[i];
      sm.b_count[nb] = sm.ccount[i];
      ++nb;
    }
    plan_ctas(plan, nb, sm.b_begin, sm.b_count, grid);
  }
}

// P1: go-left flags of every position that sits in a node being split (kind 1)
__global__ void __launch_bounds__(256) part_level_flags_kernel(BinsView view, const uint8_t* __restrict__ bins_cm,
                                                               const uint32_t* __restrict__ rows_in, int identity,
                                                               int n_rows, const LevelSegs* __restrict__ segs,
                                                               uint32_t* __restrict__ flag_bits,
                                                               int* __restrict__ tile_cnt, LevelSegs* segs_w,
                                                               TreeState* st, int level, int max_depth, NodeArrays na,
                                                               HistPlan* plan, int grid, int by_hessian) {
  __shared__ SegCache sc;
  __shared__ SegSplit ss;
  __shared__ ScanSmem scan_sm;
  __shared__ int wsum[8];
  __shared__ int s_last;
  const int n = segs->n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    sc.begin[i] = segs->begin[i]; sc.count[i] = segs->count[i];
    ss.meta[i] = segs->feat[i] | (segs->dl[i] << 16) | (segs->kind[i] << 24);
    ss.bin[i] = segs->bin[i];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int tile0 = blockIdx.x * kPartTile;
  constexpr int IT = kPartTile / 256;
  uint32_t row[IT];
  int seg[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) {            // all row indices first: 8 independent loads in flight
    const int pos = tile0 + wid * (kPartTile / 8) + it * 32 + lane;
    row[it] = 0; seg[it] = -1;
    if (pos < n_rows) {
      row[it] = identity ? (uint32_t)pos : __ldg(rows_in + pos);
      seg[it] = find_seg(sc.begin, sc.count, n, pos);
    }
  }
  int code[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) {            // then all gathers
    code[it] = -1;
    if (seg[it] >= 0 && (ss.meta[seg[it]] >> 24) == 1) code[it] = code_at(view, bins_cm, row[it], ss.meta[seg[it]] & 0xFFFF);
  }
  int cnt = 0;
#pragma unroll
  for (int it = 0; it < IT; ++it) {
    const int pos = tile0 + wid * (kPartTile / 8) + it * 32 + lane;
    bool fl = false;
    if (code[it] >= 0) {
      const int m = ss.meta[seg[it]];
      fl = code[it] == CB_MISSING_BIN ? ((m >> 16) & 1) != 0 : code[it] <= ss.bin[seg[it]];
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, fl);
    if (lane == 0 && pos < n_rows) flag_bits[pos >> 5] = bal;
    cnt += __popc(bal);
  }
  if (lane == 0) wsum[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += wsum[k];
    tile_cnt[blockIdx.x] = t;
    __threadfence();
    const unsigned prev = atomicAdd(&st->done[2 * level + 1], 1u);
    s_last = prev == gridDim.x - 1u;
  }
  __syncthreads();
  if (s_last) {            // the last tile to finish runs the scan + planning step (was a kernel of its own)
    __threadfence();
    part_scan_body(tile_cnt, (int)gridDim.x, n_rows, flag_bits, segs_w, st, level, max_depth, na, plan, grid, by_hessian,
                   scan_sm);
  }
}


// P3: stable scatter of split nodes (lefts first), leaf values of nodes that stopped here
__global__ void __launch_bounds__(256) part_level_scatter_kernel(const uint32_t* __restrict__ rows_in, int identity,
                                                                 int n_rows, const LevelSegs* __restrict__ segs,
                                                                 const uint32_t* __restrict__ flag_bits,
                                                                 const int* __restrict__ tile_ex,
                                                                 uint32_t* __restrict__ rows_out,
                                                                 float* __restrict__ margin) {
  __shared__ SegCache sc;
  const int n = segs->n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { sc.begin[i] = segs->begin[i]; sc.count[i] = segs->count[i]; }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int tile0 = blockIdx.x * kPartTile;
  const int w0 = tile0 >> 5, wpw = kPartTile / 8 / 32;
  int before = 0;
  for (int k = lane; k < wid * wpw; k += 32)
    if ((w0 + k) * 32 < n_rows) before += __popc(flag_bits[w0 + k]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  int run = tile_ex[blockIdx.x] + before;
  for (int it = 0; it < wpw; ++it) {
    const int pos = tile0 + wid * (kPartTile / 8) + it * 32 + lane;
    if (pos - lane >= n_rows) break;
    const uint32_t bal = flag_bits[pos >> 5];
    if (pos < n_rows) {
      const uint32_t row = identity ? (uint32_t)pos : __ldg(rows_in + pos);
      const int i = find_seg(sc.begin, sc.count, n, pos);
      int dest = pos;
      if (i >= 0) {
        const int kind = segs->kind[i];
        if (kind == 1) {
          const int lb = run + __popc(bal & ((1u << lane) - 1u)) - segs->Lb[i];   // lefts before, inside the node
          const int bg = sc.begin[i];
          dest = ((bal >> lane) & 1u) ? bg + lb : bg + segs->nleft[i] + (pos - bg - lb);
        } else if (kind == 0) {
          margin[row] = __fadd_rn(margin[row], segs->leafL[i]);
        }
      }
      rows_out[dest] = row;
    }
    run += __popc(bal);
  }
}

// last level: children are leaves, so rows are not moved; each row takes its child's leaf value
__global__ void __launch_bounds__(256) final_apply_kernel(BinsView view, const uint8_t* __restrict__ bins_cm,
                                                          const uint32_t* __restrict__ rows_in, int identity,
                                                          int n_rows, const LevelSegs* __restrict__ segs,
                                                          float* __restrict__ margin) {
  __shared__ SegCache sc;
  const int n = segs->n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { sc.begin[i] = segs->begin[i]; sc.count[i] = segs->count[i]; }
  __syncthreads();
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n_rows) return;
  const int i = find_seg(sc.begin, sc.count, n, pos);
  if (i < 0) return;
  const uint32_t row = identity ? (uint32_t)pos : __ldg(rows_in + pos);
  float leaf = segs->leafL[i];
  if (segs->kind[i] == 2) {
    const int c = code_at(view, bins_cm, row, segs->feat[i]);
    const bool left = c == CB_MISSING_BIN ? segs->dl[i] != 0 : c <= segs->bin[i];
    if (!left) leaf = segs->leafR[i];
  }
  margin[row] = __fadd_rn(margin[row], leaf);
}

__global__ void round_end_kernel(const TreeState* st, int32_t* n_nodes_out) { *n_nodes_out = st->n_nodes; }

}  // namespace cb
