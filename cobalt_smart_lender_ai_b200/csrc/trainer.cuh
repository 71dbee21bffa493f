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
  int tree_index;                         // forest row being grown; advanced by round_end (graph replay needs no host value)
  long long root_tot[2];                  // local sum of (q_g, q_h), filled by the gradient kernel
  unsigned int round_seq;                 // rounds finished so far (peer mode: base of the flag epochs)
  unsigned int peer_timeout;              // set when a wait on a peer flag gave up (never in a healthy run)
  unsigned long long best_key[kMaxLevelNodes];  // per level node: max over features of (chg bits << 32 | 0xFFFF - feature)
};

// the level's nodes in id order (= ascending position order): what the partition pass needs
struct LevelSegs {
  int n;
  int begin[kMaxLevelNodes], count[kMaxLevelNodes];
  int kind[kMaxLevelNodes];               // 0 leaf now, 1 split + partition, 2 split, children are leaves
  int feat[kMaxLevelNodes], bin[kMaxLevelNodes], dl[kMaxLevelNodes], child[kMaxLevelNodes];
  float leafL[kMaxLevelNodes], leafR[kMaxLevelNodes];
};

struct NodeArrays {                       // device arrays [max_nodes]
  int* begin; int* count; long long* G; long long* H;
};

struct TreeOut {                          // slices of the forest arrays for the tree being grown
  int32_t* left; int32_t* right; int32_t* parent; int32_t* split_index; int32_t* split_bin;
  float* split_cond; uint8_t* default_left; float* base_weight; float* loss_chg; float* sum_hess;
};

// ---------------------------------------------------------------------------------------------------
// Row-sharded training over NVLink peer memory (no NCCL on the critical path).  Every rank owns one
// "symmetric" region (same layout on every GPU, mapped into every peer with CUDA IPC):
//   flags[2][kMaxRanks]   epoch counters written BY peers: [0] "my level buffer is complete",
//                         [1] "my best candidates of this level are in your table"
//   tab[kMaxRanks][256]   best candidate of each level node, one row per source rank (written by peers)
//   level buffer          this rank's LOCAL histograms of the level's built nodes (+2 root totals)
// The scan kernel of rank r handles only r's slice of the features: it adds that slice of all ranks'
// level buffers (peer loads over NVLink) on the fly -- the collective is fused into the scan, the full
// histogram never travels -- and publishes one candidate per node to every peer.
// ---------------------------------------------------------------------------------------------------
constexpr int kMaxRanks = 16;
struct PeerBest { unsigned long long key; int split_bin; int dl; long long GL; long long HL; };
struct SymHeader {
  unsigned int flags[2][kMaxRanks];
  unsigned int pad_[32];
  PeerBest tab[kMaxRanks][kMaxLevelNodes];
};
struct PeerCtx {                           // by value into kernels; world <= 1 means single-GPU mode
  int world, rank;
  int f0, f1;                              // this rank's feature slice
  SymHeader* self;                         // own region (peers write into it)
  SymHeader* const* peers;                 // device array [world] of all regions (peer-mapped addresses)
  int64_t* const* peer_level_buf;          // device array [world] of all level buffers
};

__device__ __forceinline__ unsigned peer_epoch(const TreeState* st, int level, int phase) {
  return st->round_seq * 64u + (unsigned)level * 2u + (unsigned)phase + 1u;
}
// one thread: block until every peer's flag of `phase` reached `epoch` (bounded: a dead peer must not hang the GPU)
__device__ __forceinline__ void peer_wait(const PeerCtx& pc, TreeState* st, int phase, unsigned epoch) {
  for (int r = 0; r < pc.world; ++r) {
    volatile unsigned* fl = &pc.self->flags[phase][r];
    unsigned long long spins = 0;
    while ((int)(*fl - epoch) < 0) {
      if (++spins > (1ull << 25)) { st->peer_timeout = 1; break; }
      __nanosleep(64);
    }
  }
  __threadfence_system();
}
// tell every peer that `phase` of this level is done here (all earlier writes become visible first)
__global__ void peer_signal_kernel(PeerCtx pc, const TreeState* st, int level, int phase) {
  pdl_enter();
  __threadfence_system();
  const int r = threadIdx.x;
  if (r < pc.world) {
    volatile unsigned* fl = &pc.peers[r]->flags[phase][pc.rank];
    *fl = peer_epoch(st, level, phase);
  }
}

struct BestRec { unsigned long long key; int split_bin, dl; long long GL, HL; };

// `base` points at forest row 0; the row of the tree being grown comes from the device-side counter
__device__ __forceinline__ TreeOut tree_slice(const TreeOut& base, const TreeState* st, int max_nodes) {
  const size_t o = (size_t)st->tree_index * max_nodes;
  TreeOut t;
  t.left = base.left + o; t.right = base.right + o; t.parent = base.parent + o; t.split_index = base.split_index + o;
  t.split_bin = base.split_bin + o; t.split_cond = base.split_cond + o; t.default_left = base.default_left + o;
  t.base_weight = base.base_weight + o; t.loss_chg = base.loss_chg + o; t.sum_hess = base.sum_hess + o;
  return t;
}

__global__ void set_tree_index_kernel(TreeState* st, int idx) { st->tree_index = idx; }
__global__ void init_state_kernel(TreeState* st) {
  st->tree_index = 0; st->round_seq = 0; st->peer_timeout = 0; st->root_tot[0] = 0; st->root_tot[1] = 0;
}

__global__ void round_init_kernel(TreeState* st, NodeArrays na, TreeOut to_base, int max_nodes, int n_rows,
                                  HistPlan* plan, int grid) {
  pdl_enter();
  const TreeOut to = tree_slice(to_base, st, max_nodes);
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < max_nodes) {
    to.left[i] = -1; to.right[i] = -1; to.parent[i] = 2147483647; to.split_index[i] = 0; to.split_bin[i] = 0;
    to.split_cond[i] = 0.0f; to.default_left[i] = 0; to.base_weight[i] = 0.0f; to.loss_chg[i] = 0.0f;
    to.sum_hess[i] = 0.0f;
  }
  if (i == 0) {
    st->n_nodes = 1;
    for (int l = 0; l < kMaxDepth + 2; ++l) { st->lvl_first[l] = 0; st->lvl_count[l] = 0; }
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

// K9 + K5 fused, one CTA of 128 threads per (level node, feature), 2 bins per thread:
// reads the node's histogram from the level buffer (built child) or as parent - built (sibling),
// stores it in the pool for the next level, and scans it (exact int64 prefix sums, f64 gains).
__device__ __forceinline__ unsigned long long cand_key(float chg, int tiebreak) {
  return chg > 0.0f ? ((unsigned long long)__float_as_uint(chg) << 32) | (unsigned)tiebreak : 0ull;
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
                                                               int* __restrict__ group_cnt, int n_groups, PeerCtx pc,
                                                               const int32_t* __restrict__ feat_code) {
  CB_LEVEL_TRACE_SCOPE(0);
  pdl_enter();
  __shared__ long long sg[4], sh[4];
  __shared__ unsigned long long skf[4], skb[4];
  const int i = blockIdx.y;
  if (pc.world > 1) {                       // every rank's level buffer must be complete before it is read
    if (threadIdx.x == 0) peer_wait(pc, stw, 0, peer_epoch(stw, level, 0));
    __syncthreads();
  }
  if (blockIdx.x == 0 && i == 0)                        // housekeeping for the partition pass that follows
    for (int k = threadIdx.x; k < n_groups; k += blockDim.x) group_cnt[k] = 0;
  if (i >= st->lvl_count[level]) return;
  const int feat = (pc.world > 1 ? pc.f0 : 0) + (int)blockIdx.x;
  const bool no_feature = pc.world > 1 && feat >= pc.f1;     // empty slice (world > n_features): housekeeping only
  const int nid = st->lvl_first[level] + i;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  long long Gt, Ht;
  if (level == 0) {
    if (pc.world > 1) {                      // root totals: sum of the ranks' local totals (level buffer tail)
      Gt = 0; Ht = 0;
      for (int r = 0; r < pc.world; ++r) { Gt += __ldcv(pc.peer_level_buf[r] + 2 * (size_t)S); Ht += __ldcv(pc.peer_level_buf[r] + 2 * (size_t)S + 1); }
    } else {
      Gt = tail[0]; Ht = tail[1];
    }
    if (blockIdx.x == 0 && tid == 0) { na.G[0] = Gt; na.H[0] = Ht; }
  } else {
    Gt = na.G[nid]; Ht = na.H[nid];
  }
  if (no_feature) return;
  do {
  if (feature_mask && !feature_mask[(size_t)st->tree_index * f + feat]) {
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
      if (pc.world > 1) {                    // the fused collective: add this slice of every rank's histogram
        // all peers' loads are issued before the first add: one NVLink round trip per bin, not one per rank
        long long ag = 0, ah = 0;
        const size_t off = (size_t)j * S + b0 + b;
        for (int r0 = 0; r0 < pc.world; r0 += 8) {
          longlong2 q[8];
#pragma unroll
          for (int u = 0; u < 8; ++u)
            q[u] = (r0 + u < pc.world) ? __ldcv(reinterpret_cast<const longlong2*>(pc.peer_level_buf[r0 + u]) + off)
                                       : make_longlong2(0, 0);
#pragma unroll
          for (int u = 0; u < 8; ++u) { ag += q[u].x; ah += q[u].y; }
        }
        v[k] = make_longlong2(ag, ah);
      } else {
        v[k] = lb[b];
      }
    }
  }
  // member of a one-hot group column: only "member set" (bin 1) was accumulated; bin 0 = every other row of the
  // BUILT node = its totals - bin 1, exactly (int64).  Groups never hold missing values.
  if (feat_code && __ldg(feat_code + feat) != 0 && tid == 0) {
    long long Gb = Gt, Hb = Ht;              // level 0: the built node is the root
    if (level > 0) { const int bn = plan->b_node[j]; Gb = na.G[bn]; Hb = na.H[bn]; }
    v[0].x = Gb - v[1].x; v[0].y = Hb - v[1].y;
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int b = 2 * tid + k;
    if (b < nb) {
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
  const double dg = pow2neg(p.sg), dh = pow2neg(p.sh);
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
}

// peer mode, after the scan: every rank hands its best candidate per level node (over ITS feature slice)
// to all ranks, then raises flag 1.  One CTA.
__global__ void __launch_bounds__(256) peer_publish_kernel(PeerCtx pc, const TreeState* st, int level,
                                                           const FeatBest* __restrict__ feat_best, int f) {
  pdl_enter();
  const int cnt = st->lvl_count[level];
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    PeerBest pb; pb.key = st->best_key[i]; pb.split_bin = 0; pb.dl = 0; pb.GL = 0; pb.HL = 0;
    if (pb.key) {
      const FeatBest* r = feat_best + (size_t)i * f + (0xFFFF - (int)(pb.key & 0xFFFFu));
      pb.split_bin = r->split_bin; pb.dl = r->default_left; pb.GL = r->GL; pb.HL = r->HL;
    }
    for (int r = 0; r < pc.world; ++r) pc.peers[r]->tab[pc.rank][i] = pb;
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < pc.world) {
    volatile unsigned* fl = &pc.peers[threadIdx.x]->flags[1][pc.rank];
    *fl = peer_epoch(st, level, 1);
  }
}

struct SegCache {
  int begin[kMaxLevelNodes], count[kMaxLevelNodes];
};

// ===================================================================================================
// After the scan, a level needs two serial steps: "finish" (validity, child numbering in node-id
// order (R9), tree arrays, children totals) and "plan" (children segments from the left counts, which
// child to build, CTA assignment).  Neither is on the critical path: every CTA of the partition
// kernels derives the little it needs by itself (a <= 256-entry node table built from the atomic
// best-feature keys), and ONE extra CTA of the flags kernel does "finish" while the tiles compute
// flags, ONE extra CTA of the scatter kernel does "plan" while the tiles move rows.
// ===================================================================================================
struct NodeTable {                        // shared memory, one entry per node of the level
  int begin[kMaxLevelNodes], count[kMaxLevelNodes];
  int meta[kMaxLevelNodes];               // column position | default_left << 16 | kind << 24
  int bin[kMaxLevelNodes];
  int mcode[kMaxLevelNodes];              // 0: the column's code is the feature's bin; m + 1: bin = (code == m + 1) (group member)
  float leafL[kMaxLevelNodes], leafR[kMaxLevelNodes];
};

struct LevelCtx {
  const TreeState* st;
  int level, max_depth, f;
  const FeatBest* feat_best;
  cb_tree_params p;
  NodeArrays na;
  const int32_t* feat_pos;
  const int32_t* feat_code;               // nullptr: no grouped columns
  PeerCtx pc;
};

// best split candidate of level node i: the local arg-max over features, or (peer mode) the arg-max over
// the candidates every rank published for its feature slice (the key orders by gain, then lower feature)
__device__ __forceinline__ BestRec get_best(const LevelCtx& c, int i) {
  BestRec b; b.key = 0ull; b.split_bin = 0; b.dl = 0; b.GL = 0; b.HL = 0;
  if (c.pc.world <= 1) {
    b.key = __ldcg(&c.st->best_key[i]);
    if (b.key) {
      const FeatBest* r = c.feat_best + (size_t)i * c.f + (0xFFFF - (int)(b.key & 0xFFFFu));
      b.split_bin = __ldcg(&r->split_bin); b.dl = __ldcg(&r->default_left); b.GL = __ldcg(&r->GL); b.HL = __ldcg(&r->HL);
    }
    return b;
  }
  int best_r = -1;
  for (int r = 0; r < c.pc.world; ++r) {
    const unsigned long long k = __ldcg(&c.pc.self->tab[r][i].key);
    if (k > b.key) { b.key = k; best_r = r; }
  }
  if (best_r >= 0) {
    const PeerBest* t = &c.pc.self->tab[best_r][i];
    b.split_bin = __ldcg(&t->split_bin); b.dl = __ldcg(&t->dl); b.GL = __ldcg(&t->GL); b.HL = __ldcg(&t->HL);
  }
  return b;
}

// kind / split / leaf values of every node of the level, recomputed by whoever needs them
__device__ void build_node_table(NodeTable& nt, const LevelCtx& c, bool want_leaves) {
  const int first = c.st->lvl_first[c.level], cnt = c.st->lvl_count[c.level];
  const double dg = pow2neg(c.p.sg), dh = pow2neg(c.p.sh);
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int nid = first + i;
    const BestRec br = get_best(c, i);
    const unsigned long long key = br.key;
    const float bc = key ? __uint_as_float((unsigned)(key >> 32)) : 0.0f;
    const int bf = key ? 0xFFFF - (int)(key & 0xFFFFu) : -1;
    nt.begin[i] = c.na.begin[nid]; nt.count[i] = c.na.count[nid];
    bool valid = false;
    long long GL = 0, HL = 0;
    int sbin = 0, dl = 0;
    const long long G = c.na.G[nid], H = c.na.H[nid];
    if (bf >= 0) {
      GL = br.GL; HL = br.HL; sbin = br.split_bin; dl = br.dl;
      valid = bc > 1e-6f && !(bc < c.p.gamma) && c.level < c.max_depth && HL != 0 && (H - HL) != 0;
    }
    if (valid) {
      const int kind = c.level + 1 == c.max_depth ? 2 : 1;
      nt.meta[i] = c.feat_pos[bf] | (dl << 16) | (kind << 24);
      nt.bin[i] = sbin;
      nt.mcode[i] = c.feat_code ? c.feat_code[bf] : 0;
      nt.leafL[i] = 0.0f; nt.leafR[i] = 0.0f;
      if (kind == 2 && want_leaves) {
        const float bwL = __double2float_rn(calc_weight(__dmul_rn((double)GL, dg), __dmul_rn((double)HL, dh), c.p));
        const float bwR = __double2float_rn(calc_weight(__dmul_rn((double)(G - GL), dg), __dmul_rn((double)(H - HL), dh), c.p));
        nt.leafL[i] = __fmul_rn(bwL, c.p.eta); nt.leafR[i] = __fmul_rn(bwR, c.p.eta);
      }
    } else {
      nt.meta[i] = 0; nt.bin[i] = 0; nt.mcode[i] = 0;
      float leaf = 0.0f;
      if (want_leaves) {
        // same value the level above stored as this node's base weight (recomputed: the finish CTA of
        // this kernel overwrites that array entry with the leaf value while we run)
        const float bw = __double2float_rn(calc_weight(__dmul_rn((double)G, dg), __dmul_rn((double)H, dh), c.p));
        leaf = __fmul_rn(bw, c.p.eta);
      }
      nt.leafL[i] = leaf; nt.leafR[i] = leaf;
    }
  }
  __syncthreads();
}

// "finish": ONE CTA.  Numbers the children (R9), writes the tree and the segment table for the scatter pass.
__device__ void finish_level(TreeState* st, const LevelCtx& c, const float* __restrict__ cuts,
                             const float* __restrict__ min_vals, TreeOut to, LevelSegs* segs, int* s_child) {
  const int first = st->lvl_first[c.level], cnt = st->lvl_count[c.level];
  const double dg = pow2neg(c.p.sg), dh = pow2neg(c.p.sh);
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {      // validity again (cheap), 1 = splits
    const BestRec br = get_best(c, i);
    int v = 0;
    if (br.key) {
      const float bc = __uint_as_float((unsigned)(br.key >> 32));
      const long long HL = br.HL, HR = c.na.H[first + i] - HL;
      v = (bc > 1e-6f && !(bc < c.p.gamma) && c.level < c.max_depth && HL != 0 && HR != 0) ? 1 : 0;
    }
    s_child[i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (c.level == 0) {
      to.base_weight[0] = __double2float_rn(calc_weight(__dmul_rn((double)c.na.G[0], dg), __dmul_rn((double)c.na.H[0], dh), c.p));
      to.sum_hess[0] = __double2float_rn(__dmul_rn((double)c.na.H[0], dh));
    }
    int next = st->n_nodes;
    const int old = next;
    for (int i = 0; i < cnt; ++i) {
      if (s_child[i]) { s_child[i] = next; next += 2; } else s_child[i] = -1;
    }
    st->n_nodes = next;
    st->lvl_first[c.level + 1] = old;
    st->lvl_count[c.level + 1] = next - old;
    segs->n = cnt;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int nid = first + i;
    segs->begin[i] = c.na.begin[nid];
    segs->count[i] = c.na.count[nid];
    const int L = s_child[i];
    if (L < 0) {
      const float leaf = __fmul_rn(to.base_weight[nid], c.p.eta);
      to.split_cond[nid] = leaf;
      to.base_weight[nid] = leaf;
      segs->kind[i] = 0; segs->leafL[i] = leaf; segs->leafR[i] = leaf;
      segs->feat[i] = 0; segs->bin[i] = 0; segs->dl[i] = 0; segs->child[i] = -1;
      continue;
    }
    const BestRec br = get_best(c, i);
    const unsigned long long key = br.key;
    const int R = L + 1, bf = 0xFFFF - (int)(key & 0xFFFFu);
    const int sbin = br.split_bin, dl = br.dl;
    const long long GL = br.GL, HL = br.HL, GR = c.na.G[nid] - GL, HR = c.na.H[nid] - HL;
    to.left[nid] = L; to.right[nid] = R; to.parent[L] = nid; to.parent[R] = nid;
    to.split_index[nid] = bf; to.split_bin[nid] = sbin;
    to.split_cond[nid] = sbin >= 0 ? cuts[(size_t)bf * CB_CUT_STRIDE + sbin] : min_vals[bf];
    to.default_left[nid] = (uint8_t)dl;
    to.loss_chg[nid] = __uint_as_float((unsigned)(key >> 32));
    c.na.G[L] = GL; c.na.H[L] = HL; c.na.G[R] = GR; c.na.H[R] = HR;
    const float bwL = __double2float_rn(calc_weight(__dmul_rn((double)GL, dg), __dmul_rn((double)HL, dh), c.p));
    const float bwR = __double2float_rn(calc_weight(__dmul_rn((double)GR, dg), __dmul_rn((double)HR, dh), c.p));
    to.sum_hess[L] = __double2float_rn(__dmul_rn((double)HL, dh));
    to.sum_hess[R] = __double2float_rn(__dmul_rn((double)HR, dh));
    segs->feat[i] = c.feat_pos[bf]; segs->bin[i] = sbin; segs->dl[i] = dl; segs->child[i] = L;
    if (c.level + 1 == c.max_depth) {
      const float lL = __fmul_rn(bwL, c.p.eta), lR = __fmul_rn(bwR, c.p.eta);
      to.split_cond[L] = lL; to.base_weight[L] = lL;
      to.split_cond[R] = lR; to.base_weight[R] = lR;
      segs->kind[i] = 2; segs->leafL[i] = lL; segs->leafR[i] = lR;
    } else {
      to.base_weight[L] = bwL; to.base_weight[R] = bwR;
      segs->kind[i] = 1; segs->leafL[i] = 0.0f; segs->leafR[i] = 0.0f;
    }
  }
}

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
// the same for 32 consecutive positions starting at pos0 (warp uniform when the span lies in one segment)
__device__ __forceinline__ int find_seg_warp(const int* __restrict__ s_begin, const int* __restrict__ s_count, int n,
                                             int pos0, int lane) {
  const int a = find_seg(s_begin, s_count, n, pos0);
  if (a >= 0 && pos0 + 31 < s_begin[a] + s_count[a]) return a;
  return find_seg(s_begin, s_count, n, pos0 + lane);
}

// the code of column position `pos` of `row`: from the position-major copy when there is one
// (consecutive rows of a node are close together there: a level's gathers touch ~2^level bytes per
// row instead of a DRAM burst per row), else from the row/block-major matrix
__device__ __forceinline__ int code_at(const BinsView& v, const uint8_t* __restrict__ bins_cm, uint32_t row, int pos) {
  if (bins_cm) return __ldg(bins_cm + (size_t)pos * v.n_rows_cm + row);
  return __ldg(v.base + byte_index(v, row, pos));
}

constexpr int kTileGroup = 64;            // tiles per coarse left-count group

// P1: go-left ballots of every position inside a node that splits here.  Grid = tiles + 1 (finish CTA).
__global__ void __launch_bounds__(256) part_level_flags_kernel(BinsView view, const uint8_t* __restrict__ bins_cm,
                                                               const uint32_t* __restrict__ rows_in, int identity,
                                                               int n_rows, int n_tiles, LevelCtx c, TreeState* stw,
                                                               const float* __restrict__ cuts,
                                                               const float* __restrict__ min_vals, TreeOut to,
                                                               LevelSegs* segs, uint32_t* __restrict__ flag_bits,
                                                               int* __restrict__ tile_cnt, int* __restrict__ group_cnt) {
  CB_LEVEL_TRACE_SCOPE(1);
  pdl_enter();
  __shared__ NodeTable nt;
  __shared__ int wsum[8];
  if (c.pc.world > 1) {                             // all ranks' candidates of this level must have arrived
    if (threadIdx.x == 0) peer_wait(c.pc, stw, 1, peer_epoch(stw, c.level, 1));
    __syncthreads();
  }
  if (blockIdx.x == 0) {                            // the extra CTA (scheduled first): serial bookkeeping, overlapped with the tiles
    finish_level(stw, c, cuts, min_vals, tree_slice(to, stw, (1 << (c.max_depth + 1)) - 1), segs, nt.begin);
    return;
  }
  build_node_table(nt, c, false);
  const int n = c.st->lvl_count[c.level];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  constexpr int IT = kPartTile / 256;
  for (int tile = (int)blockIdx.x - 1; tile < n_tiles; tile += (int)gridDim.x - 1) {   // whole waves: no tail
  const int tile0 = tile * kPartTile;
  uint32_t row[IT];
  int seg[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) {                 // row indices first: 8 independent loads in flight
    const int pos0 = tile0 + wid * (kPartTile / 8) + it * 32;
    const int pos = pos0 + lane;
    row[it] = 0; seg[it] = -1;
    if (pos0 < n_rows) {
      const int sgm = find_seg_warp(nt.begin, nt.count, n, pos0, lane);
      if (pos < n_rows) {
        seg[it] = sgm;
        row[it] = identity ? (uint32_t)pos : __ldg(rows_in + pos);
      }
    }
  }
  int code[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) {                 // then the gathers
    code[it] = -1;
    if (seg[it] >= 0 && (nt.meta[seg[it]] >> 24) == 1) {
      code[it] = code_at(view, bins_cm, row[it], nt.meta[seg[it]] & 0xFFFF);
      const int mc = nt.mcode[seg[it]];
      if (mc) code[it] = code[it] == mc ? 1 : 0;
    }
  }
  int cnt = 0;
#pragma unroll
  for (int it = 0; it < IT; ++it) {
    const int pos = tile0 + wid * (kPartTile / 8) + it * 32 + lane;
    bool fl = false;
    if (code[it] >= 0) {
      const int m = nt.meta[seg[it]];
      fl = code[it] == CB_MISSING_BIN ? ((m >> 16) & 1) != 0 : code[it] <= nt.bin[seg[it]];
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
    tile_cnt[tile] = t;
    if (t) atomicAdd(group_cnt + tile / kTileGroup, t);
  }
  __syncthreads();
  }
}

// lefts in all tiles before tile t (coarse groups + the tiles of t's own group), computed by one warp
__device__ __forceinline__ int tile_prefix_warp(const int* __restrict__ group_cnt, const int* __restrict__ tile_cnt, int t,
                                                int lane) {
  const int g = t / kTileGroup;
  int s = 0;
  for (int k = lane; k < g; k += 32) s += group_cnt[k];
  for (int k = g * kTileGroup + lane; k < t; k += 32) s += tile_cnt[k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}
// lefts strictly before position p (global prefix over all positions), computed by one warp
__device__ __forceinline__ int lefts_before_warp(const int* __restrict__ group_cnt, const int* __restrict__ tile_cnt,
                                                 const uint32_t* __restrict__ flag_bits, int p, int n_rows, int lane) {
  if (p <= 0) return 0;
  int tile = p / kPartTile;
  int w1 = p >> 5, part_bits = p & 31;
  if (tile * kPartTile >= n_rows) { tile -= 1; w1 = (n_rows + 31) >> 5; part_bits = 0; }
  int s = 0;
  for (int w = tile * (kPartTile / 32) + lane; w < w1; w += 32) s += __popc(flag_bits[w]);
  if (lane == 0 && part_bits && w1 * 32 < n_rows) s += __popc(flag_bits[w1] & ((1u << part_bits) - 1u));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s + tile_prefix_warp(group_cnt, tile_cnt, tile, lane);
}

struct PlanSmem {
  int pick[kMaxLevelNodes], cbegin[kMaxLevelNodes], ccount[kMaxLevelNodes];
  int b_begin[kMaxBuild], b_count[kMaxBuild];
};

// "plan": ONE CTA.  Children segments from the left counts, which child to build, CTA assignment.
__device__ void plan_next_level(TreeState* st, int level, int max_depth, NodeArrays na, const LevelSegs* segs,
                                const int* __restrict__ group_cnt, const int* __restrict__ tile_cnt,
                                const uint32_t* __restrict__ flag_bits, int n_rows, HistPlan* plan, int grid,
                                int by_hessian, PlanSmem& sm) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int n = segs->n;
  for (int i = threadIdx.x; i < kMaxLevelNodes; i += blockDim.x) st->best_key[i] = 0ull;   // for the next level's scan
  for (int i = wid; i < n; i += nwarp) {
    if (segs->kind[i] != 1) { if (lane == 0) sm.pick[i] = -1; continue; }
    const int bg = segs->begin[i], ct = segs->count[i];
    const int Lb = lefts_before_warp(group_cnt, tile_cnt, flag_bits, bg, n_rows, lane);
    const int Le = lefts_before_warp(group_cnt, tile_cnt, flag_bits, bg + ct, n_rows, lane);
    if (lane == 0) {
      const int nl = Le - Lb;
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

// P3: stable scatter of the nodes that split (lefts first); rows of nodes that stopped here take their
// leaf value (K7).  Grid = tiles + 1 (plan CTA).
__global__ void __launch_bounds__(256) part_level_scatter_kernel(const uint32_t* __restrict__ rows_in, int identity,
                                                                 int n_rows, int n_tiles, const LevelSegs* __restrict__ segs,
                                                                 const uint32_t* __restrict__ flag_bits,
                                                                 const int* __restrict__ tile_cnt,
                                                                 const int* __restrict__ group_cnt,
                                                                 uint32_t* __restrict__ rows_out, float* __restrict__ margin,
                                                                 TreeState* st, int level, int max_depth, NodeArrays na,
                                                                 HistPlan* plan, int grid, int by_hessian) {
  CB_LEVEL_TRACE_SCOPE(2);
  pdl_enter();
  __shared__ SegCache sc;
  __shared__ int s_Lb[kMaxLevelNodes], s_nl[kMaxLevelNodes];
  __shared__ PlanSmem psm;
  __shared__ int s_prefix;
  if (blockIdx.x == 0) {                            // the extra CTA (scheduled first)
    plan_next_level(st, level, max_depth, na, segs, group_cnt, tile_cnt, flag_bits, n_rows, plan, grid, by_hessian, psm);
    return;
  }
  const int n = segs->n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { sc.begin[i] = segs->begin[i]; sc.count[i] = segs->count[i]; }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int tile = (int)blockIdx.x - 1; tile < n_tiles; tile += (int)gridDim.x - 1) {   // whole waves: no tail
  const int tile0 = tile * kPartTile;
  const int tile_end = min(n_rows, tile0 + kPartTile);
  // segments that overlap this tile: [s_lo, s_hi]
  int s_lo = 0;
  {
    int lo = 0, hi = n;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (sc.begin[mid] + sc.count[mid] <= tile0) lo = mid + 1; else hi = mid; }
    s_lo = lo;
  }
  if (wid == 0) {
    const int v = tile_prefix_warp(group_cnt, tile_cnt, tile, lane);
    if (lane == 0) s_prefix = v;
  }
  for (int i = s_lo + wid; i < n && sc.begin[i] < tile_end; i += 8) {      // one warp per overlapping node
    if (segs->kind[i] != 1) continue;
    const int Lb = lefts_before_warp(group_cnt, tile_cnt, flag_bits, sc.begin[i], n_rows, lane);
    const int Le = lefts_before_warp(group_cnt, tile_cnt, flag_bits, sc.begin[i] + sc.count[i], n_rows, lane);
    if (lane == 0) { s_Lb[i] = Lb; s_nl[i] = Le - Lb; }
  }
  __syncthreads();
  constexpr int wpw = kPartTile / 8 / 32;
  const int w0 = tile0 >> 5;
  int before = 0;
  for (int k = lane; k < wid * wpw; k += 32)
    if ((w0 + k) * 32 < n_rows) before += __popc(flag_bits[w0 + k]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  int run = s_prefix + before;
  uint32_t bal[wpw], row[wpw];
  int seg[wpw];
#pragma unroll
  for (int it = 0; it < wpw; ++it) {                 // all loads of the warp's 256 positions first
    const int pos0 = tile0 + wid * (kPartTile / 8) + it * 32;
    const int pos = pos0 + lane;
    bal[it] = 0; row[it] = 0; seg[it] = -2;
    if (pos0 < n_rows) {
      bal[it] = flag_bits[pos0 >> 5];
      seg[it] = find_seg_warp(sc.begin, sc.count, n, pos0, lane);
      if (pos < n_rows) row[it] = identity ? (uint32_t)pos : __ldg(rows_in + pos);
      else seg[it] = -2;
    }
  }
#pragma unroll
  for (int it = 0; it < wpw; ++it) {
    const int pos = tile0 + wid * (kPartTile / 8) + it * 32 + lane;
    if (seg[it] != -2) {
      const int i = seg[it];
      int dest = pos;
      if (i >= 0) {
        const int kind = segs->kind[i];
        if (kind == 1) {
          const int lb = run + __popc(bal[it] & ((1u << lane) - 1u)) - s_Lb[i];   // lefts before, inside the node
          const int bg = sc.begin[i];
          dest = ((bal[it] >> lane) & 1u) ? bg + lb : bg + s_nl[i] + (pos - bg - lb);
        } else if (kind == 0) {
          margin[row[it]] = __fadd_rn(margin[row[it]], segs->leafL[i]);
        }
      }
      rows_out[dest] = row[it];
    }
    run += __popc(bal[it]);
  }
  __syncthreads();
  }
}

// last level: children are leaves, so rows are not moved; each row takes its child's leaf value.
// Grid = resident tile CTAs + 1 (finish CTA); the tile CTAs loop over blocks of 1024 positions.
__global__ void __launch_bounds__(256) final_apply_kernel(BinsView view, const uint8_t* __restrict__ bins_cm,
                                                          const uint32_t* __restrict__ rows_in, int identity,
                                                          int n_rows, int n_blocks, LevelCtx c, TreeState* stw,
                                                          const float* __restrict__ cuts,
                                                          const float* __restrict__ min_vals, TreeOut to, LevelSegs* segs,
                                                          float* __restrict__ margin) {
  CB_LEVEL_TRACE_SCOPE(3);
  pdl_enter();
  __shared__ NodeTable nt;
  if (c.pc.world > 1) {
    if (threadIdx.x == 0) peer_wait(c.pc, stw, 1, peer_epoch(stw, c.level, 1));
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    finish_level(stw, c, cuts, min_vals, tree_slice(to, stw, (1 << (c.max_depth + 1)) - 1), segs, nt.begin);
    return;
  }
  build_node_table(nt, c, true);
  const int n = c.st->lvl_count[c.level];
  const int lane = threadIdx.x & 31;
  constexpr int IT = 4;                                // positions per thread, loads batched
  for (int blk = (int)blockIdx.x - 1; blk < n_blocks; blk += (int)gridDim.x - 1) {   // resident CTAs: one table build each
  const int base = blk * (blockDim.x * IT) + (threadIdx.x & ~31) * IT;
  uint32_t row[IT];
  int seg[IT], code[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) {
    const int pos0 = base + it * 32, pos = pos0 + lane;
    seg[it] = -1; row[it] = 0;
    if (pos0 < n_rows) {
      const int i = find_seg_warp(nt.begin, nt.count, n, pos0, lane);
      if (pos < n_rows && i >= 0) {
        seg[it] = i;
        row[it] = identity ? (uint32_t)pos : __ldg(rows_in + pos);
      }
    }
  }
#pragma unroll
  for (int it = 0; it < IT; ++it) {
    code[it] = 0;
    if (seg[it] >= 0 && (nt.meta[seg[it]] >> 24) == 2) {
      code[it] = code_at(view, bins_cm, row[it], nt.meta[seg[it]] & 0xFFFF);
      const int mc = nt.mcode[seg[it]];
      if (mc) code[it] = code[it] == mc ? 1 : 0;
    }
  }
  float mg[IT];
#pragma unroll
  for (int it = 0; it < IT; ++it) mg[it] = seg[it] >= 0 ? margin[row[it]] : 0.0f;
#pragma unroll
  for (int it = 0; it < IT; ++it) {
    if (seg[it] < 0) continue;
    const int m = nt.meta[seg[it]];
    float leaf = nt.leafL[seg[it]];
    if ((m >> 24) == 2) {
      const bool left = code[it] == CB_MISSING_BIN ? ((m >> 16) & 1) != 0 : code[it] <= nt.bin[seg[it]];
      if (!left) leaf = nt.leafR[seg[it]];
    }
    margin[row[it]] = __fadd_rn(mg[it], leaf);
  }
  }
}

__global__ void round_end_kernel(TreeState* st, int32_t* n_nodes_all) {
  pdl_enter();
  if (threadIdx.x == 1) { st->root_tot[0] = 0; st->root_tot[1] = 0; }   // the next round's gradient kernel adds into it
  for (int i = threadIdx.x; i < kMaxLevelNodes; i += blockDim.x) st->best_key[i] = 0ull;
  if (threadIdx.x == 0) {
    n_nodes_all[st->tree_index] = st->n_nodes;
    st->tree_index += 1;
    st->round_seq += 1;
  }
}

}  // namespace cb
