// Device kernels of the histogram-GBDT hot path (sm_100a).  Included by cobalt_gbdt.cu only.
#pragma once
#include "common.cuh"

namespace cb {

// =====================================================================================
// K2  bin-index packing: one thread produces one u32 word = 4 features of one row
// =====================================================================================
__device__ __forceinline__ uint32_t bin_of(float x, const float* __restrict__ c, int nc) {
  if (x != x) return CB_MISSING_BIN;
  int lo = 0, hi = nc;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (__ldg(c + mid) <= x) lo = mid + 1; else hi = mid;
  }
  return (uint32_t)(lo == nc ? nc - 1 : lo);
}

// Where the byte of column position p of a row lives.  Row major: base[row * ldb + p].
// Block major (conflict-free layout): column block k = p / 128 is its own [n_rows][width_k] array,
// width 128 except for the last block, so every block is read as whole aligned lines.
struct BinsView {
  const uint8_t* base;
  int64_t n_rows, ldb;
  int64_t n_rows_cm;       // row count of the position-major copy (always the true row count)
  int blocked, n_blocks, last_w;
};
__host__ __device__ inline BinsView make_view(const uint8_t* base, int64_t n_rows, int64_t ldb, int blocked) {
  BinsView v;
  v.base = base; v.n_rows = n_rows; v.ldb = ldb; v.blocked = blocked; v.n_rows_cm = n_rows;
  v.n_blocks = (int)((ldb + 127) / 128);
  v.last_w = (int)(ldb - 128 * (int64_t)(v.n_blocks - 1));
  if (v.n_blocks <= 1) v.blocked = 0;
  return v;
}
__device__ __forceinline__ size_t byte_index(const BinsView& v, int64_t row, int p) {
  if (!v.blocked) return (size_t)row * v.ldb + p;
  const int k = p >> 7;
  const int w = (k == v.n_blocks - 1) ? v.last_w : 128;
  return (size_t)v.n_rows * 128 * k + (size_t)row * w + (p & 127);
}

// pos_feature[p]: >= 0 plain feature, -1 padding (code 255), <= -2 group -2 - value of mutually exclusive 2-bin
// features packed as one code: 0 = no member in bin 1, m + 1 = member m in bin 1 (hist_cf.cuh "storage columns")
__global__ void __launch_bounds__(256) bin_pack_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx,
                                                       const float* __restrict__ cuts,
                                                       const int32_t* __restrict__ n_cuts,
                                                       const int32_t* __restrict__ pos_feature,
                                                       const int32_t* __restrict__ grp_ptr,
                                                       const int32_t* __restrict__ grp_feat,
                                                       uint8_t* __restrict__ bins, int wpr, BinsView view) {
  int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= n * wpr) return;
  int64_t row = item / wpr;
  int w = (int)(item - row * wpr);
  const float* x = X + row * ldx;
  uint32_t out = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int feat = pos_feature ? __ldg(pos_feature + 4 * w + j) : (4 * w + j < f ? 4 * w + j : -1);
    uint32_t code = CB_MISSING_BIN;
    if (feat >= 0) {
      code = bin_of(__ldg(x + feat), cuts + (size_t)feat * CB_CUT_STRIDE, __ldg(n_cuts + feat));
    } else if (feat <= -2 && grp_ptr) {
      const int g = -2 - feat;
      const int k0 = __ldg(grp_ptr + g), k1 = __ldg(grp_ptr + g + 1);
      code = 0;
      for (int k = k0; k < k1; ++k) {                 // bin 1 of a 2-bin feature <=> x >= its first cut (never NaN in a group)
        const int mf = __ldg(grp_feat + k);
        if (__ldg(x + mf) >= __ldg(cuts + (size_t)mf * CB_CUT_STRIDE)) code = (uint32_t)(k - k0 + 1);
      }
    }
    out |= code << (8 * j);
  }
  *reinterpret_cast<uint32_t*>(bins + byte_index(view, row, 4 * w)) = out;
}

// One-hot detection (one-off, before the layout is planned): for up to 64 candidate features (2-bin features, chosen by
// the host) counts over the rows how often two candidates are in bin 1 together (pair[i * nc + j], i < j) and how often
// a candidate is missing (miss[i]).  Candidates that never co-occur and are never missing can share one code byte.
__global__ void __launch_bounds__(256) onehot_pair_kernel(const float* __restrict__ X, int64_t n, int64_t ldx,
                                                          const float* __restrict__ cuts, const int32_t* __restrict__ cand,
                                                          int nc, unsigned long long* __restrict__ pair,
                                                          unsigned long long* __restrict__ miss) {
  __shared__ unsigned int s_pair[64 * 64];
  __shared__ unsigned int s_miss[64];
  __shared__ float s_cut[64];
  __shared__ int s_feat[64];
  for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) s_pair[i] = 0;
  if (threadIdx.x < 64) {
    s_miss[threadIdx.x] = 0;
    const int ft = threadIdx.x < nc ? __ldg(cand + threadIdx.x) : 0;
    s_feat[threadIdx.x] = ft;
    s_cut[threadIdx.x] = __ldg(cuts + (size_t)ft * CB_CUT_STRIDE);
  }
  __syncthreads();
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
    const float* x = X + row * ldx;
    unsigned long long on = 0ull;
    for (int i = 0; i < nc; ++i) {
      const float v = __ldg(x + s_feat[i]);
      if (v != v) atomicAdd(&s_miss[i], 1u);
      else if (v >= s_cut[i]) on |= 1ull << i;
    }
    if (__popcll(on) > 1) {
      unsigned long long a = on;
      while (a) {
        const int i = __ffsll((long long)a) - 1;
        a &= a - 1;
        unsigned long long b = a;
        while (b) {
          const int j = __ffsll((long long)b) - 1;
          b &= b - 1;
          atomicAdd(&s_pair[i * 64 + j], 1u);
        }
      }
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 64 * 64; e += blockDim.x) {
    const int i = e >> 6, j = e & 63;
    if (i < nc && j < nc && s_pair[e]) atomicAdd(pair + (size_t)i * nc + j, (unsigned long long)s_pair[e]);
  }
  if (threadIdx.x < nc && s_miss[threadIdx.x]) atomicAdd(miss + threadIdx.x, (unsigned long long)s_miss[threadIdx.x]);
}

// position-major copy of the binned matrix: out[p][row] = code of column position p (32 x 32 byte tiles)
__global__ void __launch_bounds__(256) bin_transpose_kernel(BinsView view, int64_t n, uint8_t* __restrict__ out) {
  __shared__ uint8_t t[32][33];
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int p0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k = ty; k < 32; k += 8) {
    const int64_t r = r0 + k;
    const int p = p0 + tx;
    t[k][tx] = (r < n && p < view.ldb) ? view.base[byte_index(view, r, p)] : (uint8_t)CB_MISSING_BIN;
  }
  __syncthreads();
  for (int k = ty; k < 32; k += 8) {
    const int p = p0 + k;
    const int64_t r = r0 + tx;
    if (p < view.ldb && r < n) out[(size_t)p * n + r] = t[tx][k];
  }
}

// =====================================================================================
// K3  logistic gradient / hessian, fixed point (R4)
// =====================================================================================
__global__ void __launch_bounds__(256) grad_kernel(const float* __restrict__ margin, const uint8_t* __restrict__ y,
                                                   const float* __restrict__ weight, int64_t n, float spw,
                                                   float fg, float fh, uint32_t keep_thresh, uint32_t seed,
                                                   uint32_t tree, int64_t row_offset, int2* __restrict__ gpair,
                                                   unsigned long long* __restrict__ totals,
                                                   const int* __restrict__ tree_ptr) {
  pdl_enter();
  if (tree_ptr) tree = (uint32_t)*tree_ptr;          // trainer: the device-side round counter
  int64_t G = 0, H = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float p = spec_sigmoid(__ldg(margin + i));
    uint8_t yi = __ldg(y + i);
    float wt = yi ? spw : 1.0f;
    if (weight) wt = __fmul_rn(wt, __ldg(weight + i));
    float g = __fmul_rn(__fsub_rn(p, (float)yi), wt);
    float hh = __fmul_rn(p, __fsub_rn(1.0f, p));
    hh = hh < 1e-16f ? 1e-16f : hh;
    float h = __fmul_rn(hh, wt);
    int qg = (int)rintf(__fmul_rn(g, fg)), qh = (int)rintf(__fmul_rn(h, fh));
    if (keep_thresh && spec_hash(seed, tree, (uint64_t)(row_offset + i)) >= keep_thresh) { qg = 0; qh = 0; }
    gpair[i] = make_int2(qg, qh);
    G += qg; H += qh;
  }
  G = warp_sum_i64(G); H = warp_sum_i64(H);
  __shared__ int64_t sg[8], sh[8];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sg[wid] = G; sh[wid] = H; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int64_t a = 0, b = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { a += sg[k]; b += sh[k]; }
    atomicAdd(totals, (unsigned long long)a);
    atomicAdd(totals + 1, (unsigned long long)b);
  }
}

__global__ void sigmoid_kernel(const float* __restrict__ m, int64_t n, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = spec_sigmoid(m[i]);
}

// =====================================================================================
// K4  histogram build
//   plan   : which CTA scans which slice of which node (device side, no host sync)
//   accum  : shared-memory privatised int32 histograms, one partial per <= kChunkRows rows
//   reduce : partials -> int64 level buffer
// =====================================================================================
struct TileInfo { int w0, w1, s0, s1; };  // u32 words [w0,w1) of a row, histogram slots [s0,s1)

constexpr int kMaxTiles = 16;
constexpr int kMaxBuild = 256;   // build nodes per level (depth <= 9)
constexpr int kMaxGrid = 1024;   // CTAs along x in hist_accum

struct HistPlan {               // lives in device memory
  int n_tiles;
  TileInfo tiles[kMaxTiles];
  int n_build;
  int total_rows;               // rows scanned by this level's build (sum over build nodes)
  int b_node[kMaxBuild];        // node id whose histogram is built
  int b_sibling[kMaxBuild];     // node id derived by subtraction (-1: none)
  int b_parent[kMaxBuild];
  int b_pbeg[kMaxBuild + 1];    // partial range of build slot j: [b_pbeg[j], b_pbeg[j+1])
  int lvl_src[2 * kMaxBuild];   // level node i -> build slot * 2 + (0 built | 1 sibling = parent - built)
  int cta_slot[kMaxGrid];       // build slot of CTA b (-1: idle)
  int cta_begin[kMaxGrid];      // first position (in row order) of the CTA's slice
  int cta_rows[kMaxGrid];
  int cta_pbase[kMaxGrid];      // first partial written by the CTA
  int b_cta0[kMaxBuild];        // first CTA of build slot j; its CTAs are consecutive
  int b_ncta[kMaxBuild];
  int min_rows;                 // rows a CTA should at least get (0: 256); set once at creation
  unsigned int done[kMaxBuild]; // CTAs of build slot j whose partials are complete (conflict-free kernel's in-kernel reduce)
};

// greedy feature tiling on u32-word boundaries; run once per bin layout
__global__ void plan_tiles_kernel(const int32_t* __restrict__ bin_ptr, int f, int budget_slots, HistPlan* plan) {
  if (threadIdx.x || blockIdx.x) return;
  int words = (f + 3) / 4, nt = 0, w0 = 0, s0 = 0;
  for (int w = 0; w < words; ++w) {
    int e = bin_ptr[min(4 * w + 4, f)];
    if (e - s0 > budget_slots && w > w0) {
      if (nt < kMaxTiles) plan->tiles[nt] = TileInfo{w0, w, s0, bin_ptr[4 * w]};
      ++nt; w0 = w; s0 = bin_ptr[4 * w];
    }
  }
  if (nt < kMaxTiles) plan->tiles[nt] = TileInfo{w0, words, s0, bin_ptr[f]};
  ++nt;
  plan->n_tiles = min(nt, kMaxTiles);
  plan->min_rows = 0;
}

__global__ void plan_set_min_rows_kernel(HistPlan* plan, int min_rows) { plan->min_rows = min_rows; }

// distributes `grid` CTAs over the build nodes proportionally to their row counts
__device__ inline void plan_ctas(HistPlan* plan, int n_build, const int* begins, const int* counts, int grid) {
  long long total = 0;
  for (int j = 0; j < n_build; ++j) total += counts[j];
  int spare = grid - n_build, b = 0, p = 0;
  for (int j = 0; j < n_build; ++j) {
    int nj = 1 + (total > 0 ? (int)((long long)spare * counts[j] / total) : 0);
    const int mr = plan->min_rows > 0 ? plan->min_rows : 256;
    int need = (counts[j] + mr - 1) / mr;                   // at least `mr` rows per CTA: every CTA pays a fixed zero / flush / reduce
    nj = max(1, min(nj, need));
    int per = (counts[j] + nj - 1) / nj;
    per = (per + 31) & ~31;
    plan->b_pbeg[j] = p;
    plan->b_cta0[j] = b; plan->b_ncta[j] = nj; plan->done[j] = 0u;
    for (int k = 0; k < nj; ++k) {
      int s = min(counts[j], k * per), e = min(counts[j], (k + 1) * per);
      plan->cta_slot[b] = j; plan->cta_begin[b] = begins[j] + s; plan->cta_rows[b] = e - s; plan->cta_pbase[b] = p;
      p += (e - s + kChunkRows - 1) / kChunkRows;
      ++b;
    }
  }
  plan->b_pbeg[n_build] = p;
  plan->n_build = n_build;
  plan->total_rows = (int)total;
  for (; b < grid; ++b) { plan->cta_slot[b] = -1; plan->cta_rows[b] = 0; }
}

__global__ void plan_single_kernel(HistPlan* plan, int count, int grid) {
  if (threadIdx.x || blockIdx.x) return;
  int b0 = 0, c0 = count;
  plan->b_node[0] = 0; plan->b_sibling[0] = -1; plan->b_parent[0] = -1;
  plan_ctas(plan, 1, &b0, &c0, grid);
}

// blockDim.x = 1024.  Thread t owns u32 word (t % Wt) of rows (t / Wt) + k * RPB: its four
// feature offsets are loop invariant, and within a warp every lane addresses a different
// feature, so two lanes never hit the same histogram slot.
template <int UNROLL>
__global__ void __launch_bounds__(1024, 1)
hist_accum_kernel(const uint32_t* __restrict__ bins32, int wpr, const int2* __restrict__ gpair,
                  const uint32_t* __restrict__ row_idx, int identity, const HistPlan* __restrict__ plan,
                  const int32_t* __restrict__ bin_ptr, int f, int S, int32_t* __restrict__ partials) {
  extern __shared__ int32_t smem[];
  const int tile_id = blockIdx.y;
  if (tile_id >= plan->n_tiles) return;
  const int slot = plan->cta_slot[blockIdx.x];
  if (slot < 0) return;
  const TileInfo tile = plan->tiles[tile_id];
  const int St = tile.s1 - tile.s0, Wt = tile.w1 - tile.w0;
  int32_t* sG = smem;
  int32_t* sH = smem + St;
  for (int s = threadIdx.x; s < 2 * St; s += blockDim.x) smem[s] = 0;
  const int rpb = blockDim.x / Wt;
  const int rl = threadIdx.x / Wt;
  const int w = tile.w0 + (threadIdx.x - rl * Wt);
  const bool active = rl < rpb;
  int off[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int feat = 4 * w + j;
    off[j] = (active && feat < f) ? __ldg(bin_ptr + feat) - tile.s0 : 0;
  }
  __syncthreads();
  const int pos0 = plan->cta_begin[blockIdx.x];
  const int rows = plan->cta_rows[blockIdx.x];
  int pid = plan->cta_pbase[blockIdx.x];
  for (int c0 = 0; c0 < rows; c0 += kChunkRows, ++pid) {
    const int crow = min(kChunkRows, rows - c0);
    if (active) {
      for (int base = rl; base < crow; base += rpb * UNROLL) {
        uint32_t r[UNROLL];
        uint32_t word[UNROLL];
        int2 gp[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          int i = base + u * rpb;
          bool ok = i < crow;
          int pos = pos0 + c0 + (ok ? i : 0);
          r[u] = identity ? (uint32_t)pos : __ldg(row_idx + pos);
          word[u] = 0xFFFFFFFFu;
          if (ok) word[u] = __ldg(bins32 + (size_t)r[u] * wpr + w);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) gp[u] = __ldg(gpair + r[u]);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t b = (word[u] >> (8 * j)) & 255u;
            if (b != CB_MISSING_BIN) {
              atomicAdd(sG + off[j] + b, gp[u].x);
              atomicAdd(sH + off[j] + b, gp[u].y);
            }
          }
        }
      }
    }
    __syncthreads();
    int32_t* pg = partials + (size_t)pid * 2 * S + tile.s0;
    int32_t* ph = pg + S;
    for (int s = threadIdx.x; s < St; s += blockDim.x) {
      pg[s] = sG[s]; ph[s] = sH[s];
      sG[s] = 0; sH[s] = 0;
    }
    __syncthreads();
  }
}

// level buffer [slot][S][2] int64  <-  sum of the slot's int32 partials ([2][S] each)
__global__ void __launch_bounds__(256) hist_reduce_kernel(const int32_t* __restrict__ partials,
                                                          const HistPlan* __restrict__ plan, int S,
                                                          int64_t* __restrict__ level_buf) {
  const int j = blockIdx.y;
  if (j >= plan->n_build) return;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  int64_t g = 0, h = 0;
  const int p0 = plan->b_pbeg[j], p1 = plan->b_pbeg[j + 1];
  const int32_t* pp = partials + (size_t)p0 * 2 * S + s;
  for (int p = p0; p < p1; ++p, pp += 2 * (size_t)S) {
    g += __ldg(pp);
    h += __ldg(pp + S);
  }
  reinterpret_cast<longlong2*>(level_buf)[(size_t)j * S + s] = make_longlong2(g, h);
}

__global__ void subtract_kernel(const int64_t* __restrict__ a, const int64_t* __restrict__ b, int64_t* __restrict__ c,
                                int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) c[i] = a[i] - b[i];
}

// =====================================================================================
// K5  split scan: one warp per (node, feature); 8 bins per lane; exact int64 prefix sums;
//     gains in f64 with one rounding per operation (bit-equal to the oracle)
// =====================================================================================
struct FeatBest {
  float chg;      // 0 when the feature has no candidate
  int split_bin;
  int default_left;
  int pad;
  int64_t GL, HL;
};

__device__ __forceinline__ float split_chg(int64_t gl, int64_t hl, int64_t gr, int64_t hr, double ig, double ih,
                                           float root_gain, const cb_tree_params& p) {
  double HL = __dmul_rn((double)hl, ih), HR = __dmul_rn((double)hr, ih);
  if (!(HL >= (double)p.min_child_weight && HR >= (double)p.min_child_weight)) return -1.0f;
  double gain = __dadd_rn(calc_gain(__dmul_rn((double)gl, ig), HL, p), calc_gain(__dmul_rn((double)gr, ig), HR, p));
  float chg = __double2float_rn(__dsub_rn(gain, (double)root_gain));
  return isinf(chg) ? -1.0f : chg;
}

// returns the feature's best candidate to every lane
__device__ __forceinline__ FeatBest scan_feature(const int64_t* __restrict__ hist, int nb, int64_t Gt, int64_t Ht,
                                                 const cb_tree_params& p) {
  const int lane = threadIdx.x & 31;
  const double ig = pow2neg(p.sg), ih = pow2neg(p.sh);
  const float root_gain = __double2float_rn(calc_gain(__dmul_rn((double)Gt, ig), __dmul_rn((double)Ht, ih), p));
  int64_t g[8], h[8];
  int64_t sg = 0, sh = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    int i = lane * 8 + k;
    longlong2 v = i < nb ? __ldg(reinterpret_cast<const longlong2*>(hist) + i) : make_longlong2(0, 0);
    sg += v.x; sh += v.y;
    g[k] = sg; h[k] = sh;                     // inclusive prefix inside the lane
  }
  int64_t ig_ = sg, ih_ = sh;                 // warp inclusive scan of lane totals
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t a = __shfl_up_sync(0xffffffffu, ig_, o), b = __shfl_up_sync(0xffffffffu, ih_, o);
    if (lane >= o) { ig_ += a; ih_ += b; }
  }
  const int64_t exg = ig_ - sg, exh = ih_ - sh;   // sum of the lanes below
  const int64_t Gf = __shfl_sync(0xffffffffu, ig_, 31), Hf = __shfl_sync(0xffffffffu, ih_, 31);
  // forward: missing -> right; candidate i has left = prefix_incl(i)
  float bc = 0.0f; int bi = -1;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    int i = lane * 8 + k;
    if (i < nb) {
      int64_t gl = exg + g[k], hl = exh + h[k];
      float c = split_chg(gl, hl, Gt - gl, Ht - hl, ig, ih, root_gain, p);
      if (c > bc) { bc = c; bi = i; }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {          // max chg, ties -> lower bin
    float oc = __shfl_xor_sync(0xffffffffu, bc, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oc > bc || (oc == bc && oi >= 0 && (bi < 0 || oi < bi))) { bc = oc; bi = oi; }
  }
  float best_c = bc; int best_bin = bi, best_dl = 0;
  if (Gf != Gt || Hf != Ht) {
    // backward: missing -> left; candidate i has right = sum of bins >= i; scan order is i descending
    float cc = 0.0f; int ci = -1;
#pragma unroll
    for (int k = 7; k >= 0; --k) {
      int i = lane * 8 + k;
      if (i < nb) {
        int64_t pg = exg + (k ? g[k - 1] : 0), ph = exh + (k ? h[k - 1] : 0);   // prefix_excl(i)
        int64_t gr = Gf - pg, hr = Hf - ph;
        float c = split_chg(Gt - gr, Ht - hr, gr, hr, ig, ih, root_gain, p);
        if (c > cc) { cc = c; ci = i; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {        // max chg, ties -> higher bin (earlier in scan order)
      float oc = __shfl_xor_sync(0xffffffffu, cc, o);
      int oi = __shfl_xor_sync(0xffffffffu, ci, o);
      if (oc > cc || (oc == cc && oi > ci)) { cc = oc; ci = oi; }
    }
    if (cc > best_c) { best_c = cc; best_bin = ci - 1; best_dl = 1; }
  }
  FeatBest out;
  out.chg = best_c; out.split_bin = best_bin; out.default_left = best_dl; out.pad = 0; out.GL = 0; out.HL = 0;
  if (best_c > 0.0f) {
    // left sums of the winning candidate: prefix_incl(best_bin) (+ missing when default_left)
    int src = best_bin >= 0 ? best_bin >> 3 : 0;
    int64_t gl = 0, hl = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (lane * 8 + k == best_bin) { gl = exg + g[k]; hl = exh + h[k]; }
    gl = __shfl_sync(0xffffffffu, gl, src); hl = __shfl_sync(0xffffffffu, hl, src);
    if (best_bin < 0) { gl = 0; hl = 0; }
    if (best_dl) { gl += Gt - Gf; hl += Ht - Hf; }
    out.GL = gl; out.HL = hl;
  }
  return out;
}

// =====================================================================================
// K6  stable partition, standalone single-segment form (the trainer has the level-wide form)
// =====================================================================================
__device__ __forceinline__ bool goes_left(const BinsView& v, uint32_t row, int pos, int split_bin, int default_left) {
  int c = __ldg(v.base + byte_index(v, row, pos));
  return c == CB_MISSING_BIN ? default_left != 0 : c <= split_bin;
}

__global__ void __launch_bounds__(256) part1_flags_kernel(BinsView view,
                                                          const uint32_t* __restrict__ seg, int64_t count, int feat,
                                                          int split_bin, int default_left,
                                                          uint32_t* __restrict__ flag_bits, int* __restrict__ tile_cnt) {
  __shared__ int wsum[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t tile0 = (int64_t)blockIdx.x * kPartTile;
  int cnt = 0;
  for (int it = 0; it < kPartTile / 256; ++it) {
    int64_t pos = tile0 + (int64_t)wid * (kPartTile / 8) + it * 32 + lane;
    bool fl = false;
    if (pos < count) fl = goes_left(view, seg[pos], feat, split_bin, default_left);
    uint32_t bal = __ballot_sync(0xffffffffu, fl);
    if (lane == 0 && pos < count) flag_bits[pos >> 5] = bal;
    cnt += __popc(bal);
  }
  if (lane == 0) wsum[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += wsum[k];
    tile_cnt[blockIdx.x] = t;
  }
}

// one block: exclusive scan of tile counts in place; total -> *n_left
__global__ void __launch_bounds__(1024) part2_scan_kernel(int* __restrict__ tile_cnt, int n_tiles,
                                                          int64_t* __restrict__ n_left) {
  __shared__ int part[1024];
  const int per = (n_tiles + 1023) / 1024;
  const int b = threadIdx.x * per, e = min(n_tiles, b + per);
  int s = 0;
  for (int i = b; i < e; ++i) s += tile_cnt[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  int run = part[threadIdx.x] - s;
  for (int i = b; i < e; ++i) { int c = tile_cnt[i]; tile_cnt[i] = run; run += c; }
  if (threadIdx.x == 1023 && n_left) *n_left = part[1023];
}

__global__ void __launch_bounds__(256) part3_scatter_kernel(const uint32_t* __restrict__ seg, int64_t count,
                                                            const uint32_t* __restrict__ flag_bits,
                                                            const int* __restrict__ tile_ex,
                                                            const int64_t* __restrict__ n_left,
                                                            uint32_t* __restrict__ out) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t tile0 = (int64_t)blockIdx.x * kPartTile;
  const int64_t w0 = tile0 >> 5;                 // first flag word of the tile
  const int wpw = kPartTile / 8 / 32;            // flag words per warp (8)
  // lefts in this tile before this warp's first position
  int before = 0;
  for (int k = lane; k < wid * wpw; k += 32) {
    int64_t wi = w0 + k;
    if (wi * 32 < count) before += __popc(flag_bits[wi]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
  int64_t run = (int64_t)tile_ex[blockIdx.x] + before;
  const int64_t nl = *n_left;
  for (int it = 0; it < wpw; ++it) {
    int64_t pos = tile0 + (int64_t)wid * (kPartTile / 8) + it * 32 + lane;
    if (pos - lane >= count) break;
    uint32_t bal = flag_bits[pos >> 5];
    if (pos < count) {
      bool fl = (bal >> lane) & 1u;
      int64_t lbefore = run + __popc(bal & ((1u << lane) - 1u));
      int64_t dest = fl ? lbefore : nl + (pos - lbefore);
      out[dest] = seg[pos];
    }
    run += __popc(bal);
  }
}

// =====================================================================================
// K8  ensemble traversal: 256 rows per CTA staged feature-major in shared memory
//     (address = f * 256 + row -> bank = lane: conflict free whatever the feature)
// =====================================================================================
constexpr int kPredRows = 256;

__global__ void __launch_bounds__(kPredRows) predict_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx,
                                                            const uint2* __restrict__ nodes,
                                                            const int32_t* __restrict__ tree_off, int n_trees,
                                                            float base_margin, float* __restrict__ out_margin,
                                                            float* __restrict__ out_proba) {
  extern __shared__ float xs[];   // [f][kPredRows]
  const int64_t row0 = (int64_t)blockIdx.x * kPredRows;
  const int nrow = (int)min((int64_t)kPredRows, n - row0);
  if (ldx == f) {                 // contiguous tile: coalesced read, transposing store
    const float* src = X + row0 * ldx;
    for (int i = threadIdx.x; i < nrow * f; i += kPredRows) {
      int r = i / f, c = i - r * f;
      xs[c * kPredRows + r] = __ldg(src + i);
    }
  } else {
    for (int i = threadIdx.x; i < nrow * f; i += kPredRows) {
      int r = i / f, c = i - r * f;
      xs[c * kPredRows + r] = __ldg(X + (row0 + r) * ldx + c);
    }
  }
  __syncthreads();
  const int r = threadIdx.x;
  if (r >= nrow) return;
  float m = base_margin;
  for (int t = 0; t < n_trees; ++t) {
    const uint2* tn = nodes + __ldg(tree_off + t);
    uint2 nd = __ldg(tn);
    while (nd.y >> 16) {
      float v = xs[(nd.y & 0x7FFFu) * kPredRows + r];
      bool left = (v != v) ? ((nd.y >> 15) & 1u) : (v < __uint_as_float(nd.x));
      nd = __ldg(tn + (nd.y >> 16) + (left ? 0u : 1u));
    }
    m = __fadd_rn(m, __uint_as_float(nd.x));
  }
  if (out_margin) out_margin[row0 + r] = m;
  if (out_proba) out_proba[row0 + r] = spec_sigmoid(m);
}

// ---- K8, complete-tree form: every tree is padded to a complete binary tree of depth D in heap
// order (node i -> children 2i+1, 2i+2; pass-through nodes where a branch ended early), so the
// nodes of one level are contiguous: a warp's 32 lanes read <= 32 consecutive threshold words
// -> one shared-memory wavefront, no child pointers.  Tree chunks are staged in shared memory
// next to the feature-major row tile; every thread walks 4 trees at a time (independent
// dependency chains) and adds the 4 leaves in tree order, so the f32 sum equals the oracle's.
// --- minimal mbarrier / 1-D bulk-copy (TMA) wrappers: one thread streams the next tree chunk
// from L2 into shared memory while the CTA walks the current one.
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "W_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@!p bra W_%=;\n\t}\n" ::"r"(a),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(dst)),
               "l"(src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
               : "memory");
}

// nodes: [Tp][NI] uint2 {threshold bits, byte offset of the feature's row in the tile | default_left << 31}
template <int ILP>
__global__ void __launch_bounds__(256) predict_heap_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx,
                                                           const uint2* __restrict__ nodes,    // [Tp][NI]
                                                           const float* __restrict__ leaf,     // [Tp][NL]
                                                           int n_trees, int depth, int chunk_trees, float base_margin,
                                                           float* __restrict__ out_margin, float* __restrict__ out_proba) {
  extern __shared__ __align__(128) float xs[];  // [f][R] row tile, then two tree-chunk buffers
  __shared__ __align__(8) uint64_t bars[2];
  const int R = blockDim.x;
  const int NI = (1 << depth) - 1, NL = 1 << depth;
  const uint32_t node_b = (uint32_t)chunk_trees * NI * 8, leaf_b = (uint32_t)chunk_trees * NL * 4;  // chunk % 2 == 0
  const uint32_t buf_b = node_b + leaf_b;
  char* buf0 = reinterpret_cast<char*>(xs) + (((size_t)f * R * 4 + 127) & ~(size_t)127);
  const int n_chunks = (n_trees + chunk_trees - 1) / chunk_trees;   // arrays are padded to whole chunks
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int c) {
    char* dst = buf0 + (size_t)(c & 1) * buf_b;
    uint64_t* bar = &bars[c & 1];
    mbar_expect_tx(bar, buf_b);
    bulk_g2s(dst, nodes + (size_t)c * chunk_trees * NI, node_b, bar);
    bulk_g2s(dst + node_b, leaf + (size_t)c * chunk_trees * NL, leaf_b, bar);
  };
  if (threadIdx.x == 0) {
    issue(0);
    if (n_chunks > 1) issue(1);
  }
  const int64_t row0 = (int64_t)blockIdx.x * R;
  const int nrow = (int)min((int64_t)R, n - row0);
  if (ldx == f) {
    const float* src = X + row0 * ldx;
    for (int i = threadIdx.x; i < nrow * f; i += R) {
      int r = i / f, c = i - r * f;
      xs[c * R + r] = __ldg(src + i);
    }
  } else {
    for (int i = threadIdx.x; i < nrow * f; i += R) {
      int r = i / f, c = i - r * f;
      xs[c * R + r] = __ldg(X + (row0 + r) * ldx + c);
    }
  }
  __syncthreads();
  const int r = threadIdx.x;
  float m = base_margin;
  for (int c = 0; c < n_chunks; ++c) {
    const char* buf = buf0 + (size_t)(c & 1) * buf_b;
    const uint2* s_node = reinterpret_cast<const uint2*>(buf);
    const float* s_leaf = reinterpret_cast<const float*>(buf + node_b);
    mbar_wait(&bars[c & 1], (uint32_t)((c >> 1) & 1));
    const int ct = min(chunk_trees, n_trees - c * chunk_trees);
    if (r < nrow) {
      // byte offsets inside the chunk buffer: node address a -> children at 2a + (8 - tb) (+8 for right)
      const uint32_t r4 = (uint32_t)r * 4u;     // low bits of every feature-row offset are free (R * 4 <= 1024 | offsets)
      int t = 0;
      for (; t + ILP <= ct; t += ILP) {
        uint32_t a[ILP], c1[ILP];
#pragma unroll
        for (int k = 0; k < ILP; ++k) { a[k] = (uint32_t)(t + k) * NI * 8u; c1[k] = 16u - a[k]; }
        for (int d = 0; d < depth; ++d) {
          uint2 q[ILP];
          float v[ILP];
#pragma unroll
          for (int k = 0; k < ILP; ++k) q[k] = *reinterpret_cast<const uint2*>(buf + a[k]);
#pragma unroll
          for (int k = 0; k < ILP; ++k)
            v[k] = *reinterpret_cast<const float*>(reinterpret_cast<const char*>(xs) + ((q[k].y & 0x7FFFFFFFu) | r4));
#pragma unroll
          for (int k = 0; k < ILP; ++k) {
            // left = (v < thr) | (default_left & isnan(v));  a = 2a + c1 (right child), minus 8 when left.
            // Written with predicates: the ALU pipe issues at half rate and bounds this loop.
            asm("{\n\t.reg .pred pd, pn, pl;\n\t"
                "setp.lt.s32 pd, %2, 0;\n\t"
                "setp.nan.and.f32 pn, %1, %1, pd;\n\t"
                "setp.lt.or.f32 pl, %1, %3, pn;\n\t"
                "mad.lo.u32 %0, %0, 2, %4;\n\t"
                "@pl sub.u32 %0, %0, 8;\n\t}"
                : "+r"(a[k])
                : "f"(v[k]), "r"(q[k].y), "f"(__uint_as_float(q[k].x)), "r"(c1[k]));
          }
        }
#pragma unroll
        for (int k = 0; k < ILP; ++k) {
          const uint32_t li = ((a[k] + c1[k] - 16u) >> 1) - (uint32_t)NI * 4u;   // (idx - NI) * 4
          m = __fadd_rn(m, *reinterpret_cast<const float*>(buf + node_b + (uint32_t)(t + k) * NL * 4u + li));
        }
      }
      for (; t < ct; ++t) {
        int i0 = 0;
        for (int d = 0; d < depth; ++d) {
          const uint2 q0 = s_node[t * NI + i0];
          const float v0 = *reinterpret_cast<const float*>(reinterpret_cast<const char*>(xs) + ((q0.y & 0x7FFFFFFFu) | r4));
          const uint32_t l0 = (uint32_t)(v0 < __uint_as_float(q0.x)) | ((q0.y >> 31) & (uint32_t)(v0 != v0));
          i0 = 2 * i0 + 2 - (int)l0;
        }
        m = __fadd_rn(m, s_leaf[t * NL + i0 - NI]);
      }
    }
    __syncthreads();                               // every thread is done with this buffer
    if (threadIdx.x == 0 && c + 2 < n_chunks) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(c + 2);
    }
  }
  if (r < nrow) {
    if (out_margin) out_margin[row0 + r] = m;
    if (out_proba) out_proba[row0 + r] = spec_sigmoid(m);
  }
}

// ---------------------------------------------------------------------------------------------------
// Scoring on quantised inputs.  A split "x < thr" only ever compares x with one of the forest's own
// thresholds, so a row can be reduced once to one byte per feature -- code = number of that feature's
// distinct thresholds <= x (K2's binary search against the forest's threshold table), 255 = missing --
// and "x < t_k" becomes "code <= k", bit for bit the same decision.  That makes a node 4 bytes instead
// of 8 and a feature value 1 byte instead of 4: 2 shared-memory wavefronts per (warp, level) instead of
// ~3.75, and 2 ALU-pipe instructions instead of 5 (the two resources the f32 kernel is bound by).
//   codes  : row major [n][ldc] bytes (ldc a multiple of 4), written by K2 a batch at a time and still in L2
//            when the walk reads them.
//   tile   : two byte planes [f][128]: plane 0 holds the codes (missing = 255, never <= span), plane 1
//            holds code + 1 with missing = 0 (always <= span): default-left nodes read plane 1, so the
//            missing rule costs no instruction.  Row r of a column sits at byte 4 * (r % 32) + r / 32:
//            a warp's 32 rows are in 32 different banks whatever features its lanes look at.
//   node   : u32 = span << 16 | byte offset of (plane, feature); go left iff (code << 16) <= node.
//            Heap order, 1-based: tree t occupies 8 << depth bytes, node i at byte 4i, children at
//            8i / 8i + 4, and the 2^depth leaves (f32) simply follow as i = 2^depth .. 2^(depth+1)-1:
//            after `depth` steps the walk address IS the leaf address.  A level of <= 32 nodes is
//            <= 32 consecutive words: node fetches are conflict free up to depth 6.
// K2 for scoring: codes of a 32-feature slab of X, row major.  The slab's threshold table sits in shared memory
// (32 x 257 floats: the odd stride spreads the lanes' probes over the banks), lane = feature, so X is read in
// 128-byte row segments and the 8 probes of the (branch-free, fixed-length) binary search never leave the SM --
// the general K2 kernel probes the table through L1 with 32 different sectors per warp load and takes 2.5x longer.
__global__ void __launch_bounds__(256) bin_rows_q8_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx,
                                                          const float* __restrict__ cuts, const int32_t* __restrict__ n_cuts,
                                                          uint8_t* __restrict__ codes, int64_t ldc, int rows_per_cta) {
  __shared__ float tab[32 * 257];
  const int c0 = blockIdx.y * 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 32 * 256; i += 256) {
    const int cf = i >> 8, k = i & 255;
    const bool real = c0 + cf < f && k < __ldg(n_cuts + c0 + cf);        // everything else is +inf whatever the caller padded with
    tab[cf * 257 + k] = real ? __ldg(cuts + (size_t)(c0 + cf) * CB_CUT_STRIDE + k) : __int_as_float(0x7f800000);
  }
  __syncthreads();
  const int c = c0 + lane;
  const bool on = c < f;
  const int nc = on ? __ldg(n_cuts + c) : 1;
  const int nc_max = __reduce_max_sync(0xffffffffu, nc);   // probes with a stride above every table length cannot succeed
  const float* my = tab + lane * 257;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta, r1 = min(n, r0 + rows_per_cta);
  for (int64_t rb = r0 + warp * 4; rb < r1; rb += 32) {       // 8 warps x 4 rows in flight
    float x[4];
    const float* lo[4];                                       // &table[number of thresholds <= x found so far]
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t r = min(rb + u, r1 - 1);
      x[u] = on ? __ldg(X + r * ldx + c) : 0.0f;
      lo[u] = my;
    }
    // branch-free binary search, 3 instructions per probe: the table is padded with +inf up to 256 entries, so a probe
    // past the feature's last threshold simply fails (it succeeds only for x = +inf, which the final clamp undoes)
#pragma unroll
    for (int s = 128; s > 0; s >>= 1) {
      if (s > nc_max) continue;                              // warp uniform: a slab of one-hot columns needs one probe
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (lo[u][s - 1] <= x[u]) lo[u] += s;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t r = rb + u;
      if (on && r < r1) {
        uint32_t code = (uint32_t)min((int)(lo[u] - my), nc - 1);   // same clamp as K2
        if (x[u] != x[u]) code = CB_MISSING_BIN;
        codes[r * ldc + c] = (uint8_t)code;
      }
    }
  }
}

// shared-window address of the dynamic shared memory of a kernel without static shared memory; the answer goes to a
// word in the module's device image (read back with cudaMemcpyFromSymbol: the probe allocates nothing)
__device__ uint32_t g_smem_base_probe;
__global__ void smem_base_probe_kernel() {
  extern __shared__ __align__(128) uint8_t probe_s[];
  g_smem_base_probe = (uint32_t)__cvta_generic_to_shared(probe_s);
}

// K7, standalone form: every row of a leaf's segment takes the leaf value (R10: m_r += leaf[node(r)])
__global__ void __launch_bounds__(256) leaf_update_kernel(float* __restrict__ margin, const uint32_t* __restrict__ rows,
                                                          int64_t count, float leaf) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const int64_t r = rows ? (int64_t)rows[i] : i;
  margin[r] = __fadd_rn(margin[r], leaf);
}

constexpr int kQ8Stages = 2;               // tree-chunk buffers: one being walked, one in flight
// one level of one tree walk.  a = shared address of the node; children of the node at S + 4i are at
// 2 * (S + 4i) - S (+ 4): cl = -S, cr = 4 - S.  The shift and both address updates run on the FMA pipe; the ALU
// pipe (half rate) only sees the offset mask/or and the compare.
__device__ __forceinline__ void q8_step(uint32_t& a, uint32_t x, uint32_t q, uint32_t two, uint32_t cl, uint32_t cr) {
  asm("{\n\t.reg .pred pl;\n\t.reg .u32 xs;\n\t"
      "mul.lo.u32 xs, %1, 65536;\n\t"
      "setp.le.u32 pl, xs, %2;\n\t"
      "@pl mad.lo.u32 %0, %0, %3, %4;\n\t"
      "@!pl mad.lo.u32 %0, %0, %3, %5;\n\t}"
      : "+r"(a)
      : "r"(x), "r"(q), "r"(two), "r"(cl), "r"(cr));
}

template <int ILP, int DEPTH>
__global__ void __launch_bounds__(128, 4) predict_q8_kernel(const uint8_t* __restrict__ codes, int64_t n, int f, int64_t ldc,
                                                         const uint32_t* __restrict__ forest, int n_trees,
                                                         int chunk_trees, float base_margin, uint32_t plane_base,
                                                         float* __restrict__ out_margin, float* __restrict__ out_proba) {
  // planes [2][f][128] | kQ8Stages tree-chunk buffers | mbarriers.  No static shared memory:
  // the planes start at the bottom of the CTA's shared window, whose address (plane_base, cb_predict_q8_plan) is
  // already added to every node's offset field, so "(node & 0xFFFF) | row byte" IS the shared address of the code.
  extern __shared__ __align__(128) uint8_t q8s[];
  if ((uint32_t)__cvta_generic_to_shared(q8s) != plane_base) {
    // the forest was laid out for another shared window (the host checks plane_base before launching, so this cannot
    // happen through the C ABI): answer NaN rather than walk wrong addresses or poison the context with a trap
    const int64_t rr = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (rr < n) {
      if (out_margin) out_margin[rr] = __int_as_float(0x7fc00000);
      if (out_proba) out_proba[rr] = __int_as_float(0x7fc00000);
    }
    return;
  }
  constexpr int R = 128;
  constexpr int depth = DEPTH;
  constexpr uint32_t TB = 8u << DEPTH;              // bytes per tree
  const uint32_t plane_b = (uint32_t)f * R;
  uint8_t* buf0 = q8s + ((2 * plane_b + 127) & ~127u);
  const uint32_t buf_b = (uint32_t)chunk_trees * TB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(buf0 + (size_t)kQ8Stages * buf_b);
  const int n_chunks = (n_trees + chunk_trees - 1) / chunk_trees;   // the array is padded to whole chunks
  const int64_t row0 = (int64_t)blockIdx.x * R;
  const int nrow = (int)min((int64_t)R, n - row0);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < kQ8Stages; ++k) mbar_init(&bars[k], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int c) {
    uint64_t* bar = &bars[c % kQ8Stages];
    mbar_expect_tx(bar, buf_b);
    bulk_g2s(buf0 + (size_t)(c % kQ8Stages) * buf_b, reinterpret_cast<const uint8_t*>(forest) + (size_t)c * buf_b, buf_b, bar);
  };
  if (threadIdx.x == 0) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int c = 0; c < kQ8Stages && c < n_chunks; ++c) issue(c);   // the first chunks travel while the tile is filled
  }
  const uint32_t r = threadIdx.x;
  const uint32_t pr = 4u * (r & 31u) + (r >> 5);     // this row's byte inside a feature column
  if ((int)r < nrow) {
    // tile fill straight from global memory, lane = row: a warp load touches 32 rows x 4 bytes (32 sectors), and
    // the next 7 loads of the warp hit the same sectors in L1; the byte stores go to 32 different banks
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(codes + (row0 + r) * ldc);
    for (int w0 = 0; 4 * w0 < f; w0 += 8) {
      uint32_t v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (4 * (w0 + u) < f) ? __ldg(mine + w0 + u) : 0u;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = 4 * (w0 + u) + j;
          if (c < f) {
            const uint32_t bb = (v[u] >> (8 * j)) & 255u;
            q8s[(uint32_t)c * R + pr] = (uint8_t)bb;
            q8s[plane_b + (uint32_t)c * R + pr] = (uint8_t)(bb + 1u);   // 255 (missing) wraps to 0
          }
        }
      }
    }
  }
  __syncthreads();                                   // planes complete
  float m = base_margin;
  for (int c = 0; c < n_chunks; ++c) {
    const uint8_t* buf = buf0 + (size_t)(c % kQ8Stages) * buf_b;
    mbar_wait(&bars[c % kQ8Stages], (uint32_t)((c / kQ8Stages) & 1));
    const int ct = min(chunk_trees, n_trees - c * chunk_trees);
    {
      // Every thread walks, also those past the last row of a partial tile (their codes are whatever the planes
      // hold; all addresses stay inside the planes and the tree).  Walk addresses are absolute shared-memory
      // addresses, so a visit is: load node, mask|or, load code, shift, compare, one of two multiply-adds.
      uint32_t two;
      asm volatile("mov.u32 %0, 2;" : "=r"(two));  // a register on purpose: keeps the updates IMADs (FMA pipe)
      const uint32_t sbuf = (uint32_t)__cvta_generic_to_shared(buf);
      int t = 0;
      for (; t + ILP <= ct; t += ILP) {
        uint32_t a[ILP], cl[ILP], cr[ILP];
#pragma unroll
        for (int k = 0; k < ILP; ++k) {
          const uint32_t S = sbuf + (uint32_t)(t + k) * TB;
          a[k] = S + 4u; cl[k] = 0u - S; cr[k] = 4u - S;
        }
#pragma unroll 1                                  // one level at a time: unrolled, ptxas walks tree after tree (no ILP)
        for (int d = 0; d < depth; ++d) {
          uint32_t q[ILP], x[ILP];
#pragma unroll
          for (int k = 0; k < ILP; ++k) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(q[k]) : "r"(a[k]));
#pragma unroll
          for (int k = 0; k < ILP; ++k) asm volatile("ld.shared.u8 %0, [%1];" : "=r"(x[k]) : "r"((q[k] & 0xFFFFu) | pr));
#pragma unroll
          for (int k = 0; k < ILP; ++k) q8_step(a[k], x[k], q[k], two, cl[k], cr[k]);
        }
#pragma unroll
        for (int k = 0; k < ILP; ++k) {
          float lv;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(lv) : "r"(a[k]));
          m = __fadd_rn(m, lv);
        }
      }
      for (; t < ct; ++t) {
        uint32_t a0 = (uint32_t)t * TB + 4u;
        const uint32_t cl0 = 0u - (uint32_t)t * TB;
#pragma unroll 1
        for (int d = 0; d < depth; ++d) {
          const uint32_t q0 = *reinterpret_cast<const uint32_t*>(buf + a0);
          const uint32_t x0 = q8s[((q0 & 0xFFFFu) | pr) - plane_base];
          a0 = 2u * a0 + cl0 + (((x0 << 16) <= q0) ? 0u : 4u);
        }
        m = __fadd_rn(m, *reinterpret_cast<const float*>(buf + a0));
      }
    }
    __syncthreads();                               // every thread is done with this buffer
    if (threadIdx.x == 0 && c + kQ8Stages < n_chunks) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(c + kQ8Stages);
    }
  }
  if ((int)r < nrow) {
    if (out_margin) out_margin[row0 + r] = m;
    if (out_proba) out_proba[row0 + r] = spec_sigmoid(m);
  }
}

}  // namespace cb
