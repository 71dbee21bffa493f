// K4, bank-conflict-free form.  Included by cobalt_gbdt.cu only.
//
// The shared-memory atomic pipe bounds the histogram build (ncu: l1tex data pipe 91 % busy,
// 4.1 wavefronts per ATOMS with a compact slot = bin_ptr[f] + bin layout, profiles/r01).
// Here every lane of a warp owns one shared-memory BANK: the histogram of a feature that
// lane `l` accumulates lives at word (rowbase + bin) * 32 + l, so the 32 addresses of one
// ATOMS instruction always fall in 32 different banks whatever the data -> 1 wavefront.
//
// To make that possible the columns of the binned matrix are permuted (pos_feature): a row is
// split into column blocks of <= 32 u32 words; in block k lane l reads word (l % wk) of row
// slot (l / wk), i.e. a narrow last block processes ck = 32 / wk rows per instruction into ck
// private copies that are folded before the flush.  Features are dealt to words by descending
// bin count so that bank fill is balanced.  Global histograms keep the natural feature order.
#pragma once
#include <algorithm>
#include <vector>

#include "kernels.cuh"

namespace cb {

constexpr int kCfMaxBlocks = 4;        // up to 512 features

struct CfLayout {                      // device copy lives in the trainer workspace
  int n_blocks;
  int R;                               // shared-memory rows of 32 banks (per G and per H array)
  int n_fold;                          // words to fold from row-slot copies into copy 0
  int pad_;
  int wk[kCfMaxBlocks];                // words in the block
  int ck[kCfMaxBlocks];                // rows per warp instruction (= private copies)
  int a0[kCfMaxBlocks][32][4];         // word index of bin 0 of the byte's feature, -1 = padding byte
  int nbj[kCfMaxBlocks][32][4];        // bins of the byte's feature: codes are clamped to it, row nbj = missing/dump
};

struct CfHostTables {
  CfLayout lay;
  std::vector<int32_t> slot_of;        // [R*32] global slot of a shared-memory word, -1 = none / copy
  std::vector<int32_t> fold_src, fold_dst;
};

// Storage columns.  A binned row holds one byte per COLUMN; a column is either a plain feature or a GROUP of
// mutually exclusive 2-bin features (one-hot dummies, reference src/data_preprocessing/feature_engineering.py:142-147)
// packed as one code: 0 = no member set, m + 1 = member m set.  pos_feature[p] names the column of byte p:
// >= 0 a plain feature, -1 padding, <= -2 group g = -2 - value.  A group of k members is one histogram of k + 1
// bins: one pair of shared-memory atomics per row instead of k pairs.
struct GroupTable {                    // host
  int n_groups = 0;
  const int32_t* ptr = nullptr;        // [n_groups + 1]
  const int32_t* feat = nullptr;       // members, group after group; the position inside a group + 1 is the code
  int size(int g) const { return ptr[g + 1] - ptr[g]; }
};
inline int group_key(int g) { return -2 - g; }
inline bool is_group_key(int v) { return v <= -2; }
inline int group_of_key(int v) { return -2 - v; }

// false when the groups are malformed (a feature twice, out of range, or not a 2-bin feature)
inline bool check_groups(const int32_t* n_cuts, int f, const GroupTable& gt, std::vector<int>* group_of) {
  group_of->assign(f, -1);
  for (int g = 0; g < gt.n_groups; ++g) {
    if (gt.size(g) < 1 || gt.size(g) > 254) return false;
    for (int k = gt.ptr[g]; k < gt.ptr[g + 1]; ++k) {
      const int ft = gt.feat[k];
      if (ft < 0 || ft >= f || (*group_of)[ft] >= 0 || n_cuts[ft] != 2) return false;
      (*group_of)[ft] = g;
    }
  }
  return true;
}

inline int count_columns(int f, const GroupTable& gt) {
  int members = 0;
  for (int g = 0; g < gt.n_groups; ++g) members += gt.size(g);
  return f - members + gt.n_groups;
}

// column positions: columns sorted by descending bin count, dealt round robin to the words of a block
inline bool plan_positions(const int32_t* n_cuts, int f, const GroupTable& gt, int64_t ldb, int32_t* pos_feature) {
  std::vector<int> group_of;
  if (!check_groups(n_cuts, f, gt, &group_of)) return false;
  std::vector<int> order, nb;
  for (int i = 0; i < f; ++i)
    if (group_of[i] < 0) { order.push_back(i); nb.push_back(n_cuts[i]); }
  for (int g = 0; g < gt.n_groups; ++g) { order.push_back(group_key(g)); nb.push_back(gt.size(g) + 1); }
  const int nc = (int)order.size();
  std::vector<int> idx(nc);
  for (int i = 0; i < nc; ++i) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return nb[a] > nb[b]; });
  for (int64_t p = 0; p < ldb; ++p) pos_feature[p] = -1;
  const int W = (nc + 3) / 4;
  if ((int64_t)W * 4 > ldb) return false;
  if (ldb <= 128) {
    // one column block: use every word of the row (fewer columns per lane = fewer shared-memory rows) and deal the
    // sorted columns in serpentine order, which evens out the lanes' bin totals (the tallest lane sets the row count R)
    const int Wa = (int)(ldb / 4);
    for (int r = 0; r < nc; ++r) {
      const int j = r / Wa, q = r % Wa;
      const int word = (j & 1) ? Wa - 1 - q : q;
      pos_feature[word * 4 + j] = order[idx[r]];
    }
    return true;
  }
  int next = 0;
  for (int k = 0; k * 32 < W; ++k) {
    const int wk = std::min(32, W - 32 * k);
    const int take = std::min(nc - next, 4 * wk);
    for (int r = 0; r < take; ++r) {
      const int word = r % wk, j = r / wk;
      pos_feature[(32 * k + word) * 4 + j] = order[idx[next + r]];
    }
    next += take;
  }
  return true;
}

// builds the shared-memory tables for a given column placement; false when it does not fit
inline bool build_cf_tables(const int32_t* bin_ptr, const int32_t* pos_feature, int f, const GroupTable& gt, int64_t ldb,
                            int budget_bytes, CfHostTables* out) {
  int last = -1;                                          // last byte position in use
  for (int64_t p = 0; p < ldb; ++p)
    if (pos_feature[p] != -1) last = (int)p;
  const int W = last / 4 + 1;
  if (last < 0 || (int64_t)W * 4 > ldb) return false;
  const int nb = (W + 31) / 32;
  if (nb > kCfMaxBlocks) return false;
  auto col_bins = [&](int col) { return is_group_key(col) ? gt.size(group_of_key(col)) + 1 : bin_ptr[col + 1] - bin_ptr[col]; };
  CfLayout& L = out->lay;
  memset(&L, 0, sizeof L);
  L.n_blocks = nb;
  int bank_rows[32] = {0};
  struct Rec { int k, l, j, rowbase, col; };
  std::vector<Rec> recs;
  for (int k = 0; k < nb; ++k) {
    const int wk = std::min(32, W - 32 * k);
    const int ck = 32 / wk;
    L.wk[k] = wk; L.ck[k] = ck;
    for (int l = 0; l < 32; ++l) {
      for (int j = 0; j < 4; ++j) { L.a0[k][l][j] = -1; L.nbj[k][l][j] = 0; }
      if (l >= ck * wk) continue;
      const int word = 32 * k + l % wk;
      for (int j = 0; j < 4; ++j) {
        const int col = pos_feature[word * 4 + j];
        if (col == -1) continue;
        if (col >= f || (is_group_key(col) && group_of_key(col) >= gt.n_groups)) return false;
        const int nbins = col_bins(col);
        L.a0[k][l][j] = bank_rows[l] * 32 + l;
        L.nbj[k][l][j] = nbins;
        recs.push_back(Rec{k, l, j, bank_rows[l], col});
        bank_rows[l] += nbins + 1;                      // + the row that takes missing values
      }
    }
  }
  int R = 1;
  for (int l = 0; l < 32; ++l) R = std::max(R, bank_rows[l]);
  R += 1;                                               // last row: dump slot of padding bytes
  L.R = R;
  for (int k = 0; k < nb; ++k)
    for (int l = 0; l < 32; ++l)
      for (int j = 0; j < 4; ++j)
        if (L.a0[k][l][j] < 0) { L.a0[k][l][j] = (R - 1) * 32 + l; L.nbj[k][l][j] = 0; }
  if (R > 896 || (int64_t)2 * 28672 * 4 > budget_bytes + 2048) return false;   // kCfMaxRows, kCfSmemBytes
  out->slot_of.assign((size_t)R * 32, -1);
  out->fold_src.clear(); out->fold_dst.clear();
  // global slot of code b of a column: a plain feature's bin b; for a group, code m + 1 is bin 1 of member m
  // (code 0, "no member set", has no slot: every member's bin 0 is node total - bin 1, derived in the split scan)
  auto slot_of_code = [&](int col, int b) {
    if (!is_group_key(col)) return bin_ptr[col] + b;
    if (b == 0) return -1;
    const int g = group_of_key(col);
    return bin_ptr[gt.feat[gt.ptr[g] + b - 1]] + 1;
  };
  for (const Rec& r : recs) {
    const int wk = L.wk[r.k];
    const int nbins = col_bins(r.col);
    if (r.l < wk) {
      for (int b = 0; b < nbins; ++b) out->slot_of[(size_t)(r.rowbase + b) * 32 + r.l] = slot_of_code(r.col, b);
    } else {
      const int l0 = r.l % wk;
      const int dst0 = L.a0[r.k][l0][r.j];
      for (int b = 0; b < nbins; ++b) {
        out->fold_src.push_back((r.rowbase + b) * 32 + r.l);
        out->fold_dst.push_back(dst0 + b * 32);
      }
    }
  }
  L.n_fold = (int)out->fold_src.size();
  // every column must have been placed exactly once (grouped features through their group only)
  std::vector<int> group_of;
  std::vector<int32_t> nc(f);
  for (int i = 0; i < f; ++i) nc[i] = bin_ptr[i + 1] - bin_ptr[i];
  if (!check_groups(nc.data(), f, gt, &group_of)) return false;
  std::vector<int> seen_f(f, 0), seen_g(gt.n_groups, 0);
  for (int64_t p = 0; p < (int64_t)W * 4; ++p) {
    const int col = pos_feature[p];
    if (col >= 0) seen_f[col]++;
    else if (is_group_key(col)) seen_g[group_of_key(col)]++;
  }
  for (int i = 0; i < f; ++i)
    if (seen_f[i] != (group_of[i] < 0 ? 1 : 0)) return false;
  for (int g = 0; g < gt.n_groups; ++g)
    if (seen_g[g] != 1) return false;
  return true;
}

constexpr int kCfHOffWords = 28672;            // H array at a fixed +112 KB: an immediate in the ATOMS address
constexpr int kCfMaxRows = kCfHOffWords / 32;  // 896 rows of 32 banks per array
constexpr int kCfSmemBytes = 2 * kCfHOffWords * 4;
__device__ __forceinline__ uint32_t ldg_nc_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// one byte of a bin word: clamp the code to the feature's bin count (255 = missing lands in the
// row after the last bin, which the reduce step ignores) and add (g, h) in the lane's own bank.
// 5 issue slots per byte: PRMT, VIMNMX, IMAD, 2 x ATOMS; no predicates, no branches.
template <int POS>
__device__ __forceinline__ void cf_add_byte_at(uint32_t word, uint32_t nb, uint32_t base, int g, int h) {
  asm volatile(
      "{\n\t"
      ".reg .u32 b, a;\n\t"
      "bfe.u32 b, %0, %1, 8;\n\t"
      "min.u32 b, b, %2;\n\t"
      "mad.lo.u32 a, b, 128, %3;\n\t"
      "red.shared.add.s32 [a], %4;\n\t"
      "red.shared.add.s32 [a+114688], %5;\n\t"
      "}\n"
      :
      : "r"(word), "n"(POS), "r"(nb), "r"(base), "r"(g), "r"(h)
      : "memory");
}
// 1024 threads = 32 warps; warp w takes row groups w, w + 32, ... of the CTA's slice, one column block after the other
// (each block is its own contiguous array: block-major storage), UNROLL rows in flight per warp.
//
// The reduction to the int64 level buffer happens in the same launch.  The grid never exceeds one CTA per SM, so
// all of a node's CTAs are resident together and may wait for each other:
//   * a node that one CTA covers in one chunk goes from shared memory straight to the level buffer;
//   * otherwise every CTA stores its int32 partials (shared-memory order, coalesced), the node's CTAs meet at a
//     counter, and then each of them sums ITS slice of the words over all of the node's partials
//     (int32 -> int64, un-permuted through slot_of).
// No second kernel, no zeroed buffer, no atomics on the level buffer: every slot is written exactly once.
template <int UNROLL, bool IDENTITY>
__global__ void __launch_bounds__(1024, 1)
hist_accum_cf_kernel(const uint8_t* __restrict__ bins, int64_t n_rows, int last_w, const int2* __restrict__ gpair,
                     const uint32_t* __restrict__ row_idx, const HistPlan* __restrict__ plan,
                     const CfLayout* __restrict__ lay, const int32_t* __restrict__ fold_src,
                     const int32_t* __restrict__ fold_dst, int32_t* __restrict__ partials,
                     const int32_t* __restrict__ slot_of, int S, int64_t* __restrict__ level_buf,
                     unsigned int* __restrict__ done, unsigned int* __restrict__ wait_timeout,
                     const long long* __restrict__ root_tot, int64_t* __restrict__ root_tail) {
  CB_LEVEL_TRACE_SCOPE(5);
  pdl_enter();
  extern __shared__ __align__(16) int32_t smem[];
  if (root_tail && blockIdx.x == 0 && threadIdx.x == 0) { root_tail[0] = root_tot[0]; root_tail[1] = root_tot[1]; }
  const int slot = plan->cta_slot[blockIdx.x];
  if (slot < 0) return;
  const int RW = lay->R * 32;                 // used words per array
  int32_t* sG = smem;
  int32_t* sH = smem + kCfHOffWords;
  for (int s = threadIdx.x; s < RW; s += blockDim.x) { sG[s] = 0; sH[s] = 0; }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nblk = lay->n_blocks, n_fold = lay->n_fold;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  __syncthreads();
  const int pos0 = plan->cta_begin[blockIdx.x];
  const int rows = plan->cta_rows[blockIdx.x];
  const int nct = plan->b_ncta[slot];
  const int p0 = plan->b_pbeg[slot], p1 = plan->b_pbeg[slot + 1];
  const bool direct = nct == 1 && p1 - p0 <= 1;            // this CTA alone holds the node's whole histogram
  longlong2* out = reinterpret_cast<longlong2*>(level_buf) + (size_t)slot * S;
  int pid = plan->cta_pbase[blockIdx.x];
  for (int c0 = 0; c0 < rows || (direct && c0 == 0); c0 += kChunkRows, ++pid) {
    const int crow = min(kChunkRows, rows - c0);
    for (int k = 0; k < nblk; ++k) {
      const int wk = lay->wk[k], ck = lay->ck[k];
      const bool active = lane < wk * ck;
      const int rslot = active ? lane / wk : 0;
      const int wl = active ? lane - rslot * wk : 0;
      const int bw = (k == nblk - 1) ? last_w : 128;                 // bytes per row of this block
      const uint8_t* bbase = bins + (size_t)n_rows * 128 * k + 4 * wl;
      uint32_t base[4], nbj[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        base[j] = sbase + 4u * (uint32_t)lay->a0[k][lane][j];
        nbj[j] = (uint32_t)lay->nbj[k][lane][j];
      }
      const int ilast = crow - 1;
      // gathered rows: the row ids of the NEXT step are fetched while this step's rows are in flight, so a step costs
      // one memory latency (rows + gradient pairs), not two in a row (ids, then rows)
      uint32_t rn[UNROLL];
      auto fetch_ids = [&](int g0) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const int i = (g0 + 32 * u) * ck + rslot;
          rn[u] = __ldg(row_idx + pos0 + c0 + max(min(i, ilast), 0));
        }
      };
      if (!IDENTITY && warp * ck < crow) fetch_ids(warp);
      for (int g0 = warp; g0 * ck < crow; g0 += 32 * UNROLL) {
        uint32_t wv[UNROLL];
        int2 gp[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const int i = (g0 + 32 * u) * ck + rslot;
          const int pos = pos0 + c0 + min(i, ilast);                 // clamped: always a valid row, no branch
          const uint32_t r = IDENTITY ? (uint32_t)pos : rn[u];
          wv[u] = ldg_nc_u32(reinterpret_cast<const uint32_t*>(bbase + (size_t)r * bw));
          gp[u] = __ldg(gpair + r);
          if (!active || i > ilast) gp[u] = make_int2(0, 0);         // adds zero: harmless
        }
        if (!IDENTITY && (g0 + 32 * UNROLL) * ck < crow) fetch_ids(g0 + 32 * UNROLL);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          cf_add_byte_at<0>(wv[u], nbj[0], base[0], gp[u].x, gp[u].y);
          cf_add_byte_at<8>(wv[u], nbj[1], base[1], gp[u].x, gp[u].y);
          cf_add_byte_at<16>(wv[u], nbj[2], base[2], gp[u].x, gp[u].y);
          cf_add_byte_at<24>(wv[u], nbj[3], base[3], gp[u].x, gp[u].y);
        }
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < n_fold; e += blockDim.x) {   // fold row-slot copies into copy 0
      const int s = __ldg(fold_src + e), d = __ldg(fold_dst + e);
      const int g = sG[s], h = sH[s];
      if (g) atomicAdd(sG + d, g);
      if (h) atomicAdd(sH + d, h);
    }
    __syncthreads();
    if (direct) {
      for (int s = threadIdx.x; s < RW; s += blockDim.x) {
        const int sl = __ldg(slot_of + s);
        if (sl >= 0) out[sl] = make_longlong2((long long)sG[s], (long long)sH[s]);
      }
      return;
    }
    int32_t* pp = partials + (size_t)pid * 2 * RW;
    const bool more = c0 + kChunkRows < rows;
    for (int s = threadIdx.x; s < RW; s += blockDim.x) {
      pp[s] = sG[s]; pp[RW + s] = sH[s];
      if (more) { sG[s] = 0; sH[s] = 0; }
    }
    __syncthreads();
  }
  // ---- the node's CTAs meet (one fence by the signalling thread, after the CTA barrier, publishes all of the CTA's
  // stores), then each reduces its slice of the words over all of the node's partials
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(done + slot, 1u);
    unsigned spins = 0;
    while (ld_acquire_u32(done + slot) < (unsigned)nct) {
      if (++spins > (1u << 22)) { *wait_timeout = 2u; break; }   // cannot happen while the grid is co-resident
      __nanosleep(32);
    }
  }
  __syncthreads();
  // Slice of this CTA: quads [q0, q1) of the node's words.  Threads form a (quad, partial group) grid so that both
  // a node with many partials and few quads per CTA (the root) and one with few partials and many quads keep every
  // thread busy with independent 16-byte loads; group sums meet in shared memory (free now), one thread per word
  // adds them and writes the slot.
  const int k = (int)blockIdx.x - plan->b_cta0[slot];
  const int nq = RW >> 2, np = p1 - p0;
  const int q0 = (int)((long long)nq * k / nct), q1 = (int)((long long)nq * (k + 1) / nct);
  const int32_t* pbase = partials + (size_t)p0 * 2 * RW;
  long long* sacc = reinterpret_cast<long long*>(smem);          // [G][Q][8], G * Q <= 1024
  // partial groups: enough that no thread walks more than ~8 partials in a row (two rounds of 4 loads in flight),
  // and as few as possible otherwise so that one pass covers up to 1024 quads
  int Gw = (np + 7) >> 3;
  Gw = Gw < 1 ? 1 : (Gw > 32 ? 32 : Gw);
  const int QT = (int)blockDim.x / Gw;                           // quads per pass
  for (int qa = q0; qa < q1; qa += QT) {
    const int Q = min(QT, q1 - qa);
    int G = min(np, (int)blockDim.x / Q);
    G = G < 1 ? 1 : G;
    const int t = threadIdx.x;
    const int pg = t / Q, qi = t - pg * Q;
    if (pg < G) {
      long long g[4] = {0, 0, 0, 0}, h[4] = {0, 0, 0, 0};
      const int32_t* pq = pbase + 4 * (size_t)(qa + qi);
      int p = pg;
      for (; p + 3 * G < np; p += 4 * G) {                       // 4 partials (8 loads) in flight
        int4 a[4], c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          a[u] = __ldcg(reinterpret_cast<const int4*>(pq + (size_t)(p + u * G) * 2 * RW));
          c[u] = __ldcg(reinterpret_cast<const int4*>(pq + (size_t)(p + u * G) * 2 * RW + RW));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          g[0] += a[u].x; g[1] += a[u].y; g[2] += a[u].z; g[3] += a[u].w;
          h[0] += c[u].x; h[1] += c[u].y; h[2] += c[u].z; h[3] += c[u].w;
        }
      }
      for (; p < np; p += G) {
        const int4 a = __ldcg(reinterpret_cast<const int4*>(pq + (size_t)p * 2 * RW));
        const int4 c = __ldcg(reinterpret_cast<const int4*>(pq + (size_t)p * 2 * RW + RW));
        g[0] += a.x; g[1] += a.y; g[2] += a.z; g[3] += a.w;
        h[0] += c.x; h[1] += c.y; h[2] += c.z; h[3] += c.w;
      }
      long long* d = sacc + ((size_t)pg * Q + qi) * 8;
#pragma unroll
      for (int i = 0; i < 4; ++i) { d[2 * i] = g[i]; d[2 * i + 1] = h[i]; }
    }
    __syncthreads();
    for (int w = threadIdx.x; w < 4 * Q; w += blockDim.x) {      // word w of the pass: add the groups, write the slot
      const int sl = __ldg(slot_of + 4 * qa + w);
      if (sl < 0) continue;
      const int qi2 = w >> 2, i = w & 3;
      long long gg = 0, hh = 0;
      for (int e = 0; e < G; ++e) {
        const long long* d = sacc + ((size_t)e * Q + qi2) * 8 + 2 * i;
        gg += d[0]; hh += d[1];
      }
      out[sl] = make_longlong2(gg, hh);
    }
    __syncthreads();
  }
}

}  // namespace cb
