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

// column positions: features sorted by descending bin count, dealt round robin to the words of a block
inline void plan_positions(const int32_t* n_cuts, int f, int64_t ldb, int32_t* pos_feature) {
  std::vector<int> order(f);
  for (int i = 0; i < f; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return n_cuts[a] > n_cuts[b]; });
  for (int64_t p = 0; p < ldb; ++p) pos_feature[p] = -1;
  const int W = (f + 3) / 4;
  int next = 0;
  for (int k = 0; k * 32 < W; ++k) {
    const int wk = std::min(32, W - 32 * k);
    const int take = std::min(f - next, 4 * wk);
    for (int r = 0; r < take; ++r) {
      const int word = r % wk, j = r / wk;
      pos_feature[(32 * k + word) * 4 + j] = order[next + r];
    }
    next += take;
  }
}

// builds the shared-memory tables for a given column placement; false when it does not fit
inline bool build_cf_tables(const int32_t* bin_ptr, const int32_t* pos_feature, int f, int64_t ldb, int budget_bytes,
                            CfHostTables* out) {
  const int W = (f + 3) / 4;
  if ((int64_t)W * 4 > ldb) return false;
  const int nb = (W + 31) / 32;
  if (nb > kCfMaxBlocks) return false;
  CfLayout& L = out->lay;
  memset(&L, 0, sizeof L);
  L.n_blocks = nb;
  int bank_rows[32] = {0};
  struct Rec { int k, l, j, rowbase, feat; };
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
        const int feat = pos_feature[word * 4 + j];
        if (feat < 0) continue;
        if (feat >= f) return false;
        const int nbins = bin_ptr[feat + 1] - bin_ptr[feat];
        L.a0[k][l][j] = bank_rows[l] * 32 + l;
        L.nbj[k][l][j] = nbins;
        recs.push_back(Rec{k, l, j, bank_rows[l], feat});
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
  for (const Rec& r : recs) {
    const int wk = L.wk[r.k];
    const int nbins = bin_ptr[r.feat + 1] - bin_ptr[r.feat];
    if (r.l < wk) {
      for (int b = 0; b < nbins; ++b) out->slot_of[(size_t)(r.rowbase + b) * 32 + r.l] = bin_ptr[r.feat] + b;
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
  // every feature must have been placed exactly once
  std::vector<int> seen(f, 0);
  for (int64_t p = 0; p < (int64_t)W * 4; ++p)
    if (pos_feature[p] >= 0) seen[pos_feature[p]]++;
  for (int i = 0; i < f; ++i)
    if (seen[i] != 1) return false;
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

// one byte of a bin word: clamp the code to the feature's bin count (255 = missing lands in the
// row after the last bin, which the reduce kernel ignores) and add (g, h) in the lane's own bank.
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

// 1024 threads = 32 warps; warp w takes row groups w, w + 32, ... of the CTA's slice, one column
// block after the other (each block is its own contiguous array: block-major storage).
// Partials are written in shared-memory order ([2][R*32] int32), un-permuted by the reduce kernel.
template <int UNROLL, bool IDENTITY>
__global__ void __launch_bounds__(1024, 1)
hist_accum_cf_kernel(const uint8_t* __restrict__ bins, int64_t n_rows, int last_w, const int2* __restrict__ gpair,
                     const uint32_t* __restrict__ row_idx, const HistPlan* __restrict__ plan,
                     const CfLayout* __restrict__ lay, const int32_t* __restrict__ fold_src,
                     const int32_t* __restrict__ fold_dst, int32_t* __restrict__ partials) {
  CB_LEVEL_TRACE_SCOPE(5);
  extern __shared__ int32_t smem[];
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
  int pid = plan->cta_pbase[blockIdx.x];
  for (int c0 = 0; c0 < rows; c0 += kChunkRows, ++pid) {
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
      for (int g0 = warp; g0 * ck < crow; g0 += 32 * UNROLL) {
        uint32_t wv[UNROLL];
        int2 gp[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const int i = (g0 + 32 * u) * ck + rslot;
          const int pos = pos0 + c0 + min(i, ilast);                 // clamped: always a valid row, no branch
          const uint32_t r = IDENTITY ? (uint32_t)pos : __ldg(row_idx + pos);
          wv[u] = ldg_nc_u32(reinterpret_cast<const uint32_t*>(bbase + (size_t)r * bw));
          gp[u] = __ldg(gpair + r);
          if (!active || i > ilast) gp[u] = make_int2(0, 0);         // adds zero: harmless
        }
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
    int32_t* pp = partials + (size_t)pid * 2 * RW;
    for (int s = threadIdx.x; s < RW; s += blockDim.x) {
      pp[s] = sG[s]; pp[RW + s] = sH[s];
      sG[s] = 0; sH[s] = 0;
    }
    __syncthreads();
  }
}

// level buffer [slot][S][2] int64 (+)= sum over a z-slice of the node's int32 partials, un-permuted
// through slot_of.  Each thread owns 4 consecutive shared-memory-order words of G and of H (16-byte
// loads, 4 partials in flight).  With gridDim.z > 1 the z-slices add atomically into a zeroed buffer;
// with gridDim.z == 1 the result is stored directly.
__global__ void __launch_bounds__(256) hist_reduce_cf_kernel(const int32_t* __restrict__ partials,
                                                             const HistPlan* __restrict__ plan, int R, int S,
                                                             const int32_t* __restrict__ slot_of,
                                                             int64_t* __restrict__ level_buf) {
  CB_LEVEL_TRACE_SCOPE(4);
  const int j = blockIdx.y;
  if (j >= plan->n_build) return;
  const int RW = R * 32;
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (v >= RW) return;
  const int p0 = plan->b_pbeg[j], p1 = plan->b_pbeg[j + 1];
  const int np = p1 - p0, nz = gridDim.z;
  const int per = (np + nz - 1) / nz;
  const int pa = p0 + blockIdx.z * per, pb = min(p1, pa + per);
  long long g[4] = {0, 0, 0, 0}, h[4] = {0, 0, 0, 0};
  const int32_t* pp = partials + (size_t)pa * 2 * RW + v;
  int p = pa;
  for (; p + 4 <= pb; p += 4, pp += (size_t)8 * RW) {
    int4 a[4], c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      a[u] = __ldg(reinterpret_cast<const int4*>(pp + (size_t)u * 2 * RW));
      c[u] = __ldg(reinterpret_cast<const int4*>(pp + (size_t)u * 2 * RW + RW));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      g[0] += a[u].x; g[1] += a[u].y; g[2] += a[u].z; g[3] += a[u].w;
      h[0] += c[u].x; h[1] += c[u].y; h[2] += c[u].z; h[3] += c[u].w;
    }
  }
  for (; p < pb; ++p, pp += (size_t)2 * RW) {
    const int4 a = __ldg(reinterpret_cast<const int4*>(pp));
    const int4 c = __ldg(reinterpret_cast<const int4*>(pp + RW));
    g[0] += a.x; g[1] += a.y; g[2] += a.z; g[3] += a.w;
    h[0] += c.x; h[1] += c.y; h[2] += c.z; h[3] += c.w;
  }
  const int4 so = __ldg(reinterpret_cast<const int4*>(slot_of + v));
  const int sl[4] = {so.x, so.y, so.z, so.w};
  long long* out = reinterpret_cast<long long*>(level_buf) + (size_t)j * 2 * S;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (sl[i] < 0) continue;
    if (nz == 1) {
      reinterpret_cast<longlong2*>(out)[sl[i]] = make_longlong2(g[i], h[i]);
    } else {
      if (g[i]) atomicAdd(reinterpret_cast<unsigned long long*>(out) + 2 * (size_t)sl[i], (unsigned long long)g[i]);
      if (h[i]) atomicAdd(reinterpret_cast<unsigned long long*>(out) + 2 * (size_t)sl[i] + 1, (unsigned long long)h[i]);
    }
  }
}

__global__ void zero_i64_kernel(int64_t* p, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = 0;
}

}  // namespace cb
