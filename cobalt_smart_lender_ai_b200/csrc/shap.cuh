// K10  path-dependent TreeSHAP (Lundberg, Erion & Lee 2018, Algorithm 2): the per-feature attributions behind
// shap.TreeExplainer(xgb_model).shap_values(row) in the reference's /predict (src/api/cobalt_fast_api.py:46,
// :100-105).  Node cover = sum_hess.  One thread per (row, tree) fills that tree's phi_t[f+1] slice of the
// workspace; a second kernel adds the slices in tree order, so the f32 result does not depend on scheduling and
// equals the oracle's (R13) bit for bit: same operations, same order, one rounding each (-fmad=false).
#pragma once
#include "common.cuh"

namespace cb {

constexpr int kShapMaxDepth = 12;
constexpr int kShapPathLen = (kShapMaxDepth + 2) * (kShapMaxDepth + 3) / 2;

struct ShapPath { int feat; float zero, one, pw; };
struct ShapFrame { int nid, ud, poff, feat; float zero, one; };

struct ShapForest {                        // wire-format arrays, concatenated over trees
  const int64_t* tree_off;
  const int32_t* left; const int32_t* right; const int32_t* split_index;
  const float* cond; const uint8_t* default_left; const float* cover;
};

__device__ __forceinline__ void shap_extend(ShapPath* p, int ud, float zero, float one, int feat) {
  p[ud].feat = feat; p[ud].zero = zero; p[ud].one = one; p[ud].pw = ud == 0 ? 1.0f : 0.0f;
  for (int i = ud - 1; i >= 0; --i) {
    p[i + 1].pw += one * p[i].pw * (float)(i + 1) / (float)(ud + 1);
    p[i].pw = zero * p[i].pw * (float)(ud - i) / (float)(ud + 1);
  }
}
__device__ __forceinline__ void shap_unwind(ShapPath* p, int ud, int k) {
  const float one = p[k].one, zero = p[k].zero;
  float next = p[ud].pw;
  for (int i = ud - 1; i >= 0; --i) {
    if (one != 0.0f) {
      const float tmp = p[i].pw;
      p[i].pw = next * (float)(ud + 1) / ((float)(i + 1) * one);
      next = tmp - p[i].pw * zero * (float)(ud - i) / (float)(ud + 1);
    } else {
      p[i].pw = p[i].pw * (float)(ud + 1) / (zero * (float)(ud - i));
    }
  }
  for (int i = k; i < ud; ++i) { p[i].feat = p[i + 1].feat; p[i].zero = p[i + 1].zero; p[i].one = p[i + 1].one; }
}
__device__ __forceinline__ float shap_unwound_sum(const ShapPath* p, int ud, int k) {
  const float one = p[k].one, zero = p[k].zero;
  float next = p[ud].pw, total = 0.0f;
  for (int i = ud - 1; i >= 0; --i) {
    if (one != 0.0f) {
      const float tmp = next * (float)(ud + 1) / ((float)(i + 1) * one);
      total += tmp;
      next = p[i].pw - tmp * zero * ((float)(ud - i) / (float)(ud + 1));
    } else {
      total += p[i].pw / zero / ((float)(ud - i) / (float)(ud + 1));
    }
  }
  return total;
}

// cover-weighted mean of the leaves of every tree: one thread per tree, nodes in reverse id order
// (children always carry larger ids than their parent, checked by the host)
__global__ void shap_means_kernel(ShapForest fo, int n_trees, float* __restrict__ node_mean, float* __restrict__ tree_mean) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_trees) return;
  const int64_t o = fo.tree_off[t], e = fo.tree_off[t + 1];
  for (int64_t i = e - 1; i >= o; --i) {
    if (fo.left[i] == -1) {
      node_mean[i] = fo.cond[i];
    } else {
      const int64_t l = o + fo.left[i], r = o + fo.right[i];
      node_mean[i] = (node_mean[l] * fo.cover[l] + node_mean[r] * fo.cover[r]) / fo.cover[i];
    }
  }
  tree_mean[t] = node_mean[o];
}

// phi_t of tree t for row r -> ws[(r * n_trees + t) * (f + 1) ...] (zeroed by the caller); hot child first
__global__ void __launch_bounds__(128) shap_trees_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx, ShapForest fo,
                                                         int n_trees, float* __restrict__ ws) {
  const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= n * n_trees) return;
  const int64_t r = item / n_trees;
  const int t = (int)(item - r * n_trees);
  const float* x = X + r * ldx;
  float* phi = ws + item * (f + 1);
  const int64_t o = fo.tree_off[t];
  ShapPath path[kShapPathLen];
  ShapFrame st[2 * kShapMaxDepth + 4];
  int sp = 0;
  st[sp++] = ShapFrame{0, 0, 0, -1, 1.0f, 1.0f};
  while (sp > 0) {
    const ShapFrame fr = st[--sp];
    ShapPath* p = path + fr.poff + fr.ud + (fr.ud > 0 ? 1 : 0);   // own buffer: right behind the parent's elements
    for (int i = 0; i < fr.ud; ++i) p[i] = path[fr.poff + i];
    int ud = fr.ud;
    shap_extend(p, ud, fr.zero, fr.one, fr.feat);
    const int nid = fr.nid;
    if (fo.left[o + nid] == -1) {
      const float leaf = fo.cond[o + nid];
      for (int i = 1; i <= ud; ++i) {
        const float w = shap_unwound_sum(p, ud, i);
        phi[p[i].feat] += w * (p[i].one - p[i].zero) * leaf;
      }
      continue;
    }
    const int sf = fo.split_index[o + nid];
    const float v = x[sf];
    const bool go_left = (v != v) ? fo.default_left[o + nid] != 0 : (v < fo.cond[o + nid]);
    const int hot = go_left ? fo.left[o + nid] : fo.right[o + nid];
    const int cold = go_left ? fo.right[o + nid] : fo.left[o + nid];
    const float w = fo.cover[o + nid];
    const float hot_zero = fo.cover[o + hot] / w, cold_zero = fo.cover[o + cold] / w;
    float in_zero = 1.0f, in_one = 1.0f;
    int k = 0;
    for (; k <= ud; ++k)
      if (p[k].feat == sf) break;
    if (k != ud + 1) {
      in_zero = p[k].zero; in_one = p[k].one;
      shap_unwind(p, ud, k);
      ud -= 1;
    }
    const int poff = (int)(p - path);
    st[sp++] = ShapFrame{cold, ud + 1, poff, sf, cold_zero * in_zero, 0.0f};
    st[sp++] = ShapFrame{hot, ud + 1, poff, sf, hot_zero * in_zero, in_one};
  }
}

// out[r][j] = sum over trees, in tree order, of phi_t[j]; column f = sum of the trees' means + base_margin
__global__ void shap_sum_kernel(const float* __restrict__ ws, const float* __restrict__ tree_mean, int64_t n, int f, int n_trees,
                                float base_margin, float* __restrict__ out) {
  const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= n * (f + 1)) return;
  const int64_t r = item / (f + 1);
  const int j = (int)(item - r * (f + 1));
  float s = 0.0f;
  if (j < f) {
    const float* p = ws + r * n_trees * (f + 1) + j;
    for (int t = 0; t < n_trees; ++t) s += p[(int64_t)t * (f + 1)];
  } else {
    for (int t = 0; t < n_trees; ++t) s += tree_mean[t];
    s += base_margin;
  }
  out[item] = s;
}

}  // namespace cb
