/*
 * TEST INFRASTRUCTURE -- CPU oracle for the histogram-GBDT hot path.  NOT PRODUCT CODE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this file's shared object.  The product (cobalt_smart_lender_ai_b200/) never
 * imports, links or executes anything under oracle/.
 *
 * What it restates.  The reference delegates this path to the third-party package
 * `xgboost` (requirements.txt:11, unpinned; the shipped model says 3.0.0, tree_method
 * hist, max_bin 256, depthwise -- src/api/models/xgb_model_tree.pkl).  xgboost is neither
 * vendored under /root/reference nor installable offline, so this file restates its
 * published `hist` algorithm (rules R1..R12 of SURVEY.md 8c / DESIGN.md) and is anchored
 * on the reference's call sites:
 *   fit            src/model_train_test/model_tree_train_test.py:111-118,132-159
 *   predict(_proba) src/model_train_test/model_tree_train_test.py:171-172,
 *                  src/api/cobalt_fast_api.py:90-91
 *   importance     src/model_train_test/model_tree_train_test.py:197, src/api/cobalt_fast_api.py:135-136
 * and on the shipped model file (gain / leaf / hessian-additivity known answers and the
 * traversal rule), see tests/test_oracle_golden.py.
 *
 * PARITY PIN STATUS: the reference holds no test for this path.  The gain formula, the
 * leaf formula, the tree wire format and the traversal rule are pinned by the shipped
 * model (tests/golden/fixture_forest.npz).  Training end to end (cuts, split choice on
 * data) is "parity unpinned": there is no runnable xgboost here to produce expected splits.
 *
 * Declared deviations from upstream xgboost (all mirrored by the CUDA path):
 *   - R1: exact-rank quantiles replace the approximate mergeable sketch.
 *   - R2: at most 255 real bins per feature (u8 code 255 = missing).
 *   - R4/R5: gradients are rounded to fixed point (2^-sg, 2^-sh) and histogram sums are
 *     exact int64, as north_star asks, instead of upstream's f64 sums of f32.
 *   - sigmoid uses the fixed operation sequence below (bit-reproducible on CPU and GPU).
 *   - subsample / colsample use their own counter-based hash (statistical parity only).
 *
 * Threads (it is also bench.py's timed CPU baseline): cuts, binning, gradients, histogram build (per-thread private
 * histograms kept between calls, slot-sliced merge), the split scan (contiguous feature ranges, combined in index order),
 * the stable partition of large segments (chunk counts, prefix, write), totals and leaf updates all use OpenMP; results
 * do not depend on the thread count (integer sums; same '>' tie-breaks in the same order).
 *
 * Build: see oracle/Makefile (gcc -O2 -fopenmp -ffp-contract=off; no -ffast-math).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MISSING 255
#define ORC_MAX_CUTS 255 /* real bins per feature, R2 */

int orc_version(void) { return 1; }

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ------------------------------------------------------------------ R1: cuts ------- */
static int cmp_f32(const void* a, const void* b) {
  float x = *(const float*)a, y = *(const float*)b;
  return (x > y) - (x < y);
}

/* One feature.  v: scratch of n floats.  Writes <=255 cuts, returns their count. */
static int cuts_one_feature(const float* X, int64_t n, int64_t ldx, int f, int max_bin, float* v,
                            float* cuts, float* min_val) {
  int64_t m = 0;
  for (int64_t r = 0; r < n; ++r) {
    float x = X[r * ldx + f];
    if (x != x) continue;   /* NaN = missing */
    v[m++] = x + 0.0f;      /* -0 -> +0 */
  }
  int cap = max_bin < ORC_MAX_CUTS ? max_bin : ORC_MAX_CUTS;
  if (cap < 2) cap = 2;
  if (m == 0) {             /* upstream's empty-column constants */
    float c = 1e-5f;
    *min_val = c - (fabsf(c) + 1e-5f);
    cuts[0] = c + (fabsf(c) + 1e-5f);
    return 1;
  }
  qsort(v, (size_t)m, sizeof(float), cmp_f32);
  float vmin = v[0], vmax = v[m - 1];
  *min_val = vmin - (fabsf(vmin) + 1e-5f);
  float sentinel = vmax + (fabsf(vmax) + 1e-5f);
  /* distinct count, stop once above cap */
  int64_t distinct = 1;
  for (int64_t i = 1; i < m && distinct <= cap; ++i) distinct += (v[i] != v[i - 1]);
  int nc = 0;
  if (distinct <= cap) {
    for (int64_t i = 1; i < m; ++i)
      if (v[i] != v[i - 1]) cuts[nc++] = v[i];
  } else {
    float last = vmin;
    for (int k = 1; k < cap; ++k) {
      int64_t rank = (int64_t)(((__int128)k * (__int128)m) / cap);
      float c = v[rank];
      if (c > last) { cuts[nc++] = c; last = c; }
    }
  }
  cuts[nc++] = sentinel;
  return nc;
}

/* cuts: [f][256] (padded), n_cuts[f], min_vals[f] */
int orc_cuts(const float* X, int64_t n, int f, int64_t ldx, int max_bin, float* cuts, int* n_cuts,
             float* min_vals) {
#pragma omp parallel
  {
    float* v = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
#pragma omp for schedule(dynamic, 1)
    for (int j = 0; j < f; ++j) {
      for (int k = 0; k < 256; ++k) cuts[(size_t)j * 256 + k] = INFINITY;
      n_cuts[j] = cuts_one_feature(X, n, ldx, j, max_bin, v, cuts + (size_t)j * 256, min_vals + j);
    }
    free(v);
  }
  return 0;
}

/* ------------------------------------------------------------------ R2: bins ------- */
static inline uint8_t bin_of(float x, const float* c, int nc) {
  if (x != x) return ORC_MISSING;
  int lo = 0, hi = nc; /* upper_bound: first index with c[idx] > x */
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (c[mid] <= x) lo = mid + 1; else hi = mid;
  }
  if (lo == nc) lo = nc - 1;
  return (uint8_t)lo;
}

int orc_bin(const float* X, int64_t n, int f, int64_t ldx, const float* cuts, const int* n_cuts,
            uint8_t* bins, int64_t ldb) {
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < n; ++r) {
    for (int j = 0; j < f; ++j)
      bins[r * ldb + j] = bin_of(X[r * ldx + j], cuts + (size_t)j * 256, n_cuts[j]);
    for (int64_t j = f; j < ldb; ++j) bins[r * ldb + j] = ORC_MISSING;
  }
  return 0;
}

/* ------------------------------------------------------------------ R4: gradients -- */
/* exp on [-87, 88.7]: n = rint(t*log2e); r = t - n*ln2_hi - n*ln2_lo; degree-6 Horner in
 * separately rounded f32 mul/add; scale by 2^n through the exponent bits (two steps so
 * that n = 128 stays finite). */
static inline float orc_exp(float t) {
  float nf = rintf(t * 1.44269504088896341f);
  float r = t - nf * 0.693145751953125f;
  r = r - nf * 1.42860682030941723212e-6f;
  float p = 1.0f / 720.0f;
  p = p * r + 1.0f / 120.0f;
  p = p * r + 1.0f / 24.0f;
  p = p * r + 1.0f / 6.0f;
  p = p * r + 0.5f;
  p = p * r + 1.0f;
  p = p * r + 1.0f;
  int n = (int)nf;
  int n1 = n / 2, n2 = n - n1;
  union { uint32_t u; float f; } a, b;
  a.u = (uint32_t)(n1 + 127) << 23;
  b.u = (uint32_t)(n2 + 127) << 23;
  return (p * a.f) * b.f;
}

static inline float orc_sigmoid1(float x) {
  float t = -x;
  if (t > 88.7f) t = 88.7f;
  if (t < -87.0f) t = -87.0f;
  float d = orc_exp(t) + 1.0f;
  return 1.0f / d;
}

void orc_sigmoid(const float* x, int64_t n, float* out) {
  for (int64_t i = 0; i < n; ++i) out[i] = orc_sigmoid1(x[i]);
}

/* counter-based row sampler: keep row iff hash < subsample * 2^32 */
static inline uint32_t orc_hash3(uint32_t seed, uint32_t a, uint64_t b) {
  uint32_t h = seed ^ 0x9E3779B9u;
  h ^= a * 0x85EBCA6Bu; h = (h << 13) | (h >> 19); h = h * 5u + 0xE6546B64u;
  h ^= (uint32_t)(b & 0xFFFFFFFFu) * 0xCC9E2D51u; h = (h << 13) | (h >> 19); h = h * 5u + 0xE6546B64u;
  h ^= (uint32_t)(b >> 32) * 0x1B873593u;
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}

uint32_t orc_hash(uint32_t seed, uint32_t a, uint64_t b) { return orc_hash3(seed, a, b); }

/* q: [n][2] int32 (g, h).  w may be NULL.  keep_thresh = subsample>=1 ? 0 : floor(subsample*2^32) */
int orc_grad(const float* margin, const uint8_t* y, const float* w, int64_t n, float spw, int sg,
             int sh, uint32_t keep_thresh, uint32_t seed, uint32_t tree, int64_t row_offset,
             int32_t* q) {
  float fg = ldexpf(1.0f, sg), fh = ldexpf(1.0f, sh);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    float p = orc_sigmoid1(margin[i]);
    float yy = (float)y[i];
    float wt = y[i] ? spw : 1.0f;
    if (w) wt = wt * w[i];
    float g = (p - yy) * wt;
    float hh = p * (1.0f - p);
    if (hh < 1e-16f) hh = 1e-16f;
    float h = hh * wt;
    int32_t qg = (int32_t)rintf(g * fg), qh = (int32_t)rintf(h * fh);
    if (keep_thresh && orc_hash3(seed, tree, (uint64_t)(row_offset + i)) >= keep_thresh) { qg = 0; qh = 0; }
    q[2 * i] = qg; q[2 * i + 1] = qh;
  }
  return 0;
}

/* ------------------------------------------------------------------ R5: histogram -- */
/* compact layout: slot = bin_ptr[f] + bin, hist[slot][2] int64; missing rows skipped */
static int64_t** g_hist_loc = NULL;   /* per-thread private histograms, kept between calls */
static int g_hist_loc_threads = 0;
static size_t g_hist_loc_slots = 0;

int orc_hist(const uint8_t* bins, int64_t ldb, int f, const int32_t* q, const uint32_t* row_idx,
             int64_t count, const int* bin_ptr, int64_t* hist) {
  int S = bin_ptr[f];
  int nt = omp_get_max_threads();
  if (count < 8192 || nt == 1) {                 /* small node: one thread, straight into the output */
    memset(hist, 0, sizeof(int64_t) * 2 * (size_t)S);
    for (int64_t i = 0; i < count; ++i) {
      int64_t r = row_idx ? (int64_t)row_idx[i] : i;
      int64_t g = q[2 * r], h = q[2 * r + 1];
      const uint8_t* b = bins + r * ldb;
      for (int j = 0; j < f; ++j) {
        if (b[j] == ORC_MISSING) continue;
        int s = bin_ptr[j] + b[j];
        hist[2 * s] += g; hist[2 * s + 1] += h;
      }
    }
    return 0;
  }
  if (nt > g_hist_loc_threads || (size_t)S > g_hist_loc_slots) {
    for (int k = 0; k < g_hist_loc_threads; ++k) free(g_hist_loc[k]);
    free(g_hist_loc);
    g_hist_loc = (int64_t**)malloc(sizeof(int64_t*) * (size_t)nt);
    if (!g_hist_loc) { g_hist_loc_threads = 0; return -1; }
    for (int k = 0; k < nt; ++k) g_hist_loc[k] = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)S);
    g_hist_loc_threads = nt; g_hist_loc_slots = (size_t)S;
  }
  int used = 0;
#pragma omp parallel
  {
    int tid = omp_get_thread_num();
#pragma omp single
    used = omp_get_num_threads();
    int64_t* loc = g_hist_loc[tid];
    memset(loc, 0, sizeof(int64_t) * 2 * (size_t)S);
#pragma omp for schedule(static)
    for (int64_t i = 0; i < count; ++i) {
      int64_t r = row_idx ? (int64_t)row_idx[i] : i;
      int64_t g = q[2 * r], h = q[2 * r + 1];
      const uint8_t* b = bins + r * ldb;
      for (int j = 0; j < f; ++j) {
        if (b[j] == ORC_MISSING) continue;
        int s = bin_ptr[j] + b[j];
        loc[2 * s] += g; loc[2 * s + 1] += h;
      }
    }
    /* merge: every thread sums a slice of the slots over all private copies (integer sums: order free) */
#pragma omp for schedule(static)
    for (int s = 0; s < 2 * S; ++s) {
      int64_t a = 0;
      for (int k = 0; k < used; ++k) a += g_hist_loc[k][s];
      hist[s] = a;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ R6/R7: split --- */
typedef struct {
  float eta, lambda, alpha, gamma, mcw;
  int max_depth, sg, sh;
} orc_tparams;

typedef struct {
  float loss_chg;    /* 0 when no candidate */
  int feature;       /* -1 when none */
  int split_bin;     /* rows with bin <= split_bin go left; -1 possible (backward at bin 0) */
  int default_left;
  float threshold;
  int64_t GL, HL, GR, HR;
} orc_split;

static inline double thr_l1(double g, double a) {
  if (g > a) return g - a;
  if (g < -a) return g + a;
  return 0.0;
}
static inline double calc_gain(double G, double H, const orc_tparams* p) {
  if (H < (double)p->mcw || H <= 0.0) return 0.0;
  double t = p->alpha == 0.0f ? G : thr_l1(G, (double)p->alpha);
  return (t * t) / (H + (double)p->lambda);
}
static inline double calc_weight(double G, double H, const orc_tparams* p) {
  if (H < (double)p->mcw || H <= 0.0) return 0.0;
  double t = p->alpha == 0.0f ? G : thr_l1(G, (double)p->alpha);
  return -t / (H + (double)p->lambda);
}

double orc_calc_gain(double G, double H, float lambda, float alpha, float mcw) {
  orc_tparams p; memset(&p, 0, sizeof p); p.lambda = lambda; p.alpha = alpha; p.mcw = mcw;
  return calc_gain(G, H, &p);
}
double orc_calc_weight(double G, double H, float lambda, float alpha, float mcw) {
  orc_tparams p; memset(&p, 0, sizeof p); p.lambda = lambda; p.alpha = alpha; p.mcw = mcw;
  return calc_weight(G, H, &p);
}

/* features [j0, j1) of one node, in index order; *pbest must come in initialised (feature -1, loss_chg 0) */
static void scan_range(const int64_t* hist, const int* bin_ptr, int j0, int j1, const float* cuts, const float* min_vals,
                       int64_t Gt, int64_t Ht, const orc_tparams* p, const uint8_t* feature_mask, double ig, double ih,
                       float root_gain, orc_split* pbest) {
  orc_split best = *pbest;
  for (int j = j0; j < j1; ++j) {
    if (feature_mask && !feature_mask[j]) continue;
    int nb = bin_ptr[j + 1] - bin_ptr[j];
    const int64_t* hj = hist + 2 * (size_t)bin_ptr[j];
    const float* cj = cuts + (size_t)j * 256;
    int64_t gl = 0, hl = 0;
    for (int i = 0; i < nb; ++i) {             /* forward: missing -> right */
      gl += hj[2 * i]; hl += hj[2 * i + 1];
      int64_t gr = Gt - gl, hr = Ht - hl;
      double HLd = (double)hl * ih, HRd = (double)hr * ih;
      if (HLd >= (double)p->mcw && HRd >= (double)p->mcw) {
        double gain = calc_gain((double)gl * ig, HLd, p) + calc_gain((double)gr * ig, HRd, p);
        float chg = (float)(gain - (double)root_gain);
        if (!isinf(chg) && chg > best.loss_chg) {
          best.loss_chg = chg; best.feature = j; best.split_bin = i; best.default_left = 0;
          best.threshold = cj[i]; best.GL = gl; best.HL = hl; best.GR = gr; best.HR = hr;
        }
      }
    }
    if (gl != Gt || hl != Ht) {                /* node has missing in j: backward, missing -> left */
      int64_t gr = 0, hr = 0;
      for (int i = nb - 1; i >= 0; --i) {
        gr += hj[2 * i]; hr += hj[2 * i + 1];
        int64_t gl2 = Gt - gr, hl2 = Ht - hr;
        double HLd = (double)hl2 * ih, HRd = (double)hr * ih;
        if (HRd >= (double)p->mcw && HLd >= (double)p->mcw) {
          double gain = calc_gain((double)gl2 * ig, HLd, p) + calc_gain((double)gr * ig, HRd, p);
          float chg = (float)(gain - (double)root_gain);
          if (!isinf(chg) && chg > best.loss_chg) {
            best.loss_chg = chg; best.feature = j; best.split_bin = i - 1; best.default_left = 1;
            best.threshold = i == 0 ? min_vals[j] : cj[i - 1];
            best.GL = gl2; best.HL = hl2; best.GR = gr; best.HR = hr;
          }
        }
      }
    }
  }
  *pbest = best;
}

/* hist: node's compact histogram; Gt/Ht: node totals; feature_mask may be NULL */
int orc_scan(const int64_t* hist, const int* bin_ptr, int f, const float* cuts, const float* min_vals,
             int64_t Gt, int64_t Ht, const orc_tparams* p, const uint8_t* feature_mask,
             orc_split* out) {
  double ig = ldexp(1.0, -p->sg), ih = ldexp(1.0, -p->sh);
  float root_gain = (float)calc_gain((double)Gt * ig, (double)Ht * ih, p);
  orc_split best; memset(&best, 0, sizeof best); best.feature = -1; best.loss_chg = 0.0f;
  int nt = omp_get_max_threads();
  if (nt > 64) nt = 64;
  if (f < 32 || nt == 1) {
    scan_range(hist, bin_ptr, 0, f, cuts, min_vals, Gt, Ht, p, feature_mask, ig, ih, root_gain, &best);
  } else {
    /* contiguous feature ranges per thread, combined in range order with the same strict '>' as the serial scan:
       the lowest feature index still wins ties */
    orc_split part[64];
#pragma omp parallel for schedule(static, 1) num_threads(nt)
    for (int k = 0; k < nt; ++k) {
      memset(&part[k], 0, sizeof part[k]); part[k].feature = -1; part[k].loss_chg = 0.0f;
      scan_range(hist, bin_ptr, (int)((long long)f * k / nt), (int)((long long)f * (k + 1) / nt), cuts, min_vals, Gt, Ht, p,
                 feature_mask, ig, ih, root_gain, &part[k]);
    }
    for (int k = 0; k < nt; ++k)
      if (part[k].feature >= 0 && part[k].loss_chg > best.loss_chg) best = part[k];
  }
  *out = best;
  return 0;
}

/* ------------------------------------------------------------------ R8: partition -- */
/* stable split of seg[0..count) into out (lefts first); returns number of lefts */
int64_t orc_partition(const uint8_t* bins, int64_t ldb, const uint32_t* seg, int64_t count,
                      int feature, int split_bin, int default_left, uint32_t* out) {
  int nt = omp_get_max_threads();
  if (count < 65536 || nt == 1) {
    int64_t nl = 0;
    for (int64_t i = 0; i < count; ++i) {
      uint8_t c = bins[(int64_t)seg[i] * ldb + feature];
      int left = c == ORC_MISSING ? default_left : ((int)c <= split_bin);
      nl += left;
    }
    int64_t a = 0, b = nl;
    for (int64_t i = 0; i < count; ++i) {
      uint8_t c = bins[(int64_t)seg[i] * ldb + feature];
      int left = c == ORC_MISSING ? default_left : ((int)c <= split_bin);
      if (left) out[a++] = seg[i]; else out[b++] = seg[i];
    }
    return nl;
  }
  /* large segment: contiguous chunk per thread, count, exclusive prefix over the chunks, write -- still stable */
  int64_t cnt[256];
  if (nt > 256) nt = 256;
  int64_t per = (count + nt - 1) / nt, nl = 0;
#pragma omp parallel for schedule(static, 1) num_threads(nt)
  for (int k = 0; k < nt; ++k) {
    int64_t b0 = k * per, b1 = b0 + per < count ? b0 + per : count, c0 = 0;
    for (int64_t i = b0; i < b1; ++i) {
      uint8_t c = bins[(int64_t)seg[i] * ldb + feature];
      c0 += c == ORC_MISSING ? default_left : ((int)c <= split_bin);
    }
    cnt[k] = c0;
  }
  for (int k = 0; k < nt; ++k) { int64_t c0 = cnt[k]; cnt[k] = nl; nl += c0; }
#pragma omp parallel for schedule(static, 1) num_threads(nt)
  for (int k = 0; k < nt; ++k) {
    int64_t b0 = k * per, b1 = b0 + per < count ? b0 + per : count;
    int64_t a = cnt[k], b = nl + (b0 < count ? b0 : count) - cnt[k];
    for (int64_t i = b0; i < b1; ++i) {
      uint8_t c = bins[(int64_t)seg[i] * ldb + feature];
      int left = c == ORC_MISSING ? default_left : ((int)c <= split_bin);
      if (left) out[a++] = seg[i]; else out[b++] = seg[i];
    }
  }
  return nl;
}

/* ------------------------------------------------------------------ training loop -- */
/* Forest SoA, xgboost wire format (a-9): per tree stride max_nodes = 2^(max_depth+1)-1. */
typedef struct {
  int32_t* left; int32_t* right; int32_t* parent; int32_t* split_index; int32_t* split_bin;
  float* split_cond; uint8_t* default_left; float* base_weight; float* loss_chg; float* sum_hess;
  int32_t* n_nodes;
} orc_forest;

typedef struct { int nid; int64_t begin, count; int64_t G, H; int depth; } orc_node;

/*
 * Trains n_trees rounds on pre-binned data.  margin (in/out) [n].  feature_masks: NULL or
 * [n_trees][f].  row_order_out: NULL or [n] final row_idx of the LAST tree (partition parity).
 * hist_dump: NULL or [S][2] root histogram of the LAST tree.
 */
int orc_train(const uint8_t* bins, int64_t ldb, int64_t n, int f, const int* bin_ptr,
              const float* cuts, const float* min_vals, const uint8_t* y, const float* w, float spw,
              const orc_tparams* p, int n_trees, int tree_offset, uint32_t keep_thresh, uint32_t seed,
              const uint8_t* feature_masks, float* margin, orc_forest* fo, uint32_t* row_order_out,
              int64_t* hist_dump) {
  int S = bin_ptr[f];
  int max_nodes = (1 << (p->max_depth + 1)) - 1;
  int32_t* q = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)n);
  uint32_t* ridx = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
  uint32_t* tmp = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
  int64_t* hists = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)S * (size_t)max_nodes);
  orc_node* cur = (orc_node*)malloc(sizeof(orc_node) * (size_t)max_nodes);
  orc_node* nxt = (orc_node*)malloc(sizeof(orc_node) * (size_t)max_nodes);
  float* leaf_of_node = (float*)malloc(sizeof(float) * (size_t)max_nodes);
  double ig = ldexp(1.0, -p->sg), ih = ldexp(1.0, -p->sh);
  if (!q || !ridx || !tmp || !hists || !cur || !nxt || !leaf_of_node) return -1;

  for (int t = 0; t < n_trees; ++t) {
    size_t base = (size_t)t * (size_t)max_nodes;
    const uint8_t* fmask = feature_masks ? feature_masks + (size_t)t * f : NULL;
    orc_grad(margin, y, w, n, spw, p->sg, p->sh, keep_thresh, seed, (uint32_t)(tree_offset + t), 0, q);
    int64_t G = 0, H = 0;
#pragma omp parallel for schedule(static) reduction(+ : G, H)
    for (int64_t i = 0; i < n; ++i) { G += q[2 * i]; H += q[2 * i + 1]; ridx[i] = (uint32_t)i; }
    for (int k = 0; k < max_nodes; ++k) {
      fo->left[base + k] = -1; fo->right[base + k] = -1; fo->parent[base + k] = 2147483647;
      fo->split_index[base + k] = 0; fo->split_bin[base + k] = 0; fo->split_cond[base + k] = 0.0f;
      fo->default_left[base + k] = 0; fo->base_weight[base + k] = 0.0f; fo->loss_chg[base + k] = 0.0f;
      fo->sum_hess[base + k] = 0.0f;
    }
    int n_nodes = 1, n_cur = 1;
    cur[0].nid = 0; cur[0].begin = 0; cur[0].count = n; cur[0].G = G; cur[0].H = H; cur[0].depth = 0;
    fo->base_weight[base] = (float)calc_weight((double)G * ig, (double)H * ih, p);
    fo->sum_hess[base] = (float)((double)H * ih);
    orc_hist(bins, ldb, f, q, NULL, n, bin_ptr, hists); /* node 0 */
    if (hist_dump && t == n_trees - 1) memcpy(hist_dump, hists, sizeof(int64_t) * 2 * (size_t)S);

    for (int depth = 0; n_cur > 0; ++depth) {
      int n_nxt = 0;
      for (int c = 0; c < n_cur; ++c) {
        orc_node nd = cur[c];
        int64_t* hn = hists + 2 * (size_t)S * (size_t)nd.nid;
        orc_split sp; memset(&sp, 0, sizeof sp); sp.feature = -1;
        if (depth < p->max_depth) orc_scan(hn, bin_ptr, f, cuts, min_vals, nd.G, nd.H, p, fmask, &sp);
        int valid = sp.feature >= 0 && sp.loss_chg > 1e-6f && !(sp.loss_chg < p->gamma) &&
                    depth < p->max_depth && sp.HL != 0 && sp.HR != 0;
        if (!valid) {
          float wgt = fo->base_weight[base + nd.nid];
          float leaf = wgt * p->eta;
          fo->split_cond[base + nd.nid] = leaf;
          fo->base_weight[base + nd.nid] = leaf;
          leaf_of_node[nd.nid] = leaf;
#pragma omp parallel for schedule(static) if (nd.count > 32768)
          for (int64_t i = 0; i < nd.count; ++i) margin[ridx[nd.begin + i]] += leaf;   /* rows are distinct: no race */
          continue;
        }
        int L = n_nodes, R = n_nodes + 1; n_nodes += 2;
        fo->left[base + nd.nid] = L; fo->right[base + nd.nid] = R;
        fo->parent[base + L] = nd.nid; fo->parent[base + R] = nd.nid;
        fo->split_index[base + nd.nid] = sp.feature; fo->split_bin[base + nd.nid] = sp.split_bin;
        fo->split_cond[base + nd.nid] = sp.threshold; fo->default_left[base + nd.nid] = (uint8_t)sp.default_left;
        fo->loss_chg[base + nd.nid] = sp.loss_chg;
        fo->base_weight[base + L] = (float)calc_weight((double)sp.GL * ig, (double)sp.HL * ih, p);
        fo->base_weight[base + R] = (float)calc_weight((double)sp.GR * ig, (double)sp.HR * ih, p);
        fo->sum_hess[base + L] = (float)((double)sp.HL * ih);
        fo->sum_hess[base + R] = (float)((double)sp.HR * ih);
        int64_t nl = orc_partition(bins, ldb, ridx + nd.begin, nd.count, sp.feature, sp.split_bin,
                                   sp.default_left, tmp + nd.begin);
        memcpy(ridx + nd.begin, tmp + nd.begin, sizeof(uint32_t) * (size_t)nd.count);
        /* children histograms: build the one with fewer rows, derive the sibling (R5/K9) */
        int64_t* hl = hists + 2 * (size_t)S * (size_t)L;
        int64_t* hr = hists + 2 * (size_t)S * (size_t)R;
        if (depth + 1 < p->max_depth) {
          if (nl <= nd.count - nl) {
            orc_hist(bins, ldb, f, q, ridx + nd.begin, nl, bin_ptr, hl);
            for (int s = 0; s < 2 * S; ++s) hr[s] = hn[s] - hl[s];
          } else {
            orc_hist(bins, ldb, f, q, ridx + nd.begin + nl, nd.count - nl, bin_ptr, hr);
            for (int s = 0; s < 2 * S; ++s) hl[s] = hn[s] - hr[s];
          }
        }
        orc_node a = {L, nd.begin, nl, sp.GL, sp.HL, depth + 1};
        orc_node b = {R, nd.begin + nl, nd.count - nl, sp.GR, sp.HR, depth + 1};
        nxt[n_nxt++] = a; nxt[n_nxt++] = b;
      }
      orc_node* sw = cur; cur = nxt; nxt = sw; n_cur = n_nxt;
    }
    fo->n_nodes[t] = n_nodes;
    if (row_order_out && t == n_trees - 1) memcpy(row_order_out, ridx, sizeof(uint32_t) * (size_t)n);
  }
  free(q); free(ridx); free(tmp); free(hists); free(cur); free(nxt); free(leaf_of_node);
  return 0;
}

/* ------------------------------------------------------------------ R11: predict --- */
/* forest arrays concatenated with tree_off[t] node offsets (any tree shape). out: margin */
int orc_predict_margin(const float* X, int64_t n, int f, int64_t ldx, const int64_t* tree_off,
                       int n_trees, const int32_t* left, const int32_t* right,
                       const int32_t* split_index, const float* split_cond,
                       const uint8_t* default_left, float base_margin, float* out) {
  (void)f;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < n; ++r) {
    const float* x = X + r * ldx;
    float m = base_margin;
    for (int t = 0; t < n_trees; ++t) {
      int64_t o = tree_off[t];
      int nid = 0;
      while (left[o + nid] != -1) {
        float v = x[split_index[o + nid]];
        int go_left = (v != v) ? default_left[o + nid] : (v < split_cond[o + nid]);
        nid = go_left ? left[o + nid] : right[o + nid];
      }
      m += split_cond[o + nid];
    }
    out[r] = m;
  }
  return 0;
}

/* ------------------------------------------------------------- R13: TreeSHAP (8f-3) ---
 * Path-dependent TreeSHAP, Lundberg, Erion & Lee 2018 ("Consistent individualized feature
 * attribution for tree ensembles", Algorithm 2): what shap.TreeExplainer(xgb_model).shap_values(row)
 * returns for an xgboost model (src/api/cobalt_fast_api.py:46, :100-105; shap hands xgboost models to
 * xgboost's pred_contribs, xgboost-3.0.0 src/predictor, un-vendored).  Node cover = sum_hess.
 * Per tree the contributions go to a zeroed phi_t[f+1] (slot f = the tree's expected value, the
 * cover-weighted mean of its leaves); the forest's phi is the sum of the phi_t in tree order, plus
 * base_margin in slot f.  All arithmetic is f32, one rounding per operation, in the order written
 * here (the CUDA kernel repeats it instruction for instruction).
 * PARITY PIN STATUS: "parity unpinned" against shap/xgboost (neither is installable here); pinned
 * instead to the DEFINITION: tests/test_oracle_golden.py enumerates all feature subsets of small
 * forests (Shapley value of the path-dependent conditional expectation, f64) and compares. */
#define ORC_SHAP_MAXD 12
typedef struct { int feat; float zero, one, pw; } orc_path_t;

static void shap_extend(orc_path_t* p, int ud, float zero, float one, int feat) {
  p[ud].feat = feat; p[ud].zero = zero; p[ud].one = one; p[ud].pw = ud == 0 ? 1.0f : 0.0f;
  for (int i = ud - 1; i >= 0; --i) {
    p[i + 1].pw += one * p[i].pw * (float)(i + 1) / (float)(ud + 1);
    p[i].pw = zero * p[i].pw * (float)(ud - i) / (float)(ud + 1);
  }
}
static void shap_unwind(orc_path_t* p, int ud, int k) {
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
static float shap_unwound_sum(const orc_path_t* p, int ud, int k) {
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
/* cover-weighted mean of the leaves below `nid` (post-order, f32) */
static float shap_mean(int64_t o, int nid, const int32_t* left, const int32_t* right, const float* cond,
                       const float* cover) {
  if (left[o + nid] == -1) return cond[o + nid];
  const int l = left[o + nid], r = right[o + nid];
  const float ml = shap_mean(o, l, left, right, cond, cover), mr = shap_mean(o, r, left, right, cond, cover);
  return (ml * cover[o + l] + mr * cover[o + r]) / cover[o + nid];
}
typedef struct { int nid, ud, poff, feat; float zero, one; } orc_frame_t;

/* phi_t[f+1] of one tree for one row; explicit stack, hot child first (as the recursion would) */
static void shap_tree(const float* x, int f, int64_t o, const int32_t* left, const int32_t* right,
                      const int32_t* split_index, const float* cond, const uint8_t* default_left,
                      const float* cover, float* phi) {
  orc_path_t path[(ORC_SHAP_MAXD + 2) * (ORC_SHAP_MAXD + 3) / 2];
  orc_frame_t st[2 * ORC_SHAP_MAXD + 4];
  int sp = 0;
  st[sp++] = (orc_frame_t){0, 0, 0, -1, 1.0f, 1.0f};
  while (sp > 0) {
    const orc_frame_t fr = st[--sp];
    orc_path_t* p = path + fr.poff + fr.ud + (fr.ud > 0);   /* own buffer: right behind the parent's ud+1 elements */
    if (fr.ud > 0) memcpy(p, path + fr.poff, sizeof(orc_path_t) * (size_t)fr.ud);
    int ud = fr.ud;
    shap_extend(p, ud, fr.zero, fr.one, fr.feat);
    const int nid = fr.nid;
    if (left[o + nid] == -1) {
      for (int i = 1; i <= ud; ++i) {
        const float w = shap_unwound_sum(p, ud, i);
        phi[p[i].feat] += w * (p[i].one - p[i].zero) * cond[o + nid];
      }
      continue;
    }
    const int sf = split_index[o + nid];
    const float v = x[sf];
    const int go_left = (v != v) ? default_left[o + nid] : (v < cond[o + nid]);
    const int hot = go_left ? left[o + nid] : right[o + nid];
    const int cold = go_left ? right[o + nid] : left[o + nid];
    const float w = cover[o + nid];
    const float hot_zero = cover[o + hot] / w, cold_zero = cover[o + cold] / w;
    float in_zero = 1.0f, in_one = 1.0f;
    int k = 0;
    for (; k <= ud; ++k) if (p[k].feat == sf) break;
    if (k != ud + 1) {
      in_zero = p[k].zero; in_one = p[k].one;
      shap_unwind(p, ud, k);
      ud -= 1;
    }
    const int poff = (int)(p - path);
    st[sp++] = (orc_frame_t){cold, ud + 1, poff, sf, cold_zero * in_zero, 0.0f};
    st[sp++] = (orc_frame_t){hot, ud + 1, poff, sf, hot_zero * in_zero, in_one};
  }
  (void)f;
}

/* out: [n][f+1] (last column = expected value incl. base_margin); returns -1 if a tree is deeper than ORC_SHAP_MAXD */
int orc_tree_shap(const float* X, int64_t n, int f, int64_t ldx, const int64_t* tree_off, int n_trees,
                  const int32_t* left, const int32_t* right, const int32_t* split_index,
                  const float* split_cond, const uint8_t* default_left, const float* sum_hess,
                  float base_margin, float* out) {
  for (int t = 0; t < n_trees; ++t) {            /* depth check: child ids are larger than the parent's */
    const int64_t o = tree_off[t], cnt = tree_off[t + 1] - o;
    int* depth = (int*)calloc((size_t)cnt, sizeof(int));
    int bad = 0;
    for (int64_t i = 0; i < cnt; ++i)
      if (left[o + i] != -1) {
        depth[left[o + i]] = depth[i] + 1; depth[right[o + i]] = depth[i] + 1;
        if (depth[i] + 1 > ORC_SHAP_MAXD) bad = 1;
      }
    free(depth);
    if (bad) return -1;
  }
  float* means = (float*)malloc(sizeof(float) * (size_t)(n_trees > 0 ? n_trees : 1));
  for (int t = 0; t < n_trees; ++t) means[t] = shap_mean(tree_off[t], 0, left, right, split_cond, sum_hess);
#pragma omp parallel
  {
    float* phi_t = (float*)malloc(sizeof(float) * (size_t)(f + 1));
#pragma omp for schedule(static)
    for (int64_t r = 0; r < n; ++r) {
      float* phi = out + r * (f + 1);
      for (int j = 0; j <= f; ++j) phi[j] = 0.0f;
      for (int t = 0; t < n_trees; ++t) {
        for (int j = 0; j <= f; ++j) phi_t[j] = 0.0f;
        shap_tree(X + r * ldx, f, tree_off[t], left, right, split_index, split_cond, default_left, sum_hess, phi_t);
        phi_t[f] = means[t];
        for (int j = 0; j <= f; ++j) phi[j] += phi_t[j];
      }
      phi[f] += base_margin;
    }
    free(phi_t);
  }
  free(means);
  return 0;
}
