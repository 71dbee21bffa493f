// K1  quantile sketch (R1): every feature column is sorted exactly on the device
// (order-preserving u32 keys, 4-pass LSD radix sort, all columns in one launch per step),
// then one CTA per feature extracts the cut array.  Included by cobalt_gbdt.cu only.
#pragma once
#include "common.cuh"

namespace cb {

constexpr int kSortTile = 4096;          // keys per CTA per pass (256 threads x 16)
constexpr uint32_t kNanKey = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t key_of(float x) {
  if (x != x) return kNanKey;
  uint32_t u = __float_as_uint(x + 0.0f);   // -0 -> +0
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float val_of(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

// X [n][ldx] f32  ->  keys [f][n] u32 (column major), 32x32 tiles through shared memory
__global__ void __launch_bounds__(256) sk_keys_kernel(const float* __restrict__ X, int64_t n, int f, int64_t ldx,
                                                      uint32_t* __restrict__ keys) {
  __shared__ uint32_t t[32][33];
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  for (int k = ty; k < 32; k += 8) {
    int64_t r = r0 + k;
    int c = c0 + tx;
    t[k][tx] = (r < n && c < f) ? key_of(__ldg(X + r * ldx + c)) : kNanKey;
  }
  __syncthreads();
  for (int k = ty; k < 32; k += 8) {
    int c = c0 + k;
    int64_t r = r0 + tx;
    if (c < f && r < n) keys[(size_t)c * n + r] = t[tx][k];
  }
}

// per (tile, feature) digit histogram -> counts[f][digit][tile]
__global__ void __launch_bounds__(256) sk_count_kernel(const uint32_t* __restrict__ keys, int64_t n, int n_tiles,
                                                       int shift, uint32_t* __restrict__ counts) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int f = blockIdx.y;
  const int64_t base = (int64_t)blockIdx.x * kSortTile;
  const uint32_t* src = keys + (size_t)f * n;
  for (int k = 0; k < kSortTile / 256; ++k) {
    int64_t i = base + k * 256 + threadIdx.x;
    if (i < n) atomicAdd(&h[(src[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[((size_t)f * 256 + threadIdx.x) * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// one CTA per feature: exclusive scan over [digit][tile] (in place)
__global__ void __launch_bounds__(1024) sk_scan_kernel(uint32_t* __restrict__ counts, int n_tiles) {
  __shared__ uint32_t part[1024];
  uint32_t* c = counts + (size_t)blockIdx.x * 256 * n_tiles;
  const int total = 256 * n_tiles;
  const int per = (total + 1023) / 1024;
  const int b = threadIdx.x * per, e = min(total, b + per);
  uint32_t s = 0;
  for (int i = b; i < e; ++i) s += c[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    uint32_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  uint32_t run = part[threadIdx.x] - s;
  for (int i = b; i < e; ++i) { uint32_t v = c[i]; c[i] = run; run += v; }
}

// stable scatter of one tile: tile order = (warp, round, lane); ranks by warp match + per-warp counters
__global__ void __launch_bounds__(256) sk_scatter_kernel(const uint32_t* __restrict__ keys_in, int64_t n, int n_tiles,
                                                         int shift, const uint32_t* __restrict__ offsets,
                                                         uint32_t* __restrict__ keys_out) {
  __shared__ uint32_t cnt[8][256];    // per warp digit counts, then running offsets
  const int f = blockIdx.y;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t* src = keys_in + (size_t)f * n;
  uint32_t* dst = keys_out + (size_t)f * n;
  const int64_t base = (int64_t)blockIdx.x * kSortTile + (int64_t)wid * (kSortTile / 8);
  for (int d = lane; d < 256; d += 32) cnt[wid][d] = 0;
  __syncwarp();
  uint32_t key[16];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    int64_t i = base + r * 32 + lane;
    bool ok = i < n;
    key[r] = ok ? src[i] : 0u;
    uint32_t d = (key[r] >> shift) & 255u;
    uint32_t act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      uint32_t peers = __match_any_sync(act, d);
      if ((peers & lt) == 0) cnt[wid][d] += __popc(peers);
    }
    __syncwarp();
  }
  __syncthreads();
  // digit d (one thread each): turn per-warp counts into global start offsets
  {
    const int d = threadIdx.x;
    uint32_t run = offsets[((size_t)f * 256 + d) * n_tiles + blockIdx.x];
#pragma unroll
    for (int w = 0; w < 8; ++w) { uint32_t c = cnt[w][d]; cnt[w][d] = run; run += c; }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    int64_t i = base + r * 32 + lane;
    bool ok = i < n;
    uint32_t d = (key[r] >> shift) & 255u;
    uint32_t act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      uint32_t peers = __match_any_sync(act, d);
      uint32_t pos = cnt[wid][d] + __popc(peers & lt);
      dst[pos] = key[r];
      __syncwarp(act);
      if ((peers & lt) == 0) cnt[wid][d] += __popc(peers);
    }
    __syncwarp();
  }
}

// one CTA per feature over its sorted keys: R1 cut extraction
__global__ void __launch_bounds__(1024) sk_cuts_kernel(const uint32_t* __restrict__ keys, int64_t n, int max_bin,
                                                       float* __restrict__ cuts, int32_t* __restrict__ n_cuts,
                                                       float* __restrict__ min_vals) {
  __shared__ uint32_t part[1024];
  __shared__ float dist[CB_MAX_BINS + 1];
  __shared__ int64_t s_m;
  const int f = blockIdx.x;
  const uint32_t* k = keys + (size_t)f * n;
  float* cj = cuts + (size_t)f * CB_CUT_STRIDE;
  int cap = max_bin < CB_MAX_BINS ? max_bin : CB_MAX_BINS;
  if (cap < 2) cap = 2;
  if (threadIdx.x == 0) {                      // m = number of non-missing keys
    int64_t lo = 0, hi = n;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (k[mid] < kNanKey) lo = mid + 1; else hi = mid; }
    s_m = lo;
  }
  for (int i = threadIdx.x; i < CB_CUT_STRIDE; i += blockDim.x) cj[i] = __int_as_float(0x7f800000);
  __syncthreads();
  const int64_t m = s_m;
  if (m == 0) {
    if (threadIdx.x == 0) {
      const float c = 1e-5f;
      min_vals[f] = __fsub_rn(c, __fadd_rn(fabsf(c), 1e-5f));
      cj[0] = __fadd_rn(c, __fadd_rn(fabsf(c), 1e-5f));
      n_cuts[f] = 1;
    }
    return;
  }
  // distinct values: boundaries i in [1, m) with k[i] != k[i-1]; thread t owns a contiguous chunk
  const int64_t per = (m + 1023) / 1024;
  const int64_t b = min(m, (int64_t)threadIdx.x * per), e = min(m, b + per);
  uint32_t c = 0;
  for (int64_t i = max(b, (int64_t)1); i < e; ++i) c += (k[i] != k[i - 1]);
  part[threadIdx.x] = c;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    uint32_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  const uint32_t distinct = part[1023] + 1;
  const float vmin = val_of(k[0]), vmax = val_of(k[m - 1]);
  const float sentinel = __fadd_rn(vmax, __fadd_rn(fabsf(vmax), 1e-5f));
  if (distinct <= (uint32_t)cap) {
    uint32_t o = part[threadIdx.x] - c;         // boundaries before this chunk
    for (int64_t i = max(b, (int64_t)1); i < e; ++i)
      if (k[i] != k[i - 1]) dist[o++] = val_of(k[i]);
    __syncthreads();
    for (int i = threadIdx.x; i < (int)distinct - 1; i += blockDim.x) cj[i] = dist[i];
    if (threadIdx.x == 0) {
      cj[distinct - 1] = sentinel;
      n_cuts[f] = (int)distinct;
      min_vals[f] = __fsub_rn(vmin, __fadd_rn(fabsf(vmin), 1e-5f));
    }
    return;
  }
  if (threadIdx.x == 0) {
    int nc = 0;
    float last = vmin;
    for (int q = 1; q < cap; ++q) {
      // floor(q * m / cap) without overflow: m < 2^31, q < 256
      int64_t rank = ((int64_t)q * m) / cap;
      float v = val_of(k[rank]);
      if (v > last) { cj[nc++] = v; last = v; }
    }
    cj[nc++] = sentinel;
    n_cuts[f] = nc;
    min_vals[f] = __fsub_rn(vmin, __fadd_rn(fabsf(vmin), 1e-5f));
  }
}

}  // namespace cb

extern "C" {

int cb_sketch_workspace_bytes(int64_t n, int f, size_t* bytes) {
  if (!bytes || n < 0 || f <= 0) return cb::fail(CB_ERR_INVALID, "cb_sketch_workspace_bytes: bad arguments");
  cb::Bump b(nullptr);
  const int64_t nt = (n + cb::kSortTile - 1) / cb::kSortTile;
  b.take<uint32_t>((size_t)n * f);
  b.take<uint32_t>((size_t)n * f);
  b.take<uint32_t>((size_t)f * 256 * (nt > 0 ? nt : 1));
  *bytes = b.used();
  return CB_OK;
}

int cb_sketch_cuts(const float* X, int64_t n, int f, int64_t ldx, int max_bin, float* cuts, int32_t* n_cuts,
                   float* min_vals, void* ws, size_t ws_bytes, void* stream) {
  using namespace cb;
  if (!X || !cuts || !n_cuts || !min_vals || !ws || n <= 0 || f <= 0 || ldx < f || max_bin < 2 || n > 2147483647LL ||
      f > 65535)
    return fail(CB_ERR_INVALID, "cb_sketch_cuts: bad arguments");
  size_t need = 0;
  cb_sketch_workspace_bytes(n, f, &need);
  if (ws_bytes < need) return fail(CB_ERR_WORKSPACE, "cb_sketch_cuts: workspace %zu < %zu bytes", ws_bytes, need);
  cudaStream_t s = (cudaStream_t)stream;
  const int nt = (int)((n + kSortTile - 1) / kSortTile);
  Bump b(ws);
  uint32_t* ka = b.take<uint32_t>((size_t)n * f);
  uint32_t* kb = b.take<uint32_t>((size_t)n * f);
  uint32_t* counts = b.take<uint32_t>((size_t)f * 256 * nt);
  sk_keys_kernel<<<dim3((unsigned)((n + 31) / 32), (f + 31) / 32), 256, 0, s>>>(X, n, f, ldx, ka);
  CB_LAUNCH_CHECK();
  for (int pass = 0; pass < 4; ++pass) {
    sk_count_kernel<<<dim3(nt, f), 256, 0, s>>>(ka, n, nt, pass * 8, counts);
    sk_scan_kernel<<<f, 1024, 0, s>>>(counts, nt);
    sk_scatter_kernel<<<dim3(nt, f), 256, 0, s>>>(ka, n, nt, pass * 8, counts, kb);
    CB_LAUNCH_CHECK();
    uint32_t* t = ka; ka = kb; kb = t;
  }
  sk_cuts_kernel<<<f, 1024, 0, s>>>(ka, n, max_bin, cuts, n_cuts, min_vals);
  CB_LAUNCH_CHECK();
  return CB_OK;
}

}  // extern "C"
