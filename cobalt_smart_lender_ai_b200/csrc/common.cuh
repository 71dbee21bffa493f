// Shared helpers for libcobalt_gbdt.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/cobalt_gbdt.h"

namespace cb {

extern thread_local char g_err[512];

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

#define CB_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return cb::fail(CB_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

#define CB_LAUNCH_CHECK()                                                                          \
  do {                                                                                             \
    cudaError_t e_ = cudaGetLastError();                                                           \
    if (e_ != cudaSuccess)                                                                         \
      return cb::fail(CB_ERR_CUDA, "%s:%d kernel launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
  } while (0)

// bump allocator over a caller-provided workspace (256-byte aligned blocks)
struct Bump {
  char* base;
  size_t off = 0;
  explicit Bump(void* p) : base(static_cast<char*>(p)) {}
  template <typename T>
  T* take(size_t count) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
  size_t used() const { return (off + 255) & ~size_t(255); }
};

struct DeviceInfo {
  int sm_count = 0;
  int smem_optin = 0;
  int device = -1;
};
int device_info(DeviceInfo* out);

__device__ __forceinline__ int64_t warp_sum_i64(int64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- R4: the specified sigmoid.  Same operation sequence as the oracle; every
// operation is a separately rounded IEEE f32 op (no fma contraction, no fast math).
__device__ __forceinline__ float spec_exp(float t) {
  float nf = rintf(__fmul_rn(t, 1.44269504088896341f));
  float r = __fsub_rn(t, __fmul_rn(nf, 0.693145751953125f));
  r = __fsub_rn(r, __fmul_rn(nf, 1.42860682030941723212e-6f));
  float p = 1.0f / 720.0f;
  p = __fadd_rn(__fmul_rn(p, r), 1.0f / 120.0f);
  p = __fadd_rn(__fmul_rn(p, r), 1.0f / 24.0f);
  p = __fadd_rn(__fmul_rn(p, r), 1.0f / 6.0f);
  p = __fadd_rn(__fmul_rn(p, r), 0.5f);
  p = __fadd_rn(__fmul_rn(p, r), 1.0f);
  p = __fadd_rn(__fmul_rn(p, r), 1.0f);
  int n = (int)nf;
  int n1 = n / 2, n2 = n - n1;
  float a = __uint_as_float((uint32_t)(n1 + 127) << 23);
  float b = __uint_as_float((uint32_t)(n2 + 127) << 23);
  return __fmul_rn(__fmul_rn(p, a), b);
}

__device__ __forceinline__ float spec_sigmoid(float x) {
  float t = -x;
  t = t > 88.7f ? 88.7f : t;
  t = t < -87.0f ? -87.0f : t;
  float d = __fadd_rn(spec_exp(t), 1.0f);
  return __fdiv_rn(1.0f, d);
}

__device__ __forceinline__ uint32_t spec_hash(uint32_t seed, uint32_t a, uint64_t b) {
  uint32_t h = seed ^ 0x9E3779B9u;
  h ^= a * 0x85EBCA6Bu; h = (h << 13) | (h >> 19); h = h * 5u + 0xE6546B64u;
  h ^= (uint32_t)(b & 0xFFFFFFFFu) * 0xCC9E2D51u; h = (h << 13) | (h >> 19); h = h * 5u + 0xE6546B64u;
  h ^= (uint32_t)(b >> 32) * 0x1B873593u;
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}

// 2^-s as a double, exactly (s in [-1000, 1000]): replaces a library ldexp call in per-thread prologues
__device__ __forceinline__ double pow2neg(int s) { return __longlong_as_double((long long)(1023 - s) << 52); }

// ---- R6 gain / weight in f64, one rounding per operation
__device__ __forceinline__ double thr_l1(double g, double a) {
  if (g > a) return __dsub_rn(g, a);
  if (g < -a) return __dadd_rn(g, a);
  return 0.0;
}
__device__ __forceinline__ double calc_gain(double G, double H, const cb_tree_params& p) {
  if (H < (double)p.min_child_weight || H <= 0.0) return 0.0;
  double t = p.reg_alpha == 0.0f ? G : thr_l1(G, (double)p.reg_alpha);
  return __ddiv_rn(__dmul_rn(t, t), __dadd_rn(H, (double)p.reg_lambda));
}
__device__ __forceinline__ double calc_weight(double G, double H, const cb_tree_params& p) {
  if (H < (double)p.min_child_weight || H <= 0.0) return 0.0;
  double t = p.reg_alpha == 0.0f ? G : thr_l1(G, (double)p.reg_alpha);
  return __ddiv_rn(-t, __dadd_rn(H, (double)p.reg_lambda));
}

#ifdef CB_LEVEL_TRACE
// debug build only (tools/level_trace.py): %globaltimer at the start / end of every CTA of the level-tail kernels
__device__ unsigned long long g_level_trace[6][2048][2];
__device__ __forceinline__ unsigned long long level_trace_gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
struct LevelTraceScope {
  int k;
  __device__ LevelTraceScope(int kernel) : k(kernel) { if (threadIdx.x == 0 && blockIdx.x + blockIdx.y * gridDim.x < 2048) g_level_trace[k][blockIdx.x + blockIdx.y * gridDim.x][0] = level_trace_gt(); }
  __device__ ~LevelTraceScope() { if (threadIdx.x == 0 && blockIdx.x + blockIdx.y * gridDim.x < 2048) g_level_trace[k][blockIdx.x + blockIdx.y * gridDim.x][1] = level_trace_gt(); }
};
#define CB_LEVEL_TRACE_SCOPE(kernel) LevelTraceScope level_trace_scope_(kernel)
#else
#define CB_LEVEL_TRACE_SCOPE(kernel)
#endif

// ---- programmatic dependent launch: every kernel of a boosting round is launched with the stream-serialisation
// attribute (cobalt_gbdt.cu launch_dep) and starts with pdl_enter(): it lets ITS successor be scheduled as soon as all
// of this grid's CTAs run, then blocks until its predecessor has completed and flushed its memory.  The launch gap and
// the scheduling ramp of kernel k + 1 overlap the tail of kernel k; ordering is unchanged (wait = full completion).
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

constexpr int kChunkRows = 16352;  // rows per int32 partial histogram: 16352 * 2^17 < 2^31
constexpr int kPartTile = 1024;    // positions per partition tile

}  // namespace cb
