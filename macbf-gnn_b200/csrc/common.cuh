// Shared device/host helpers for libmacbf_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/macbf_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libmacbf_b200 is written for sm_100a (B200) only"
#endif

namespace macbf {

// ---- flat parameter offsets (state_dict order; cbf.py:13-19, controller.py:16-29) --------
constexpr int CBF_W1 = 0, CBF_B1 = 384, CBF_W2 = 448, CBF_B2 = 8640, CBF_W3 = 8768, CBF_B3 = 16960,
              CBF_W4 = 17024, CBF_B4 = 17088;
constexpr int CT_EW1 = 0, CT_EB1 = 320, CT_EW2 = 384, CT_EB2 = 8576, CT_NW1 = 8704, CT_NB1 = 17152,
              CT_NW2 = 17216, CT_NB2 = 25408, CT_NW3 = 25536, CT_NB3 = 33728, CT_NW4 = 33792,
              CT_NB4 = 34048;

constexpr int NUM_SMS = 148;

// ---- error plumbing ------------------------------------------------------------------------
void set_error(const char* fmt, ...);
// Device flag consulted by the rollout kernels (neighbour search, controller): when it points to a
// non-zero int the kernel returns at once.  macbf_rollout_forward sets it around its launches so that
// the steps after every env has reached its goal cost (almost) nothing and no host sync is needed.
extern thread_local const int* tl_skip_flag;
// Query partition consulted by the neighbour-search launchers: part p of n handles a contiguous n-th of the queries
// (of the cell-sorted order on the grid path, of the agent index otherwise) and leaves the other rows of idx
// untouched.  macbf_neighbor_topk_part sets it around its launch (row-sharding one large env over several GPUs).
// When set, the tensor-core controller forward also records the arg-max slot of every (agent, channel) there
// (macbf_step_forward_ex with MACBF_STEP_SAVE_WINNERS: the training forward keeps what the backward needs).
extern thread_local uint8_t* tl_save_winners;
struct QueryPart { int part, nparts; };
extern thread_local QueryPart tl_query_part;
// Row range consulted by the launchers of the forward kernels (macbf_step_forward_rows_*): when active (end >= 0) only the
// agents [begin, end) of a single env are processed as queries / rows - their neighbour search, controller, CBF and
// loss partial sums - while every agent stays a candidate.  One large env row-sharded over several GPUs.
// both neighbour graphs of a step in one launch (neighbor_grid.cu: neighbor_topk_grid_pair): grids [B_split, 2 B_split)
// are the envs one Euler step (dt) later and write idx_next
struct GraphPair { int B_split; float dt; int32_t* idx_next; };
extern thread_local GraphPair tl_graph_pair;
struct RowRange { int begin, end; };
extern thread_local RowRange tl_row_range;
inline bool rows_active() { return tl_row_range.end >= 0; }
inline void query_range(int total, int* begin, int* end) {
  if (rows_active()) { *begin = tl_row_range.begin; *end = tl_row_range.end < total ? tl_row_range.end : total; return; }
  *begin = (int)((long long)total * tl_query_part.part / tl_query_part.nparts);
  *end = (int)((long long)total * (tl_query_part.part + 1) / tl_query_part.nparts);
}
#define MACBF_SKIP_IF_SET(skip) do { if ((skip) != nullptr && *(skip) != 0) return; } while (0)
void count_launch(int n = 1);
// Per-launch device timing (macbf_timing_begin / macbf_timing_end): when switched on, every launch site records a
// CUDA event on the profiled stream right after its kernel; off (the default) this is one thread-local test.
void timing_mark(const char* name);
#define MACBF_CHECK_ARG(cond, ...)      \
  do {                                  \
    if (!(cond)) {                      \
      macbf::set_error(__VA_ARGS__);    \
      return 1;                         \
    }                                   \
  } while (0)
#define MACBF_CHECK_LAUNCH(name)                                                     \
  do {                                                                               \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess) {                                                        \
      macbf::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));      \
      return 2;                                                                      \
    }                                                                                \
    macbf::count_launch();                                                           \
    macbf::timing_mark(name);                                                        \
  } while (0)
#define MACBF_CHECK_CUDA(expr)                                                       \
  do {                                                                               \
    cudaError_t e__ = (expr);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      macbf::set_error("%s: %s", #expr, cudaGetErrorString(e__));                    \
      return 2;                                                                      \
    }                                                                                \
  } while (0)

// ---- exact (non-contracted) fp32 arithmetic: every op is one IEEE round-to-nearest op,
//      as in the reference's eager torch elementwise kernels -----------
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }

// Candidate state j of env b: agents then obstacle points (ox, oy, 0, 0).
__device__ __forceinline__ float4 load_candidate(const float* __restrict__ states,
                                                 const float* __restrict__ obstacles, int b, int j,
                                                 int N, int M) {
  if (j < N) return __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + j);
  float2 o = __ldg(reinterpret_cast<const float2*>(obstacles) + (size_t)b * M + (j - N));
  return make_float4(o.x, o.y, 0.f, 0.f);
}

struct EdgeGeom {
  float dx, dy, dvx, dvy, eye;
};
// cbf.py:23 / controller.py:33 / core.py:188: s_i - s_j, plus the identity indicator.
__device__ __forceinline__ EdgeGeom edge_geom(const float4 si, const float4 sj, int i, int j) {
  EdgeGeom g;
  g.dx = xsub(si.x, sj.x);
  g.dy = xsub(si.y, sj.y);
  g.dvx = xsub(si.z, sj.z);
  g.dvy = xsub(si.w, sj.w);
  g.eye = (i == j) ? 1.f : 0.f;
  return g;
}
// cbf.py:24-25,34-35: sqrt(sum(square(dxy) + 1e-4))
__device__ __forceinline__ float cbf_dist(float dx, float dy) {
  return __fsqrt_rn(xadd(xadd(xmul(dx, dx), 1e-4f), xadd(xmul(dy, dy), 1e-4f)));
}
// controller.py:38: torch.norm over 2 elements = sqrt(fma(dy, dy, dx*dx)) on AVX2/AVX512 builds
__device__ __forceinline__ float ctrl_dist(float dx, float dy) {
  return __fsqrt_rn(__fmaf_rn(dy, dy, xmul(dx, dx)));
}
// core.py:193-208 on one edge
__device__ __forceinline__ bool ttc_dangerous(const EdgeGeom& g, float r2, float ttc) {
  float x = xadd(g.dx, g.eye), y = xadd(g.dy, g.eye);
  float alpha = xadd(xmul(g.dvx, g.dvx), xmul(g.dvy, g.dvy));
  float beta = xmul(2.f, xadd(xmul(x, g.dvx), xmul(y, g.dvy)));
  float gamma = xsub(xadd(xmul(x, x), xmul(y, y)), r2);
  float disc = xsub(xmul(beta, beta), xmul(xmul(4.f, alpha), gamma));
  float at = xmul(xmul(2.f, alpha), ttc);
  bool two_pos = (disc > 0.f) && (gamma > 0.f) && (beta < 0.f);
  float bp = xadd(beta, at);
  bool root = (xsub(-beta, at) < 0.f) || (xmul(bp, bp) < disc);
  return (gamma < 0.f) || (two_pos && root);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace macbf
