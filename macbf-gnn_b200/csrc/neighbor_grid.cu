// K1 (grid path) - exact K nearest candidates through a uniform grid (K <= 128).
//
// Same contract as neighbor.cu (core.py:234-250; lexicographic (key, index) order, keys computed
// with the reference's op sequence), but a query only looks at the candidates of the grid cells
// around it instead of all N + M:
//
//   grid_build_kernel  one CTA per env: bounding box -> G x G cells -> counting sort of the
//                      candidates by cell (shared-memory histogram, scan, scatter) into
//                      sorted[b][.] = (x, y, index) plus cell_start[b][G*G+1].
//   knn_grid_kernel    one warp per query: ring r = 0, 1, 2, ... of cells around the query's cell.
//                      The <= 4r cell ranges of a ring are flattened with a warp prefix sum so that
//                      every lane evaluates one candidate per step; candidates that pass the exact
//                      pre-filter are merged into the register-resident sorted list (entry e in
//                      lane e) by a warp bitonic merge, or inserted one by one when only a few pass.
//                      The search stops once the K-th key is smaller than the distance from the
//                      query to the edge of the visited square (everything unvisited is farther).
//
// Exactness does not depend on how the agents are distributed: a clustered swarm only makes
// the rings grow until the stop test holds (worst case: the whole grid = brute force).
#include "common.cuh"

namespace macbf {

constexpr int GRID_GMAX = 160;                // cells per axis (dynamic shared-memory histogram of G*G + 1 ints <= 100 KB)
constexpr int GRID_BUILD_THREADS = 1024;
constexpr int GRID_QUERY_THREADS = 256;

struct GridHeader {
  float ox, oy, inv_cs, cs;
};

__device__ __forceinline__ int cell_coord(float x, float o, float inv_cs, int G) {
  int c = (int)floorf((x - o) * inv_cs);
  return min(max(c, 0), G - 1);
}

__global__ void __launch_bounds__(GRID_BUILD_THREADS)
grid_build_kernel(const float* __restrict__ states, const float* __restrict__ obstacles, int N, int M, int G,
                  float4* __restrict__ sorted, int* __restrict__ cell_start, GridHeader* __restrict__ header,
                  const int* __restrict__ skip, int Bs, float dt) {
  MACBF_SKIP_IF_SET(skip);
  extern __shared__ int cnt[];                  // G*G + 1: histogram -> cell starts -> scatter cursors
  __shared__ float red[4][32];
  __shared__ int wsum[32];
  __shared__ GridHeader hd;
  const int b = blockIdx.x, C = N + M, cells = G * G;
  // Bs > 0 (both graphs of a step in one launch): grid b >= Bs is the grid of env b - Bs one Euler step later. The
  // neighbour search sees positions only, and p' = p + v dt does not depend on the controller (core.py:133-137 with
  // core.py:74-86): the same two ops as the controller's Euler epilogue, so the points are the s_next it writes bit for bit.
  const bool adv = Bs > 0 && b >= Bs;
  const int b_src = adv ? b - Bs : b;
  auto cand = [&](int t) {
    float4 sv = load_candidate(states, obstacles, b_src, t, N, M);
    if (adv && t < N) { sv.x = xadd(sv.x, xmul(sv.z, dt)); sv.y = xadd(sv.y, xmul(sv.w, dt)); }
    return sv;
  };
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // ---- bounding box ----
  float xmin = 3.0e38f, ymin = 3.0e38f, xmax = -3.0e38f, ymax = -3.0e38f;
  for (int t0 = threadIdx.x; t0 < C; t0 += 4 * GRID_BUILD_THREADS) {
    float4 sv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * GRID_BUILD_THREADS;
      sv[u] = cand(t < C ? t : t0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      xmin = fminf(xmin, sv[u].x); xmax = fmaxf(xmax, sv[u].x);
      ymin = fminf(ymin, sv[u].y); ymax = fmaxf(ymax, sv[u].y);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
    ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
    xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
    ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
  }
  if (lane == 0) { red[0][warp] = xmin; red[1][warp] = ymin; red[2][warp] = xmax; red[3][warp] = ymax; }
  for (int t = threadIdx.x; t <= cells; t += GRID_BUILD_THREADS) cnt[t] = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < GRID_BUILD_THREADS / 32; ++w) {
      red[0][0] = fminf(red[0][0], red[0][w]); red[1][0] = fminf(red[1][0], red[1][w]);
      red[2][0] = fmaxf(red[2][0], red[2][w]); red[3][0] = fmaxf(red[3][0], red[3][w]);
    }
    float ext = fmaxf(red[2][0] - red[0][0], red[3][0] - red[1][0]);
    float cs = ext / (float)G;
    if (!(cs > 0.f) || !(cs < 3.0e38f)) cs = 1.f;          // all candidates coincide (or non-finite input)
    hd.ox = red[0][0]; hd.oy = red[1][0]; hd.cs = cs; hd.inv_cs = 1.f / cs;
    header[b] = hd;
  }
  __syncthreads();
  const GridHeader g = hd;
  // ---- histogram ----
  for (int t0 = threadIdx.x; t0 < C; t0 += 4 * GRID_BUILD_THREADS) {
    float4 sv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * GRID_BUILD_THREADS;
      sv[u] = cand(t < C ? t : t0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (t0 + u * GRID_BUILD_THREADS < C)
        atomicAdd(&cnt[cell_coord(sv[u].y, g.oy, g.inv_cs, G) * G + cell_coord(sv[u].x, g.ox, g.inv_cs, G)], 1);
  }
  __syncthreads();
  // ---- exclusive scan of cnt[0..cells) (each thread owns a contiguous chunk) ----
  const int per = (cells + GRID_BUILD_THREADS - 1) / GRID_BUILD_THREADS;
  const int c_begin = min(threadIdx.x * per, cells), c_end = min(c_begin + per, cells);
  int local = 0;
  for (int c = c_begin; c < c_end; ++c) local += cnt[c];
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = wsum[lane], s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int u = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += u;
    }
    wsum[lane] = s - v;   // exclusive prefix of the warp sums
  }
  __syncthreads();
  int run = wsum[warp] + incl - local;
  for (int c = c_begin; c < c_end; ++c) {
    const int n = cnt[c];
    cnt[c] = run;
    run += n;
  }
  if (threadIdx.x == GRID_BUILD_THREADS - 1) cnt[cells] = C;
  __syncthreads();
  for (int t = threadIdx.x; t <= cells; t += GRID_BUILD_THREADS) cell_start[(size_t)b * (GRID_GMAX * GRID_GMAX + 1) + t] = cnt[t];
  __syncthreads();                              // the starts are out: cnt now serves as the scatter cursors
  // ---- scatter (order inside a cell is arbitrary: the query orders by (key, index) explicitly); four candidates in
  //      flight per thread - a single CTA owns the whole env, so the loads' latency is what this loop costs ----
  for (int t0 = threadIdx.x; t0 < C; t0 += 4 * GRID_BUILD_THREADS) {
    float4 sv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * GRID_BUILD_THREADS;
      sv[u] = cand(t < C ? t : t0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * GRID_BUILD_THREADS;
      if (t < C) {
        const int cell = cell_coord(sv[u].y, g.oy, g.inv_cs, G) * G + cell_coord(sv[u].x, g.ox, g.inv_cs, G);
        const int pos = atomicAdd(&cnt[cell], 1);
        sorted[(size_t)b * C + pos] = make_float4(sv[u].x, sv[u].y, __int_as_float(t), 0.f);
      }
    }
  }
}

// ---- the same grid for LARGE envs, built by many CTAs per env (one CTA would spend ~100 us on 65,536 points) ----
//   grid_big_bbox_kernel     per-CTA bounding box -> ordered-integer atomicMin / atomicMax per env
//   grid_big_hist_kernel     header from the box; cell histogram with global atomics
//   grid_big_scan_kernel     one CTA per env: exclusive scan -> cell_start, and a copy as the scatter cursors
//   grid_big_scatter_kernel  sorted[cursor[cell]++] = (x, y, index)
// Scratch per env (behind the header array): bbox[4] as ordered uints, cursors[G*G + 1].
constexpr int GRID_BIG_THREADS = 256;
constexpr int GRID_BIG_PER_CTA = 2048;                 // candidates per CTA
constexpr int GRID_BIG_MIN_C = 8192;                   // envs with at least this many candidates take this path

__device__ __forceinline__ unsigned f2ord(float f) {   // order-preserving float -> uint
  const unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ GridHeader header_from_box(const unsigned* __restrict__ box, int G) {
  const float x0 = ord2f(box[0]), y0 = ord2f(box[1]), x1 = ord2f(box[2]), y1 = ord2f(box[3]);
  float cs = fmaxf(x1 - x0, y1 - y0) / (float)G;
  if (!(cs > 0.f) || !(cs < 3.0e38f)) cs = 1.f;          // all candidates coincide (or non-finite input)
  GridHeader hd;
  hd.ox = x0; hd.oy = y0; hd.cs = cs; hd.inv_cs = 1.f / cs;
  return hd;
}

__global__ void __launch_bounds__(GRID_BIG_THREADS)
grid_big_init_kernel(unsigned* __restrict__ box, int* __restrict__ cursors, int cells_p1, const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y;
  for (int t = blockIdx.x * GRID_BIG_THREADS + threadIdx.x; t < cells_p1; t += gridDim.x * GRID_BIG_THREADS)
    cursors[(size_t)b * (GRID_GMAX * GRID_GMAX + 1) + t] = 0;
  if (blockIdx.x == 0 && threadIdx.x < 4) box[b * 4 + threadIdx.x] = threadIdx.x < 2 ? 0xffffffffu : 0u;
}

__global__ void __launch_bounds__(GRID_BIG_THREADS)
grid_big_bbox_kernel(const float* __restrict__ states, const float* __restrict__ obstacles, int N, int M,
                     unsigned* __restrict__ box, const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y, C = N + M;
  const int t_begin = blockIdx.x * GRID_BIG_PER_CTA, t_end = min(C, t_begin + GRID_BIG_PER_CTA);
  float xmin = 3.0e38f, ymin = 3.0e38f, xmax = -3.0e38f, ymax = -3.0e38f;
  for (int t = t_begin + threadIdx.x; t < t_end; t += GRID_BIG_THREADS) {
    const float4 sv = load_candidate(states, obstacles, b, t, N, M);
    xmin = fminf(xmin, sv.x); xmax = fmaxf(xmax, sv.x);
    ymin = fminf(ymin, sv.y); ymax = fmaxf(ymax, sv.y);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
    ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
    xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
    ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
  }
  if ((threadIdx.x & 31) == 0 && xmin <= xmax) {
    atomicMin(&box[b * 4 + 0], f2ord(xmin)); atomicMin(&box[b * 4 + 1], f2ord(ymin));
    atomicMax(&box[b * 4 + 2], f2ord(xmax)); atomicMax(&box[b * 4 + 3], f2ord(ymax));
  }
}

__global__ void __launch_bounds__(GRID_BIG_THREADS)
grid_big_hist_kernel(const float* __restrict__ states, const float* __restrict__ obstacles, int N, int M, int G,
                     const unsigned* __restrict__ box, int* __restrict__ cursors, GridHeader* __restrict__ header,
                     const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y, C = N + M;
  const GridHeader g = header_from_box(box + b * 4, G);
  if (blockIdx.x == 0 && threadIdx.x == 0) header[b] = g;
  int* cnt = cursors + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  const int t_begin = blockIdx.x * GRID_BIG_PER_CTA, t_end = min(C, t_begin + GRID_BIG_PER_CTA);
  for (int t = t_begin + threadIdx.x; t < t_end; t += GRID_BIG_THREADS) {
    const float4 sv = load_candidate(states, obstacles, b, t, N, M);
    atomicAdd(&cnt[cell_coord(sv.y, g.oy, g.inv_cs, G) * G + cell_coord(sv.x, g.ox, g.inv_cs, G)], 1);
  }
}

__global__ void __launch_bounds__(GRID_BUILD_THREADS)
grid_big_scan_kernel(int C, int G, int* __restrict__ cursors, int* __restrict__ cell_start, const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  extern __shared__ int cnt[];                  // the env's histogram, staged with coalesced loads
  __shared__ int wsum[32];
  const int b = blockIdx.x, cells = G * G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int* cur = cursors + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  int* cst = cell_start + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  for (int t = threadIdx.x; t < cells; t += GRID_BUILD_THREADS) cnt[t] = cur[t];
  __syncthreads();
  const int per = (cells + GRID_BUILD_THREADS - 1) / GRID_BUILD_THREADS;
  const int c_begin = min(threadIdx.x * per, cells), c_end = min(c_begin + per, cells);
  int local = 0;
  for (int c = c_begin; c < c_end; ++c) local += cnt[c];
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = wsum[lane], s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int u = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += u;
    }
    wsum[lane] = s - v;
  }
  __syncthreads();
  int run = wsum[warp] + incl - local;
  for (int c = c_begin; c < c_end; ++c) {
    const int n = cnt[c];
    cnt[c] = run;
    run += n;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < cells; t += GRID_BUILD_THREADS) {
    const int v = cnt[t];
    cur[t] = v;                                   // scatter cursor
    cst[t] = v;
  }
  if (threadIdx.x == 0) cst[cells] = C;
}

__global__ void __launch_bounds__(GRID_BIG_THREADS)
grid_big_scatter_kernel(const float* __restrict__ states, const float* __restrict__ obstacles, int N, int M, int G,
                        const GridHeader* __restrict__ header, int* __restrict__ cursors, float4* __restrict__ sorted,
                        const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y, C = N + M;
  const GridHeader g = header[b];
  int* cur = cursors + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  const int t_begin = blockIdx.x * GRID_BIG_PER_CTA, t_end = min(C, t_begin + GRID_BIG_PER_CTA);
  for (int t = t_begin + threadIdx.x; t < t_end; t += GRID_BIG_THREADS) {
    const float4 sv = load_candidate(states, obstacles, b, t, N, M);
    const int cell = cell_coord(sv.y, g.oy, g.inv_cs, G) * G + cell_coord(sv.x, g.ox, g.inv_cs, G);
    const int pos = atomicAdd(&cur[cell], 1);
    sorted[(size_t)b * C + pos] = make_float4(sv.x, sv.y, __int_as_float(t), 0.f);
  }
}

// ---- packed (key, index) entries: key bits in the high word (keys are positive floats, so the
//      unsigned order of the bits is the numeric order), candidate index in the low word ----------
typedef unsigned long long u64;
constexpr u64 ENTRY_INF = 0x7f8000007fffffffull;

__device__ __forceinline__ u64 shfl64(u64 v, int src) {
  return ((u64)__shfl_sync(0xffffffffu, (unsigned)(v >> 32), src) << 32) | __shfl_sync(0xffffffffu, (unsigned)v, src);
}
__device__ __forceinline__ u64 shfl64_xor(u64 v, int m) {
  return ((u64)__shfl_xor_sync(0xffffffffu, (unsigned)(v >> 32), m) << 32) | __shfl_xor_sync(0xffffffffu, (unsigned)v, m);
}
__device__ __forceinline__ u64 shfl64_up(u64 v, int d) {
  return ((u64)__shfl_up_sync(0xffffffffu, (unsigned)(v >> 32), d) << 32) | __shfl_up_sync(0xffffffffu, (unsigned)v, d);
}
__device__ __forceinline__ u64 umin64(u64 a, u64 b) { return a < b ? a : b; }
__device__ __forceinline__ u64 umax64(u64 a, u64 b) { return a < b ? b : a; }

// ascending bitonic sort of one entry per lane
__device__ __forceinline__ u64 warp_sort32(u64 x, int lane) {
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const u64 p = shfl64_xor(x, j);
      const bool up = (lane & k) == 0 || k == 32;
      const bool lower = (lane & j) == 0;
      x = (lower == up) ? umin64(x, p) : umax64(x, p);
    }
  }
  return x;
}
// list, fresh: ascending; returns the 32 smallest of the 64, ascending
__device__ __forceinline__ u64 warp_merge32(u64 list, u64 fresh, int lane) {
  u64 y = umin64(list, shfl64(fresh, 31 - lane));   // bitonic
#pragma unroll
  for (int j = 16; j > 0; j >>= 1) {
    const u64 p = shfl64_xor(y, j);
    y = ((lane & j) == 0) ? umin64(y, p) : umax64(y, p);
  }
  return y;
}

// S = ceil(K / 32) entries of the sorted best list per lane: entry e lives in lane e % 32, slot e / 32 (S = 1: K <= 32,
// with the bitonic merge for dense batches; S = 2, 4: K <= 64, 128 - the dense-neighbourhood end of the C5 sweep, where
// the brute-force search costs O(N^2)).
template <int S>
__global__ void __launch_bounds__(GRID_QUERY_THREADS)
knn_grid_kernel(const float* __restrict__ states, const float4* __restrict__ sorted,
                const int* __restrict__ cell_start, const GridHeader* __restrict__ header, int N, int C, int G,
                int K, int32_t* __restrict__ idx_out, const int* __restrict__ skip, int q_begin, int q_end) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int i = q_begin + blockIdx.x * (GRID_QUERY_THREADS / 32) + (threadIdx.x >> 5);
  if (i >= q_end) return;                              // whole warp
  const GridHeader g = header[b];
  const int* cs = cell_start + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  const float4* pts = sorted + (size_t)b * C;
  const float4 sq = __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + i);
  const float qx = sq.x, qy = sq.y;
  const int cx = cell_coord(qx, g.ox, g.inv_cs, G), cy = cell_coord(qy, g.oy, g.inv_cs, G);

  u64 list[S];                                         // entry e of the sorted best list: lane e % 32, slot e / 32
#pragma unroll
  for (int sl = 0; sl < S; ++sl) list[sl] = ENTRY_INF;
  float vthr = __int_as_float(0x7f800000);             // a candidate with v >= vthr cannot enter the list
  const int kslot = (K - 1) >> 5, klane = (K - 1) & 31;
  auto kth_entry = [&]() -> u64 {
    u64 v = list[0];
#pragma unroll
    for (int sl = 1; sl < S; ++sl)
      if (kslot == sl) v = list[sl];
    return shfl64(v, klane);
  };

  for (int r = 0; r < G; ++r) {
    // ---- the cell ranges of ring r, one per lane (chunks of 32 ranges) ----
    const int nranges = (r == 0) ? 1 : 4 * r;
    for (int rb = 0; rb < nranges; rb += 32) {
      const int t = rb + lane;
      int start = 0, len = 0;
      if (t < nranges) {
        int y, x0, x1;
        if (r == 0) { y = cy; x0 = cx; x1 = cx; }
        else if (t == 0) { y = cy - r; x0 = cx - r; x1 = cx + r; }
        else if (t == 1) { y = cy + r; x0 = cx - r; x1 = cx + r; }
        else {
          const int m = (t - 2) >> 1;
          y = cy - r + 1 + m;
          x0 = x1 = ((t - 2) & 1) ? cx + r : cx - r;
        }
        x0 = max(x0, 0); x1 = min(x1, G - 1);
        if (y >= 0 && y < G && x0 <= x1 && (r == 0 || t < 2 || y < cy + r)) {
          start = __ldg(cs + y * G + x0);
          len = __ldg(cs + y * G + x1 + 1) - start;
        }
      }
      int incl = len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      const int excl = incl - len;
      for (int t0 = 0; t0 < total; t0 += 32) {
        // ---- load-balanced fetch: virtual candidate t0 + lane -> (range, offset) ----
        const int tv = t0 + lane;
        int lo = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
          const int v = __shfl_sync(0xffffffffu, incl, lo + step - 1);
          if (v <= tv) lo += step;
        }
        lo = min(lo, 31);
        const int rs = __shfl_sync(0xffffffffu, start, lo), re = __shfl_sync(0xffffffffu, excl, lo);
        const bool valid = tv < total;
        float v = __int_as_float(0x7f800000);
        int cid = 0x7fffffff;
        if (valid) {
          const float4 p = __ldg(pts + rs + (tv - re));
          cid = __float_as_int(p.z);
          // core.py:238, op by op
          const float dx = xsub(qx, p.x), dy = xsub(qy, p.y);
          v = xadd(xadd(xmul(dx, dx), 1e-6f), xadd(xmul(dy, dy), 1e-6f));
        }
        const bool passf = valid && (v < vthr);
        const unsigned pass = __ballot_sync(0xffffffffu, passf);
        if (pass == 0u) continue;
        const u64 cand = passf ? (((u64)__float_as_uint(__fsqrt_rn(v)) << 32) | (unsigned)cid) : ENTRY_INF;
        if (S == 1 && __popc(pass) >= 6) {
          list[0] = warp_merge32(list[0], warp_sort32(cand, lane), lane);
        } else {
          unsigned rem = pass;
          while (rem) {
            const int src = __ffs(rem) - 1;
            rem &= rem - 1;
            const u64 c = shfl64(cand, src);
            if (c < kth_entry()) {                      // exact lexicographic test (key, index)
              int pos = 0;
#pragma unroll
              for (int sl = 0; sl < S; ++sl) pos += __popc(__ballot_sync(0xffffffffu, list[sl] < c));
#pragma unroll
              for (int sl = S - 1; sl >= 0; --sl) {
                u64 up = shfl64_up(list[sl], 1);
                if (sl > 0) {
                  const u64 wrap = shfl64(list[sl - 1], 31);
                  if (lane == 0) up = wrap;
                }
                const int e = sl * 32 + lane;
                if (e > pos) list[sl] = up;
                else if (e == pos) list[sl] = c;
              }
            }
          }
        }
        // v >= RU(next(key_K)^2)  =>  sqrt_rn(v) > key_K: such a candidate can never enter
        const float kK = __uint_as_float((unsigned)(kth_entry() >> 32));
        const float kn = __uint_as_float(__float_as_uint(kK) + 1u);
        vthr = (kK < __int_as_float(0x7f800000)) ? __fmul_ru(kn, kn) : kK;
      }
    }
    // ---- stop test: everything not yet visited lies outside the square of rings 0..r ----
    const float kK = __uint_as_float((unsigned)(kth_entry() >> 32));
    float bound = __int_as_float(0x7f800000);
    if (cx - r > 0) bound = fminf(bound, qx - (g.ox + (float)(cx - r) * g.cs));
    if (cx + r < G - 1) bound = fminf(bound, (g.ox + (float)(cx + r + 1) * g.cs) - qx);
    if (cy - r > 0) bound = fminf(bound, qy - (g.oy + (float)(cy - r) * g.cs));
    if (cy + r < G - 1) bound = fminf(bound, (g.oy + (float)(cy + r + 1) * g.cs) - qy);
    if (bound == __int_as_float(0x7f800000)) break;     // the square already covers the whole grid
    // margin: cell assignment and the edge coordinates carry a few ulp of the coordinate range
    if (kK < bound * 0.9999f - 1e-5f * g.cs) break;
  }
#pragma unroll
  for (int sl = 0; sl < S; ++sl)
    if (sl * 32 + lane < K) idx_out[((size_t)b * N + i) * K + sl * 32 + lane] = (int32_t)(unsigned)list[sl];
}

// ---- one THREAD per query (K a compile-time constant) ------------------------------------------
// Queries are taken in cell-sorted order, so the 32 lanes of a warp sit in the same or adjacent
// cells, walk nearly the same candidates and stay converged.  Each lane keeps its own sorted best
// list in registers (packed (key, index) words) and streams the candidates of its rings one at a
// time; an instruction now serves 32 queries instead of one, which is ~3x fewer issue slots per
// query than the warp-per-query kernel above.  Same rings, same stop test, same exact pre-filter.
// The kernel is bound by the serial compare-exchange chain of the sorted insertion (~85 instructions per accepted
// candidate, ~100 loop trips per warp): staging the candidates in shared memory bought 4 %, and two lanes per query
// with a bitonic merge of their lists per ring was slower (47 vs 43 us) - more parallelism, but more instructions.
// STAGE: the env's sorted candidates and cell starts are first copied into shared memory (coalesced), so that the
// dependent per-candidate loads of the loop below are LDS round trips (~30 cycles) instead of L1/L2 ones (the cell
// starts of 64 envs do not fit L1: every ring paid two ~300-cycle L2 loads per cell range).
template <int K, bool STAGE>
__global__ void __launch_bounds__(128)
knn_grid_tpq_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const GridHeader* __restrict__ header, int N, int C, int G, int32_t* __restrict__ idx_out,
                    const int* __restrict__ skip, int part, int nparts, int row_begin, int row_end, int Bs,
                    int32_t* __restrict__ idx_out2) {
  MACBF_SKIP_IF_SET(skip);
  const int b = blockIdx.y;                            // Bs > 0: grids b >= Bs are the envs one Euler step later -> idx_out2
  const int p0 = blockIdx.x * blockDim.x + threadIdx.x;
  const GridHeader g = header[b];
  const int* cs = cell_start + (size_t)b * (GRID_GMAX * GRID_GMAX + 1);
  const float4* pts = sorted + (size_t)b * C;
  if constexpr (STAGE) {
    extern __shared__ float4 s_pts[];                  // [C] candidates, then [G*G + 1] cell starts
    int* s_cs = reinterpret_cast<int*>(s_pts + C);
    for (int t = threadIdx.x; t < C; t += blockDim.x) s_pts[t] = __ldg(pts + t);
    for (int t = threadIdx.x; t <= G * G; t += blockDim.x) s_cs[t] = __ldg(cs + t);
    __syncthreads();
    pts = s_pts;
    cs = s_cs;
  }
  if (p0 >= C) return;
  const float4 me = pts[p0];
  const int i = __float_as_int(me.z);
  if (i >= N) return;                                  // obstacle points are candidates, not queries
  const float qx = me.x, qy = me.y;
  const int cx = cell_coord(qx, g.ox, g.inv_cs, G), cy = cell_coord(qy, g.oy, g.inv_cs, G);
  // row sharding: a query belongs to the part that owns its CELL (by the cell's start in the sorted order) - the order
  // of the candidates inside a cell is not reproducible between two builds of the grid, cell membership is
  if (nparts > 1 && (int)((long long)cs[cy * G + cx] * nparts / C) != part) return;
  if (row_end >= 0 && (i < row_begin || i >= row_end)) return;   // contiguous row range (macbf_step_forward_rows_*)

  u64 list[K];
#pragma unroll
  for (int e = 0; e < K; ++e) list[e] = ENTRY_INF;
  float vthr = __int_as_float(0x7f800000);

  for (int r = 0; r < G; ++r) {
    const int nranges = (r == 0) ? 1 : 4 * r;
    for (int t = 0; t < nranges; ++t) {
      int y, x0, x1;
      if (r == 0) { y = cy; x0 = cx; x1 = cx; }
      else if (t == 0) { y = cy - r; x0 = cx - r; x1 = cx + r; }
      else if (t == 1) { y = cy + r; x0 = cx - r; x1 = cx + r; }
      else {
        y = cy - r + 1 + ((t - 2) >> 1);
        x0 = x1 = ((t - 2) & 1) ? cx + r : cx - r;
      }
      x0 = max(x0, 0); x1 = min(x1, G - 1);
      if (y < 0 || y >= G || x0 > x1) continue;
      int p = cs[y * G + x0];
      const int pend = cs[y * G + x1 + 1];
      for (; p < pend; ++p) {
        const float4 c = pts[p];
        // core.py:238, op by op
        const float dx = xsub(qx, c.x), dy = xsub(qy, c.y);
        const float v = xadd(xadd(xmul(dx, dx), 1e-6f), xadd(xmul(dy, dy), 1e-6f));
        if (v < vthr) {
          u64 cand = ((u64)__float_as_uint(__fsqrt_rn(v)) << 32) | (unsigned)__float_as_int(c.z);
          if (cand < list[K - 1]) {                    // exact lexicographic test (key, index)
            list[K - 1] = cand;
#pragma unroll
            for (int e = K - 1; e > 0; --e) {
              const u64 a = list[e - 1], bb = list[e];
              const bool sw = bb < a;
              list[e - 1] = sw ? bb : a;
              list[e] = sw ? a : bb;
            }
            const float kK = __uint_as_float((unsigned)(list[K - 1] >> 32));
            const float kn = __uint_as_float(__float_as_uint(kK) + 1u);
            vthr = (kK < __int_as_float(0x7f800000)) ? __fmul_ru(kn, kn) : kK;
          }
        }
      }
    }
    const float kK = __uint_as_float((unsigned)(list[K - 1] >> 32));
    float bound = __int_as_float(0x7f800000);
    if (cx - r > 0) bound = fminf(bound, qx - (g.ox + (float)(cx - r) * g.cs));
    if (cx + r < G - 1) bound = fminf(bound, (g.ox + (float)(cx + r + 1) * g.cs) - qx);
    if (cy - r > 0) bound = fminf(bound, qy - (g.oy + (float)(cy - r) * g.cs));
    if (cy + r < G - 1) bound = fminf(bound, (g.oy + (float)(cy + r + 1) * g.cs) - qy);
    if (bound == __int_as_float(0x7f800000)) break;
    if (kK < bound * 0.9999f - 1e-5f * g.cs) break;
  }
  int32_t* o = (Bs > 0 && b >= Bs) ? idx_out2 + ((size_t)(b - Bs) * N + i) * K : idx_out + ((size_t)b * N + i) * K;
#pragma unroll
  for (int e = 0; e < K; ++e) o[e] = (int32_t)(unsigned)list[e];
}

template <int K>
static int launch_tpq(const float4* sorted, const int* cell_start, const GridHeader* header, int B, int N, int C, int G,
                      int32_t* idx, cudaStream_t st) {
  const size_t stage_bytes = (size_t)C * sizeof(float4) + (size_t)(G * G + 1) * sizeof(int);
  dim3 grid((C + 127) / 128, B);
  if (stage_bytes <= 96 * 1024) {                      // small envs: candidates + cell starts through shared memory
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(knn_grid_tpq_kernel<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    knn_grid_tpq_kernel<K, true><<<grid, 128, stage_bytes, st>>>(sorted, cell_start, header, N, C, G, idx, tl_skip_flag,
                                                                 tl_query_part.part, tl_query_part.nparts, tl_row_range.begin,
                                                                 tl_row_range.end, tl_graph_pair.B_split, tl_graph_pair.idx_next);
  } else {
    knn_grid_tpq_kernel<K, false><<<grid, 128, 0, st>>>(sorted, cell_start, header, N, C, G, idx, tl_skip_flag,
                                                        tl_query_part.part, tl_query_part.nparts, tl_row_range.begin,
                                                        tl_row_range.end, tl_graph_pair.B_split, tl_graph_pair.idx_next);
  }
  MACBF_CHECK_LAUNCH("knn_grid_tpq_kernel");
  return 0;
}

size_t neighbor_grid_workspace_bytes(int B, int N, int M) {
  const size_t C = (size_t)N + M;
  size_t bytes = (size_t)B * C * sizeof(float4) + (size_t)B * (GRID_GMAX * GRID_GMAX + 1) * sizeof(int) +
                 (size_t)B * sizeof(GridHeader) + 256;
  if (C >= (size_t)GRID_BIG_MIN_C)                      // multi-CTA build: scatter cursors + bounding boxes
    bytes += (size_t)B * (GRID_GMAX * GRID_GMAX + 1) * sizeof(int) + (size_t)B * 16 + 256;
  return bytes;
}

// true when the grid path applies (the caller falls back to the brute-force kernel otherwise)
bool neighbor_grid_applicable(int N, int M, int top_k) {
  const int C = N + M;
  return top_k <= 128 && C > top_k && C >= 256;
}

int neighbor_topk_grid(const float* states, const float* obstacles, int B, int N, int M, int K, int32_t* idx,
                       void* workspace, size_t workspace_bytes, cudaStream_t st) {
  const int C = N + M;
  MACBF_CHECK_ARG(workspace_bytes >= neighbor_grid_workspace_bytes(B, N, M), "neighbor_topk: workspace too small (%zu < %zu)",
                  workspace_bytes, neighbor_grid_workspace_bytes(B, N, M));
  MACBF_CHECK_ARG((((uintptr_t)workspace) & 15) == 0, "neighbor_topk: workspace must be 16-byte aligned");
  float4* sorted = (float4*)workspace;
  int* cell_start = (int*)((char*)workspace + (size_t)B * C * sizeof(float4));
  GridHeader* header = (GridHeader*)((char*)cell_start + (size_t)B * (GRID_GMAX * GRID_GMAX + 1) * sizeof(int));
  // p = candidates per cell.  Measured on 1024 agents x 64 envs, K = 12: p = 2 / 3 / 4 / 5 / 7 / 9 -> 64 / 59 / 63 / 65 /
  // 67 / 73 us: small cells with an occasional third ring beat big cells that always stop after ring 1.
  const float per_cell = fmaxf(3.0f, 0.25f * (float)K);
  int G = (int)floorf(sqrtf((float)C / per_cell));
  G = G < 1 ? 1 : (G > GRID_GMAX ? GRID_GMAX : G);
  const size_t build_smem = (size_t)(G * G + 1) * sizeof(int);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)((GRID_GMAX * GRID_GMAX + 1) * sizeof(int))));
  if (C >= GRID_BIG_MIN_C) {
    char* extra = (char*)header + (((size_t)B * sizeof(GridHeader) + 255) & ~(size_t)255);
    int* cursors = (int*)extra;
    unsigned* box = (unsigned*)(extra + (size_t)B * (GRID_GMAX * GRID_GMAX + 1) * sizeof(int));
    const dim3 gpts((C + GRID_BIG_PER_CTA - 1) / GRID_BIG_PER_CTA, B);
    grid_big_init_kernel<<<dim3(8, B), GRID_BIG_THREADS, 0, st>>>(box, cursors, G * G + 1, tl_skip_flag);
    MACBF_CHECK_LAUNCH("grid_big_init_kernel");
    grid_big_bbox_kernel<<<gpts, GRID_BIG_THREADS, 0, st>>>(states, obstacles, N, M, box, tl_skip_flag);
    MACBF_CHECK_LAUNCH("grid_big_bbox_kernel");
    grid_big_hist_kernel<<<gpts, GRID_BIG_THREADS, 0, st>>>(states, obstacles, N, M, G, box, cursors, header, tl_skip_flag);
    MACBF_CHECK_LAUNCH("grid_big_hist_kernel");
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(grid_big_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)((GRID_GMAX * GRID_GMAX + 1) * sizeof(int))));
    grid_big_scan_kernel<<<B, GRID_BUILD_THREADS, build_smem, st>>>(C, G, cursors, cell_start, tl_skip_flag);
    MACBF_CHECK_LAUNCH("grid_big_scan_kernel");
    grid_big_scatter_kernel<<<gpts, GRID_BIG_THREADS, 0, st>>>(states, obstacles, N, M, G, header, cursors, sorted, tl_skip_flag);
    MACBF_CHECK_LAUNCH("grid_big_scatter_kernel");
  } else {
    grid_build_kernel<<<B, GRID_BUILD_THREADS, build_smem, st>>>(states, obstacles, N, M, G, sorted, cell_start, header,
                                                                 tl_skip_flag, tl_graph_pair.B_split, tl_graph_pair.dt);
    MACBF_CHECK_LAUNCH("grid_build_kernel");
  }
  {
    switch (K) {   // thread-per-query instantiations for the common neighbour counts
      case 4: return launch_tpq<4>(sorted, cell_start, header, B, N, C, G, idx, st);
      case 8: return launch_tpq<8>(sorted, cell_start, header, B, N, C, G, idx, st);
      case 12: return launch_tpq<12>(sorted, cell_start, header, B, N, C, G, idx, st);
      case 16: return launch_tpq<16>(sorted, cell_start, header, B, N, C, G, idx, st);
      case 24: return launch_tpq<24>(sorted, cell_start, header, B, N, C, G, idx, st);
      case 32: return launch_tpq<32>(sorted, cell_start, header, B, N, C, G, idx, st);
      default: break;
    }
  }
  int q_begin, q_end;
  query_range(N, &q_begin, &q_end);
  if (q_end <= q_begin) return 0;
  dim3 grid((q_end - q_begin + GRID_QUERY_THREADS / 32 - 1) / (GRID_QUERY_THREADS / 32), B);
  if (K <= 32)
    knn_grid_kernel<1><<<grid, GRID_QUERY_THREADS, 0, st>>>(states, sorted, cell_start, header, N, C, G, K, idx, tl_skip_flag,
                                                            q_begin, q_end);
  else if (K <= 64)
    knn_grid_kernel<2><<<grid, GRID_QUERY_THREADS, 0, st>>>(states, sorted, cell_start, header, N, C, G, K, idx, tl_skip_flag,
                                                            q_begin, q_end);
  else
    knn_grid_kernel<4><<<grid, GRID_QUERY_THREADS, 0, st>>>(states, sorted, cell_start, header, N, C, G, K, idx, tl_skip_flag,
                                                            q_begin, q_end);
  MACBF_CHECK_LAUNCH("knn_grid_kernel");
  return 0;
}

// ---- both graphs of a step in ONE pair of launches -----------------------------------------------------------------
// graph(S') needs the positions of S' only, and p' = p + v dt is known before the controller has run: the grid build
// and the query kernel take 2B grids - env b from `states`, env B + b from the advanced positions - and write idx and
// idx_next.  One build + one query launch at the start of the step instead of two of each (the second pair used to
// run on a side stream under CBF(S), which it slowed down by about as much as it hid).
bool neighbor_grid_pair_applicable(int N, int M, int top_k) {
  const int C = N + M, K = C < top_k ? C : top_k;
  const bool tpq = K == 4 || K == 8 || K == 12 || K == 16 || K == 24 || K == 32;
  return neighbor_grid_applicable(N, M, top_k) && C < GRID_BIG_MIN_C && tpq && !rows_active() && tl_query_part.nparts == 1;
}
size_t neighbor_grid_pair_workspace_bytes(int B, int N, int M) { return neighbor_grid_workspace_bytes(2 * B, N, M); }

int neighbor_topk_grid_pair(const float* states, const float* obstacles, int B, int N, int M, int K, double dt, int32_t* idx,
                            int32_t* idx_next, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  MACBF_CHECK_ARG(neighbor_grid_pair_applicable(N, M, K), "neighbor_topk_grid_pair: not applicable to N=%d M=%d K=%d", N, M, K);
  tl_graph_pair = GraphPair{B, (float)dt, idx_next};
  const int rc = neighbor_topk_grid(states, obstacles, 2 * B, N, M, K, idx, workspace, workspace_bytes, st);
  tl_graph_pair = GraphPair{0, 0.f, nullptr};
  return rc;
}

}  // namespace macbf
