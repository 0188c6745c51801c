// K1 - neighbour graph: per agent the K candidates with the smallest key, never building the
// dense (N, N, C) pair tensor of cbf.py:23 / controller.py:33 / core.py:188.
//
// Replaces core.py:234-250 (remove_distant_agents): key (core.py:238)
//   d = sqrt((dx^2 + 1e-6) + (dy^2 + 1e-6)),  topk(-d, k) sorted by ascending d.
// Order contract: lexicographic (key asc, candidate index asc).
//
// One warp owns Q query agents.  Candidates of the env are staged through shared memory in
// tiles; each lane evaluates one candidate per step for all Q queries (plain, uncontracted
// fp32 ops, so the bits equal the reference's eager kernels).  The sorted best-K list of a
// query lives in registers, entry e in lane e%32 slot e/32.  A candidate reaches the (rare)
// insertion path only if its un-rooted key v is below RU(key_K^2): v >= key_K^2 implies
// sqrt_rn(v) >= key_K by monotonicity of correctly-rounded sqrt, so the filter never drops
// a candidate the exact (rooted) comparison would keep.  Candidates are visited in ascending
// index order, so "insert after all entries with key <= new key" is exactly the
// lexicographic order.
#include "common.cuh"

namespace macbf {

constexpr int KNN_THREADS = 256;
constexpr int KNN_WARPS = KNN_THREADS / 32;
constexpr int KNN_TILE = 2048;  // candidates per shared-memory tile (16 KB of float2)

template <int S>
struct TopList {
  float key[S];
  int idx[S];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      key[s] = __int_as_float(0x7f800000);
      idx[s] = 0x7fffffff;
    }
  }
  // key of entry e (warp-uniform e)
  __device__ __forceinline__ float key_at(int e) const {
    float v = key[0];
#pragma unroll
    for (int s = 1; s < S; ++s)
      if ((e >> 5) == s) v = key[s];
    return __shfl_sync(0xffffffffu, v, e & 31);
  }
  // stable insert of (kc, ic): after every entry with key <= kc
  __device__ __forceinline__ void insert(float kc, int ic, int lane) {
    int pos = 0;
#pragma unroll
    for (int s = 0; s < S; ++s) pos += __popc(__ballot_sync(0xffffffffu, key[s] <= kc));
#pragma unroll
    for (int s = S - 1; s >= 0; --s) {
      float pk = __shfl_up_sync(0xffffffffu, key[s], 1);
      int pi = __shfl_up_sync(0xffffffffu, idx[s], 1);
      if (s > 0) {
        float wk = __shfl_sync(0xffffffffu, key[s - 1], 31);
        int wi = __shfl_sync(0xffffffffu, idx[s - 1], 31);
        if (lane == 0) {
          pk = wk;
          pi = wi;
        }
      }
      int e = s * 32 + lane;
      if (e > pos) {
        key[s] = pk;
        idx[s] = pi;
      } else if (e == pos) {
        key[s] = kc;
        idx[s] = ic;
      }
    }
  }
};

template <int S, int Q>
__global__ void __launch_bounds__(KNN_THREADS)
knn_kernel(const float* __restrict__ states, const float* __restrict__ obstacles, int N, int M, int K,
           int32_t* __restrict__ idx_out, const int* __restrict__ skip, int q_begin, int q_end) {
  MACBF_SKIP_IF_SET(skip);
  __shared__ float2 tile[KNN_TILE];
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q0 = q_begin + (blockIdx.x * KNN_WARPS + warp) * Q;
  const int C = N + M;

  float qx[Q], qy[Q], vthr[Q];
  TopList<S> top[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    int i = min(q0 + q, N - 1);
    float4 s = __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + i);
    qx[q] = s.x;
    qy[q] = s.y;
    vthr[q] = __int_as_float(0x7f800000);
    top[q].init();
  }

  for (int c0 = 0; c0 < C; c0 += KNN_TILE) {
    const int cnt = min(KNN_TILE, C - c0);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += KNN_THREADS) {
      float4 s = load_candidate(states, obstacles, b, c0 + t, N, M);
      tile[t] = make_float2(s.x, s.y);
    }
    __syncthreads();
    for (int t0 = 0; t0 < cnt; t0 += 32) {
      const int t = t0 + lane;
      const bool valid = t < cnt;
      const float2 c = tile[valid ? t : 0];
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        // core.py:238, op by op (square, + 1e-6, sum of the two terms)
        float dx = xsub(qx[q], c.x), dy = xsub(qy[q], c.y);
        float v = xadd(xadd(xmul(dx, dx), 1e-6f), xadd(xmul(dy, dy), 1e-6f));
        unsigned pass = __ballot_sync(0xffffffffu, valid && (v < vthr[q]));
        while (pass) {
          const int src = __ffs(pass) - 1;
          pass &= pass - 1;
          const float vc = __shfl_sync(0xffffffffu, v, src);
          const float kc = __fsqrt_rn(vc);
          const float kK = top[q].key_at(K - 1);
          if (kc < kK) {  // warp-uniform
            top[q].insert(kc, c0 + t0 + src, lane);
            const float nk = top[q].key_at(K - 1);
            vthr[q] = __fmul_ru(nk, nk);
            // candidates later in this batch that no longer pass are dropped by the exact test
          }
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int i = q0 + q;
    if (i < q_end) {
      int32_t* o = idx_out + ((size_t)b * N + i) * K;
#pragma unroll
      for (int s = 0; s < S; ++s) {
        int e = s * 32 + lane;
        if (e < K) o[e] = top[q].idx[s];
      }
    }
  }
}

// Dense variant for the reference's own call shape (core.py:234-250 takes the materialised
// (N, C, F) pair tensor): keys come from x[i, j, 0:2], one warp per row, same list and order.
template <int S>
__global__ void __launch_bounds__(KNN_THREADS)
knn_dense_kernel(const float* __restrict__ x, int N, int C, int F, int K, int32_t* __restrict__ idx_out) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * KNN_WARPS + (threadIdx.x >> 5);
  if (i >= N) return;
  const float* row = x + (size_t)i * C * F;
  TopList<S> top;
  top.init();
  float vthr = __int_as_float(0x7f800000);
  for (int t0 = 0; t0 < C; t0 += 32) {
    const int t = t0 + lane;
    const bool valid = t < C;
    float v = 0.f;
    if (valid) {
      const float dx = __ldg(row + (size_t)t * F), dy = __ldg(row + (size_t)t * F + 1);
      v = xadd(xadd(xmul(dx, dx), 1e-6f), xadd(xmul(dy, dy), 1e-6f));   // core.py:238
    }
    unsigned pass = __ballot_sync(0xffffffffu, valid && (v < vthr));
    while (pass) {
      const int src = __ffs(pass) - 1;
      pass &= pass - 1;
      const float kc = __fsqrt_rn(__shfl_sync(0xffffffffu, v, src));
      if (kc < top.key_at(K - 1)) {
        top.insert(kc, t0 + src, lane);
        const float nk = top.key_at(K - 1);
        vthr = __fmul_ru(nk, nk);
      }
    }
  }
  int32_t* o = idx_out + (size_t)i * K;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int e = s * 32 + lane;
    if (e < K) o[e] = top.idx[s];
  }
}

int dense_topk(const float* x, int N, int C, int F, int K, int32_t* idx, cudaStream_t st) {
  const int grid = (N + KNN_WARPS - 1) / KNN_WARPS;
  if (K <= 32) knn_dense_kernel<1><<<grid, KNN_THREADS, 0, st>>>(x, N, C, F, K, idx);
  else if (K <= 64) knn_dense_kernel<2><<<grid, KNN_THREADS, 0, st>>>(x, N, C, F, K, idx);
  else knn_dense_kernel<4><<<grid, KNN_THREADS, 0, st>>>(x, N, C, F, K, idx);
  MACBF_CHECK_LAUNCH("knn_dense_kernel");
  return 0;
}

// N + M <= top_k: every candidate, in index order (core.py:236-237, repair F3).  Only the rows [q_begin, q_end) of
// every env are written, so the parts of macbf_neighbor_topk_part stay a disjoint cover here too.
__global__ void knn_all_kernel(long long total, int N, int C, int q_begin, int q_end, int32_t* __restrict__ idx_out) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int row = (int)((t / C) % N);
  if (row >= q_begin && row < q_end) idx_out[t] = (int32_t)(t % C);
}

template <int S, int Q>
static int launch_knn(const float* states, const float* obstacles, int B, int N, int M, int K,
                      int32_t* idx, cudaStream_t st) {
  int q_begin, q_end;
  query_range(N, &q_begin, &q_end);
  if (q_end <= q_begin) return 0;
  dim3 grid((q_end - q_begin + KNN_WARPS * Q - 1) / (KNN_WARPS * Q), B);
  knn_kernel<S, Q><<<grid, KNN_THREADS, 0, st>>>(states, obstacles, N, M, K, idx, tl_skip_flag, q_begin, q_end);
  MACBF_CHECK_LAUNCH("knn_kernel");
  return 0;
}

int neighbor_topk(const float* states, const float* obstacles, int B, int N, int M, int top_k,
                  int32_t* idx, cudaStream_t st) {
  const int C = N + M;
  if (C <= top_k) {
    const long long total = (long long)B * N * C;
    int q_begin, q_end;
    query_range(N, &q_begin, &q_end);
    knn_all_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(total, N, C, q_begin, q_end, idx);
    MACBF_CHECK_LAUNCH("knn_all_kernel");
    return 0;
  }
  const int K = top_k;
  // more queries per warp amortise the shared-memory candidate reads; fewer keep small
  // problems spread over the SMs
  const long long queries = (long long)B * N;
  const bool big = queries >= (long long)NUM_SMS * KNN_WARPS * 4 * 2;
  if (K <= 32) return big ? launch_knn<1, 4>(states, obstacles, B, N, M, K, idx, st)
                          : launch_knn<1, 1>(states, obstacles, B, N, M, K, idx, st);
  if (K <= 64) return big ? launch_knn<2, 2>(states, obstacles, B, N, M, K, idx, st)
                          : launch_knn<2, 1>(states, obstacles, B, N, M, K, idx, st);
  return big ? launch_knn<4, 2>(states, obstacles, B, N, M, K, idx, st)
             : launch_knn<4, 1>(states, obstacles, B, N, M, K, idx, st);
}

}  // namespace macbf
