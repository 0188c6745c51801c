// Evaluation-time dense time-to-collision check (core.py:212-231, ttc_dangerous_mask_np; the safety-rate
// computation of train.py:74-75): for every ordered pair (i, j) of one env, is j within DIST_MIN_CHECK of i or
// on a collision course within TIME_TO_COLLISION_CHECK?
//
// The reference evaluates it in numpy on a float32 state array, and numpy's type promotion makes the
// arithmetic MIXED: s_i - s_j in float32; x, y become float64 when the float64 identity matrix is added;
// alpha = vx^2 + vy^2, 4*alpha and 2*alpha*ttc stay float32 (python scalars are weak); everything they meet
// is float64.  The kernel follows that promotion op by op (no contraction), so the mask is bit-exact against
// the reference's own output (tests/golden/ref_*.npz: ttc_np).
//
// One thread per row i, candidates j staged through shared memory; outputs the per-row count of dangerous
// j (safety_rate_i = 1 - count_i / N) and, optionally, the dense (N, N) mask.
#include "common.cuh"

namespace macbf {

constexpr int TD_THREADS = 128;
constexpr int TD_TILE = 1024;

__global__ void __launch_bounds__(TD_THREADS)
ttc_dense_kernel(const float* __restrict__ states, int N, double r2, float ttc32, int32_t* __restrict__ counts,
                 uint8_t* __restrict__ mask) {
  __shared__ float4 tile[TD_TILE];
  const int b = blockIdx.y;
  const int i = blockIdx.x * TD_THREADS + threadIdx.x;
  const float4* sb = reinterpret_cast<const float4*>(states) + (size_t)b * N;
  const float4 si = sb[min(i, N - 1)];
  int count = 0;
  for (int j0 = 0; j0 < N; j0 += TD_TILE) {
    const int cnt = min(TD_TILE, N - j0);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += TD_THREADS) tile[t] = __ldg(sb + j0 + t);
    __syncthreads();
    if (i >= N) continue;
    uint8_t* mrow = mask ? mask + ((size_t)b * N + i) * N + j0 : nullptr;
    for (int t = 0; t < cnt; ++t) {
      const float4 sj = tile[t];
      const float dx = xsub(si.x, sj.x), dy = xsub(si.y, sj.y), dvx = xsub(si.z, sj.z), dvy = xsub(si.w, sj.w);
      const double eye = (j0 + t == i) ? 1.0 : 0.0;
      const double x = __dadd_rn((double)dx, eye), y = __dadd_rn((double)dy, eye);
      const double vx = (double)dvx, vy = (double)dvy;
      const float alpha = xadd(xmul(dvx, dvx), xmul(dvy, dvy));                                  // float32
      const double beta = __dmul_rn(2.0, __dadd_rn(__dmul_rn(x, vx), __dmul_rn(y, vy)));
      const double gamma = __dsub_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), r2);
      const double disc = __dsub_rn(__dmul_rn(beta, beta), __dmul_rn((double)xmul(4.f, alpha), gamma));
      const double at = (double)xmul(xmul(2.f, alpha), ttc32);                                   // float32 product
      const bool two_pos = (disc > 0.0) && (gamma > 0.0) && (beta < 0.0);
      const double bp = __dadd_rn(beta, at);
      const bool root = (__dsub_rn(-beta, at) < 0.0) || (__dmul_rn(bp, bp) < disc);
      const bool dang = (gamma < 0.0) || (two_pos && root);
      count += dang ? 1 : 0;
      if (mrow) mrow[t] = dang ? 1 : 0;
    }
  }
  if (i < N && counts) counts[(size_t)b * N + i] = count;
}

int ttc_dense(const float* states, int B, int N, double r2, double ttc, int32_t* counts, uint8_t* mask, cudaStream_t st) {
  dim3 grid((N + TD_THREADS - 1) / TD_THREADS, B);
  ttc_dense_kernel<<<grid, TD_THREADS, 0, st>>>(states, N, r2, (float)ttc, counts, mask);
  MACBF_CHECK_LAUNCH("ttc_dense_kernel");
  return 0;
}

}  // namespace macbf
