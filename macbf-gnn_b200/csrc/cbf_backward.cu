// K2b - backward of the fused CBF head (autograd of cbf.py:21-45, REPAIRED semantics): entry point and the
// deterministic reduction of the per-CTA gradient slices.  The kernels are in cbf_bwd_tc.cu (tensor cores).
#include "common.cuh"

namespace macbf {

// out[p] += sum over slices, in slice order (deterministic)
__global__ void reduce_partials_kernel(const float* __restrict__ partial, int nslices, int P,
                                       float* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  float s = 0.f;
  for (int c = 0; c < nslices; ++c) s += partial[(size_t)c * P + p];
  out[p] += s;
}

size_t cbf_backward_tc_scratch_bytes(int B, int N, int K);
int cbf_backward_tc(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, double,
                    double, float*, float*, void*, size_t, cudaStream_t);

size_t cbf_backward_scratch_bytes(int B, int N, int K) { return cbf_backward_tc_scratch_bytes(B, N, K); }

int cbf_backward(const float* states, const float* obstacles, const int32_t* idx, const float* params,
                 const float* grad_h, int B, int N, int M, int K, double dist_min, double obs_radius,
                 float* grad_params, float* grad_states, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "cbf_backward: B*N must be < 2^31");
  return cbf_backward_tc(states, obstacles, idx, params, grad_h, B, N, M, K, dist_min, obs_radius, grad_params,
                         grad_states, scratch, scratch_bytes, st);
}

}  // namespace macbf
