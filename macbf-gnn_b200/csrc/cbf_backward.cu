// K2b - backward of the fused CBF head (autograd of cbf.py:21-45, REPAIRED semantics): entry point and the
// deterministic reduction of the per-CTA gradient slices.  The kernels are in cbf_bwd_tc.cu (tensor cores).
#include "common.cuh"

namespace macbf {

// out[p] += sum over slices, in a FIXED order (deterministic): 64 parameters x 8 slice groups per CTA; group g adds the
// slices g, g + 8, ... in order, then the 8 group sums are added in order.  (One thread per parameter walking all
// slices serially took 20 us for 10 MB: 67 CTAs of dependent loads.)
constexpr int RP_PARAMS = 64, RP_GROUPS = 8;
__global__ void __launch_bounds__(RP_PARAMS * RP_GROUPS)
reduce_partials_kernel(const float* __restrict__ partial, int nslices, int P, float* __restrict__ out) {
  __shared__ float red[RP_GROUPS][RP_PARAMS];
  const int pl = threadIdx.x % RP_PARAMS, g = threadIdx.x / RP_PARAMS;
  const int p = blockIdx.x * RP_PARAMS + pl;
  float s = 0.f;
  if (p < P)
    for (int c = g; c < nslices; c += RP_GROUPS) s += partial[(size_t)c * P + p];
  red[g][pl] = s;
  __syncthreads();
  if (g == 0 && p < P) {
    float t = red[0][pl];
#pragma unroll
    for (int q = 1; q < RP_GROUPS; ++q) t += red[q][pl];
    out[p] += t;
  }
}
int reduce_partials(const float* partial, int nslices, int P, float* out, cudaStream_t st) {
  reduce_partials_kernel<<<(P + RP_PARAMS - 1) / RP_PARAMS, RP_PARAMS * RP_GROUPS, 0, st>>>(partial, nslices, P, out);
  MACBF_CHECK_LAUNCH("reduce_partials_kernel");
  return 0;
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
