// K2 - fused CBF head (cbf.py:21-45): edge features -> 6->64->128->64->1 MLP -> h * mask.
// FP32 FFMA path.  A persistent CTA keeps the 17,089 parameters in shared memory and walks
// tiles of TA = floor(128 / K) agents (TA*K <= 128 edge rows); the per-edge feature vectors and
// all hidden activations live in shared memory only - per edge the kernel reads 4 B of idx plus
// two (L2-resident) states and writes 4 B of h.
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace macbf {

namespace cbf_smem {
constexpr int R = 128;
constexpr int WS1 = 8, WS2 = 68, WS3 = 132;
constexpr int oW1 = 0;
constexpr int oB1 = oW1 + 64 * WS1;
constexpr int oW2 = oB1 + 64;
constexpr int oB2 = oW2 + 128 * WS2;
constexpr int oW3 = oB2 + 128;
constexpr int oB3 = oW3 + 64 * WS3;
constexpr int oW4 = oB3 + 64;
constexpr int oB4 = oW4 + 64;
constexpr int oMask = oB4 + 4;
constexpr int oA0 = oMask + R;
constexpr int oA1 = oA0 + 8 * R;
constexpr int oA2 = oA1 + 64 * R;
constexpr int total = oA2 + 128 * R;
}  // namespace cbf_smem

template <bool kLiteral>
__global__ void __launch_bounds__(MLP_THREADS, 1)
cbf_forward_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                   const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M,
                   int K, long long agents_total, int TA, float dist_min, float obs_radius,
                   float* __restrict__ h) {
  using namespace cbf_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* W3 = smem + oW3; float* B3 = smem + oB3;
  float* W4 = smem + oW4; float* B4 = smem + oB4;
  float* mask = smem + oMask;
  float* A0 = smem + oA0; float* A1 = smem + oA1; float* A2 = smem + oA2;

  load_weight<64, 6, 8, WS1>(params + CBF_W1, W1);
  load_weight<128, 64, 64, WS2>(params + CBF_W2, W2);
  load_weight<64, 128, 128, WS3>(params + CBF_W3, W3);
  for (int t = threadIdx.x; t < 64; t += blockDim.x) {
    B1[t] = __ldg(params + CBF_B1 + t);
    B3[t] = __ldg(params + CBF_B3 + t);
    W4[t] = __ldg(params + CBF_W4 + t);
  }
  for (int t = threadIdx.x; t < 128; t += blockDim.x) B2[t] = __ldg(params + CBF_B2 + t);
  if (threadIdx.x == 0) B4[0] = __ldg(params + CBF_B4);

  const long long ntiles = (agents_total + TA - 1) / TA;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long a0 = tile * TA;
    const int na = (int)min((long long)TA, agents_total - a0);
    const int rows = na * K;
    __syncthreads();  // previous tile's readers of A0/mask are done; weights are visible
    // ---- edge features (cbf.py:23-37) ----
    if (threadIdx.x < R) {
      const int r = threadIdx.x;
      float f[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (r < rows) {
        const long long ga = a0 + r / K;
        const int b = (int)(ga / N), i = (int)(ga % N);
        const int j = __ldg(idx + a0 * K + r);
        const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        const float4 sj = load_candidate(states, obstacles, b, j, N, M);
        const EdgeGeom g = edge_geom(si, sj, i, j);
        const float d = cbf_dist(g.dx, g.dy);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        f[5] = xsub(d, dist_min);
        mk = (d <= obs_radius) ? 1.f : 0.f;
      }
      mask[r] = mk;
      if (kLiteral) {
        // cbf.py:40 reshapes (N,K,6) to (N,6,K): stage the flat per-agent vector, re-read below
#pragma unroll
        for (int c = 0; c < 6; ++c) A2[r * 6 + c] = f[c];
      } else {
#pragma unroll
        for (int c = 0; c < 6; ++c) A0[c * R + r] = f[c];
      }
      A0[6 * R + r] = 0.f;
      A0[7 * R + r] = 0.f;
    }
    if (kLiteral) {
      __syncthreads();
      if (threadIdx.x < R) {
        const int r = threadIdx.x;
        const int a = r / K, kp = r % K;
#pragma unroll
        for (int c = 0; c < 6; ++c) A0[c * R + r] = (r < rows) ? A2[a * 6 * K + c * K + kp] : 0.f;
      }
    }
    __syncthreads();
    layer_nk<R, 8, 64, R, R, WS1, true>(A0, W1, B1, A1);
    __syncthreads();
    layer_nk<R, 64, 128, R, R, WS2, true>(A1, W2, B2, A2);
    __syncthreads();
    layer_nk<R, 128, 64, R, R, WS3, true>(A2, W3, B3, A1);
    __syncthreads();
    // ---- 64 -> 1 and mask (cbf.py:43) ----
    {
      const int r = threadIdx.x & (R - 1), half = threadIdx.x >> 7;
      float s = 0.f;
#pragma unroll 8
      for (int c = 0; c < 32; ++c) s = fmaf(W4[half * 32 + c], A1[(half * 32 + c) * R + r], s);
      A0[half * R + r] = s;
    }
    __syncthreads();
    if (threadIdx.x < rows) {
      const int r = threadIdx.x;
      h[a0 * K + r] = (A0[r] + A0[R + r] + B4[0]) * mask[r];
    }
  }
}

int cbf_forward_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                   int M, int K, double dist_min, double obs_radius, float* h, cudaStream_t st);

int cbf_forward(const float* states, const float* obstacles, const int32_t* idx, const float* params,
                int B, int N, int M, int K, int semantics, double dist_min, double obs_radius,
                float* h, cudaStream_t st) {
  // REPAIRED runs on the tensor cores (cbf_tc.cu).  The FP32 kernel above serves LITERAL only: its feature scramble
  // (cbf.py:40) crosses rows, which the tensor-core tile layout does not express.
  if (semantics == MACBF_REPAIRED)
    return cbf_forward_tc(states, obstacles, idx, params, B, N, M, K, dist_min, obs_radius, h, st);
  const long long agents = (long long)B * N;
  const int TA = max(1, cbf_smem::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  const size_t smem = cbf_smem::total * sizeof(float);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_forward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cbf_forward_kernel<true><<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                           (float)dist_min, (float)obs_radius, h);
  MACBF_CHECK_LAUNCH("cbf_forward_kernel");
  return 0;
}

}  // namespace macbf
