// K3 - controller (controller.py:31-63) + double-integrator Euler step (core.py:74-86, train.py:70).
//   kernel A (edges):  features [dx,dy,dvx,dvy,eye] -> 5->64->128 MLP -> * mask -> max over the K
//                      slots of each agent; only the pooled (agents,128) vector leaves the SM.
//   kernel B (agents): [pool, p-g, v] (132) -> 64 -> 128 -> 64 -> 4, gains 2*sigmoid+0.2, action,
//                      s' = s + [v, u] * dt.
// FP32 FFMA path; parameters resident in shared memory of persistent CTAs.
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace macbf {

namespace ce_smem {  // edge kernel
constexpr int R = 128;
constexpr int WS1 = 8, WS2 = 68;
constexpr int oW1 = 0;
constexpr int oB1 = oW1 + 64 * WS1;
constexpr int oW2 = oB1 + 64;
constexpr int oB2 = oW2 + 128 * WS2;
constexpr int oMask = oB2 + 128;
constexpr int oA0 = oMask + R;
constexpr int oA1 = oA0 + 8 * R;
constexpr int oA2 = oA1 + 64 * R;
constexpr int RS2 = 132;  // padded so the pooling reads (lanes along channels) spread over banks
constexpr int total = oA2 + 128 * RS2;
}  // namespace ce_smem

template <bool kLiteral>
__global__ void __launch_bounds__(MLP_THREADS, 1)
ctrl_edge_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                 const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int K,
                 long long agents_total, int TA, float obs_radius, float* __restrict__ pool) {
  using namespace ce_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* mask = smem + oMask;
  float* A0 = smem + oA0; float* A1 = smem + oA1; float* A2 = smem + oA2;

  load_weight<64, 5, 8, WS1>(params + CT_EW1, W1);
  load_weight<128, 64, 64, WS2>(params + CT_EW2, W2);
  for (int t = threadIdx.x; t < 64; t += blockDim.x) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 128; t += blockDim.x) B2[t] = __ldg(params + CT_EB2 + t);

  const long long ntiles = (agents_total + TA - 1) / TA;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long a0 = tile * TA;
    const int na = (int)min((long long)TA, agents_total - a0);
    const int rows = na * K;
    __syncthreads();
    if (threadIdx.x < R) {
      const int r = threadIdx.x;
      float f[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (r < rows) {
        const long long ga = a0 + r / K;
        const int b = (int)(ga / N), i = (int)(ga % N);
        const int j = __ldg(idx + a0 * K + r);
        const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        const float4 sj = load_candidate(states, obstacles, b, j, N, M);
        const EdgeGeom g = edge_geom(si, sj, i, j);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        mk = (ctrl_dist(g.dx, g.dy) < obs_radius) ? 1.f : 0.f;  // controller.py:38-39
      }
      mask[r] = mk;
      if (kLiteral) {
#pragma unroll
        for (int c = 0; c < 5; ++c) A2[r * 5 + c] = f[c];
      } else {
#pragma unroll
        for (int c = 0; c < 5; ++c) A0[c * R + r] = f[c];
      }
      A0[5 * R + r] = 0.f; A0[6 * R + r] = 0.f; A0[7 * R + r] = 0.f;
    }
    if (kLiteral) {  // controller.py:43: (N,K,5).reshape(N,5,K)
      __syncthreads();
      if (threadIdx.x < R) {
        const int r = threadIdx.x;
        const int a = r / K, kp = r % K;
#pragma unroll
        for (int c = 0; c < 5; ++c) A0[c * R + r] = (r < rows) ? A2[a * 5 * K + c * K + kp] : 0.f;
      }
    }
    __syncthreads();
    layer_nk<R, 8, 64, R, R, WS1, true>(A0, W1, B1, A1);
    __syncthreads();
    layer_nk<R, 64, 128, R, RS2, WS2, true>(A1, W2, B2, A2);
    __syncthreads();
    // ---- mask + pooling over the K slots (controller.py:46) ----
    for (int item = threadIdx.x; item < na * 128; item += MLP_THREADS) {
      const int c = item & 127, a = item >> 7;
      const float* row = A2 + c * RS2 + a * K;
      const float* mrow = mask + a * K;
      if (kLiteral) {
        // torch.max(dim)[1]: index of the first maximum
        float best = row[0] * mrow[0];
        int bi = 0;
        for (int k = 1; k < K; ++k) {
          float v = row[k] * mrow[k];
          if (v > best) { best = v; bi = k; }
        }
        pool[(a0 + a) * 128 + c] = (float)bi;
      } else {
        const int rot = (c >> 3) & 3;  // lanes c and c+8 would share a bank: start them apart
        float best = 0.f;              // relu(.) * mask >= 0 and slot 0 (self) is always unmasked
        for (int k = 0; k < K; ++k) {
          int kk = k + rot; kk = kk >= K ? kk - K : kk;
          best = fmaxf(best, row[kk] * mrow[kk]);
        }
        pool[(a0 + a) * 128 + c] = best;
      }
    }
  }
}

int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk = nullptr);
int ctrl_node_tc(const float* states, const float* goals, const float* pool, const float* params, long long agents,
                 double dt, float* u, float* s_next, cudaStream_t st);

// The REPAIRED edge net and the node net run on the tensor cores (ctrl_tc.cu, ctrl_node_tc.cu).  LITERAL keeps the FP32
// edge kernel above: its feature scramble (controller.py:43) crosses rows and it pools arg-max INDICES
// (controller.py:46), which the tensor-core tile layout does not express; its node net is the tensor-core one
// (LITERAL only changes what `pool` holds).
bool controller_forward_is_tensor_core(int semantics) { return semantics == MACBF_REPAIRED; }

int controller_forward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                       const float* params, int B, int N, int M, int K, int semantics, double obs_radius,
                       double dt, float* pool, float* u, float* s_next, cudaStream_t st) {
  const long long agents = (long long)B * N;
  if (semantics == MACBF_REPAIRED) {
    int rc = ctrl_edge_tc(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st, tl_save_winners);
    if (rc) return rc;
  } else {
    const int TA = max(1, ce_smem::R / K);
    const long long ntiles = (agents + TA - 1) / TA;
    const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
    const size_t smem = ce_smem::total * sizeof(float);
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctrl_edge_kernel<true><<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                           (float)obs_radius, pool);
    MACBF_CHECK_LAUNCH("ctrl_edge_kernel");
  }
  return ctrl_node_tc(states, goals, pool, params, agents, dt, u, s_next, st);
}

}  // namespace macbf
