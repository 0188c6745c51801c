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

namespace cn_smem {  // node kernel
constexpr int R = 128;
constexpr int KP1 = 136;              // 132 inputs padded to a multiple of 4
constexpr int WS1 = 140, WS2 = 68, WS3 = 132, WS4 = 68;
constexpr int RS0 = 132;
constexpr int oW1 = 0;
constexpr int oB1 = oW1 + 64 * WS1;
constexpr int oW2 = oB1 + 64;
constexpr int oB2 = oW2 + 128 * WS2;
constexpr int oW3 = oB2 + 128;
constexpr int oB3 = oW3 + 64 * WS3;
constexpr int oW4 = oB3 + 64;
constexpr int oB4 = oW4 + 4 * WS4;
constexpr int oX = oB4 + 4;            // A0 [136][132], later A2 [128][128]
constexpr int oY = oX + KP1 * RS0;     // A1 [64][128], later A3 [64][128]
constexpr int total = oY + 64 * R;
}  // namespace cn_smem

__global__ void __launch_bounds__(MLP_THREADS, 1)
ctrl_node_kernel(const float* __restrict__ states, const float* __restrict__ goals,
                 const float* __restrict__ pool, const float* __restrict__ params, long long agents_total,
                 float dt, float* __restrict__ u, float* __restrict__ s_next) {
  using namespace cn_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* W3 = smem + oW3; float* B3 = smem + oB3;
  float* W4 = smem + oW4; float* B4 = smem + oB4;
  float* X = smem + oX; float* Y = smem + oY;

  load_weight<64, 132, KP1, WS1>(params + CT_NW1, W1);
  load_weight<128, 64, 64, WS2>(params + CT_NW2, W2);
  load_weight<64, 128, 128, WS3>(params + CT_NW3, W3);
  load_weight<4, 64, 64, WS4>(params + CT_NW4, W4);
  for (int t = threadIdx.x; t < 64; t += blockDim.x) {
    B1[t] = __ldg(params + CT_NB1 + t);
    B3[t] = __ldg(params + CT_NB3 + t);
  }
  for (int t = threadIdx.x; t < 128; t += blockDim.x) B2[t] = __ldg(params + CT_NB2 + t);
  if (threadIdx.x < 4) B4[threadIdx.x] = __ldg(params + CT_NB4 + threadIdx.x);

  const long long ntiles = (agents_total + R - 1) / R;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long a0 = tile * R;
    const int rows = (int)min((long long)R, agents_total - a0);
    __syncthreads();
    // ---- A0 = [pool(128), p - g, v, 0-pad]^T  (controller.py:47) ----
    {
      const int rr = lane & 7, cc = lane >> 3;  // 8 rows x 4 float4 chunks per warp step
      for (int it = warp; it < 16 * 8; it += MLP_THREADS / 32) {
        const int r = (it >> 3) * 8 + rr, c4 = (it & 7) * 4 + cc;  // channel chunk of 4
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < rows) v = __ldg(reinterpret_cast<const float4*>(pool + (a0 + r) * 128) + c4);
        X[(4 * c4 + 0) * RS0 + r] = v.x; X[(4 * c4 + 1) * RS0 + r] = v.y;
        X[(4 * c4 + 2) * RS0 + r] = v.z; X[(4 * c4 + 3) * RS0 + r] = v.w;
      }
      if (threadIdx.x < R) {
        const int r = threadIdx.x;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        float2 g = make_float2(0.f, 0.f);
        if (r < rows) {
          s = __ldg(reinterpret_cast<const float4*>(states) + a0 + r);
          g = __ldg(reinterpret_cast<const float2*>(goals) + a0 + r);
        }
        X[128 * RS0 + r] = xsub(s.x, g.x); X[129 * RS0 + r] = xsub(s.y, g.y);
        X[130 * RS0 + r] = s.z; X[131 * RS0 + r] = s.w;
        X[132 * RS0 + r] = 0.f; X[133 * RS0 + r] = 0.f; X[134 * RS0 + r] = 0.f; X[135 * RS0 + r] = 0.f;
      }
    }
    __syncthreads();
    layer_nk<R, KP1, 64, RS0, R, WS1, true>(X, W1, B1, Y);
    __syncthreads();
    layer_nk<R, 64, 128, R, R, WS2, true>(Y, W2, B2, X);
    __syncthreads();
    layer_nk<R, 128, 64, R, R, WS3, true>(X, W3, B3, Y);
    __syncthreads();
    // ---- 64 -> 4, gains, action, Euler step (controller.py:50-61, train.py:70) ----
    if (threadIdx.x < rows) {
      const int r = threadIdx.x;
      float o[4] = {B4[0], B4[1], B4[2], B4[3]};
#pragma unroll 4
      for (int c = 0; c < 64; ++c) {
        const float a = Y[c * R + r];
#pragma unroll
        for (int q = 0; q < 4; ++q) o[q] = fmaf(W4[q * WS4 + c], a, o[q]);
      }
      float kk[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) kk[q] = 2.0f * (1.0f / (1.0f + expf(-o[q]))) + 0.2f;
      const float4 s = __ldg(reinterpret_cast<const float4*>(states) + a0 + r);
      const float2 g = __ldg(reinterpret_cast<const float2*>(goals) + a0 + r);
      const float ex = xsub(s.x, g.x), ey = xsub(s.y, g.y);
      const float ax = xadd(xmul(ex, -kk[0]), xmul(s.z, -kk[1]));
      const float ay = xadd(xmul(ey, -kk[2]), xmul(s.w, -kk[3]));
      reinterpret_cast<float2*>(u)[a0 + r] = make_float2(ax, ay);
      if (s_next != nullptr) {
        float4 n;
        n.x = xadd(s.x, xmul(s.z, dt)); n.y = xadd(s.y, xmul(s.w, dt));
        n.z = xadd(s.z, xmul(ax, dt)); n.w = xadd(s.w, xmul(ay, dt));
        reinterpret_cast<float4*>(s_next)[a0 + r] = n;
      }
    }
  }
}

int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk = nullptr);
int ctrl_node_tc(const float* states, const float* goals, const float* pool, const float* params, long long agents,
                 double dt, float* u, float* s_next, cudaStream_t st);

// REPAIRED edge net runs on the tensor cores (ctrl_tc.cu); MACBF_CTRL_FFMA=1 forces the FP32-FFMA
// kernel for A/B checks.  LITERAL always uses the FFMA kernel (row-crossing feature scramble).
static bool ctrl_force_ffma() {
  static const bool v = [] {
    const char* e = getenv("MACBF_CTRL_FFMA");
    return e != nullptr && e[0] == '1';
  }();
  return v;
}

bool controller_forward_is_tensor_core(int semantics) { return semantics == MACBF_REPAIRED && !ctrl_force_ffma(); }

// edge kernel only (REPAIRED): rebuilds the pooled features for the backward pass
int controller_forward_pool_only(const float* states, const float* obstacles, const int32_t* idx,
                                 const float* params, int B, int N, int M, int K, double obs_radius, float* pool,
                                 cudaStream_t st) {
  if (!ctrl_force_ffma()) return ctrl_edge_tc(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st);
  const long long agents = (long long)B * N;
  const int TA = max(1, ce_smem::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  const size_t smem = ce_smem::total * sizeof(float);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ctrl_edge_kernel<false><<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                          (float)obs_radius, pool);
  MACBF_CHECK_LAUNCH("ctrl_edge_kernel");
  return 0;
}

int controller_forward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                       const float* params, int B, int N, int M, int K, int semantics, double obs_radius,
                       double dt, float* pool, float* u, float* s_next, cudaStream_t st) {
  const long long agents = (long long)B * N;
  if (semantics == MACBF_REPAIRED && !ctrl_force_ffma()) {
    int rc = ctrl_edge_tc(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st, tl_save_winners);
    if (rc) return rc;
  } else {
    const int TA = max(1, ce_smem::R / K);
    const long long ntiles = (agents + TA - 1) / TA;
    const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
    const size_t smem = ce_smem::total * sizeof(float);
    if (semantics == MACBF_LITERAL) {
      MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      ctrl_edge_kernel<true><<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                             (float)obs_radius, pool);
    } else {
      MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      ctrl_edge_kernel<false><<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                              (float)obs_radius, pool);
    }
    MACBF_CHECK_LAUNCH("ctrl_edge_kernel");
  }
  if (!ctrl_force_ffma()) {
    // node net on the tensor cores (either semantics: LITERAL only changes what `pool` holds)
    return ctrl_node_tc(states, goals, pool, params, agents, dt, u, s_next, st);
  } else {
    const long long ntiles = (agents + cn_smem::R - 1) / cn_smem::R;
    const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
    const size_t smem = cn_smem::total * sizeof(float);
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_node_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctrl_node_kernel<<<grid, MLP_THREADS, smem, st>>>(states, goals, pool, params, agents, (float)dt, u, s_next);
    MACBF_CHECK_LAUNCH("ctrl_node_kernel");
  }
  return 0;
}

}  // namespace macbf
