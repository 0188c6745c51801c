// K2b - backward of the fused CBF head (autograd of cbf.py:21-45, REPAIRED semantics).
//
// Nothing is saved by the forward: a tile's activations are recomputed from states + idx in
// shared memory, then the chain rule runs layer by layer in place (delta_l overwrites a_l).
// Weight gradients of a CTA accumulate in its private slice of a global partial buffer (flat
// state_dict layout, L2-resident); a second kernel adds the slices in CTA order, so the result
// does not depend on scheduling.  d/dstates goes through the edge features
// [dx,dy,dvx,dvy,.,d-r] (indices and the radius mask are constants of the backward pass, as in
// eager autograd): +grad to agent i, -grad to neighbour j (fp32 atomics; obstacles get none).
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace macbf {

namespace cbfb_smem {
constexpr int R = 128;
constexpr int RS = 132;  // == 4 (mod 32): conflict-free LDS.128 across channel rows in grad_w
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
constexpr int oD4 = oMask + R;
constexpr int oGeo = oD4 + R;          // dx, dy, d per row
constexpr int oIJ = oGeo + 3 * R;      // global agent id of i, of j (or -1), as ints
constexpr int oA0 = oIJ + 2 * R;
constexpr int oD0 = oA0 + 8 * RS;
constexpr int oA1 = oD0 + 8 * RS;
constexpr int oA2 = oA1 + 64 * RS;
constexpr int oA3 = oA2 + 128 * RS;
constexpr int total = oA3 + 64 * RS;
}  // namespace cbfb_smem

__global__ void __launch_bounds__(MLP_THREADS, 1)
cbf_backward_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                    const int32_t* __restrict__ idx, const float* __restrict__ params,
                    const float* __restrict__ grad_h, int N, int M, int K, long long agents_total, int TA,
                    float dist_min, float obs_radius, float* __restrict__ partial, float* __restrict__ grad_states) {
  using namespace cbfb_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* W3 = smem + oW3; float* B3 = smem + oB3;
  float* W4 = smem + oW4; float* B4 = smem + oB4;
  float* mask = smem + oMask; float* D4 = smem + oD4; float* geo = smem + oGeo;
  int* ij = reinterpret_cast<int*>(smem + oIJ);
  float* A0 = smem + oA0; float* D0 = smem + oD0;
  float* A1 = smem + oA1; float* A2 = smem + oA2; float* A3 = smem + oA3;
  float* gp = partial + (size_t)blockIdx.x * MACBF_CBF_PARAMS;   // this CTA's gradient slice (zeroed by the host)

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
    __syncthreads();
    // ---- recompute features (as in cbf_forward_kernel) ----
    if (threadIdx.x < R) {
      const int r = threadIdx.x;
      float f[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f, gd = 1.f, gh = 0.f;
      int gi = -1, gj = -1;
      if (r < rows) {
        const long long ga = a0 + r / K;
        const int b = (int)(ga / N), i = (int)(ga % N);
        const int j = __ldg(idx + a0 * K + r);
        const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        const float4 sj = load_candidate(states, obstacles, b, j, N, M);
        const EdgeGeom g = edge_geom(si, sj, i, j);
        gd = cbf_dist(g.dx, g.dy);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye; f[5] = xsub(gd, dist_min);
        mk = (gd <= obs_radius) ? 1.f : 0.f;
        gh = __ldg(grad_h + a0 * K + r);
        gi = (int)ga;                               // agents_total < 2^31 is checked by the host
        gj = (j < N) ? (int)((long long)b * N + j) : -1;
      }
      mask[r] = mk;
      D4[r] = gh * mk;                              // d h / d y4 = mask (cbf.py:43)
      geo[r] = f[0]; geo[R + r] = f[1]; geo[2 * R + r] = gd;
      ij[r] = gi; ij[R + r] = gj;
#pragma unroll
      for (int c = 0; c < 6; ++c) A0[c * RS + r] = f[c];
      A0[6 * RS + r] = 0.f; A0[7 * RS + r] = 0.f;
    }
    __syncthreads();
    layer_nk<R, 8, 64, RS, RS, WS1, true>(A0, W1, B1, A1);
    __syncthreads();
    layer_nk<R, 64, 128, RS, RS, WS2, true>(A1, W2, B2, A2);
    __syncthreads();
    layer_nk<R, 128, 64, RS, RS, WS3, true>(A2, W3, B3, A3);
    __syncthreads();
    // ---- layer 4 (64 -> 1): dW4, db4, delta3 in place over a3 ----
    if (threadIdx.x < 64) {
      const int c = threadIdx.x;
      float s = 0.f;
      for (int r = 0; r < R; ++r) s = fmaf(D4[r], A3[c * RS + r], s);
      gp[CBF_W4 + c] += s;
    } else if (threadIdx.x == 64) {
      float s = 0.f;
      for (int r = 0; r < R; ++r) s += D4[r];
      gp[CBF_B4] += s;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 64 * R; e += MLP_THREADS) {
      const int c = e >> 7, r = e & (R - 1);
      const float a = A3[c * RS + r];
      A3[c * RS + r] = (a > 0.f) ? D4[r] * W4[c] : 0.f;
    }
    __syncthreads();
    // ---- layer 3 ----
    grad_w<R, 128, 64, RS, RS, 128, 128>(A3, A2, gp + CBF_W3, gp + CBF_B3);
    __syncthreads();
    layer_kn<R, 64, 128, RS, RS, WS3, true>(A3, W3, A2, A2);     // delta2 over a2
    __syncthreads();
    // ---- layer 2 ----
    grad_w<R, 64, 128, RS, RS, 64, 64>(A2, A1, gp + CBF_W2, gp + CBF_B2);
    __syncthreads();
    layer_kn<R, 128, 64, RS, RS, WS2, true>(A2, W2, A1, A1);     // delta1 over a1
    __syncthreads();
    // ---- layer 1 (6 -> 64): dW1, db1, delta0 ----
    for (int o = threadIdx.x; o < 64 * 7; o += MLP_THREADS) {
      const int n = o / 7, k = o % 7;                             // k == 6: bias
      float s = 0.f;
      if (k < 6) {
        for (int r = 0; r < R; ++r) s = fmaf(A1[n * RS + r], A0[k * RS + r], s);
        gp[CBF_W1 + n * 6 + k] += s;
      } else {
        for (int r = 0; r < R; ++r) s += A1[n * RS + r];
        gp[CBF_B1 + n] += s;
      }
    }
    if (grad_states != nullptr) {
      for (int o = threadIdx.x; o < 6 * R; o += MLP_THREADS) {
        const int k = o >> 7, r = o & (R - 1);
        float s = 0.f;
#pragma unroll 8
        for (int n = 0; n < 64; ++n) s = fmaf(W1[n * WS1 + k], A1[n * RS + r], s);
        D0[k * RS + r] = s;
      }
      __syncthreads();
      // ---- features -> states (cbf.py:23-30) ----
      if (threadIdx.x < rows) {
        const int r = threadIdx.x;
        const float dx = geo[r], dy = geo[R + r], inv_d = 1.f / geo[2 * R + r];
        const float g5 = D0[5 * RS + r];
        const float gx = D0[0 * RS + r] + g5 * dx * inv_d;
        const float gy = D0[1 * RS + r] + g5 * dy * inv_d;
        const float gvx = D0[2 * RS + r], gvy = D0[3 * RS + r];
        const int gi = ij[r], gj = ij[R + r];
        if (gi != gj) {   // the self edge (i == j) contributes +g - g = 0
          float* pi = grad_states + (size_t)gi * 4;
          atomicAdd(pi + 0, gx); atomicAdd(pi + 1, gy); atomicAdd(pi + 2, gvx); atomicAdd(pi + 3, gvy);
          if (gj >= 0) {
            float* pj = grad_states + (size_t)gj * 4;
            atomicAdd(pj + 0, -gx); atomicAdd(pj + 1, -gy); atomicAdd(pj + 2, -gvx); atomicAdd(pj + 3, -gvy);
          }
        }
      }
    }
  }
}

// out[p] += sum over slices, in slice order (deterministic)
__global__ void reduce_partials_kernel(const float* __restrict__ partial, int nslices, int P,
                                       float* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  float s = 0.f;
  for (int c = 0; c < nslices; ++c) s += partial[(size_t)c * P + p];
  out[p] += s;
}

static inline int cbf_bwd_grid(int B, int N, int K) {
  const long long agents = (long long)B * N;
  const int TA = max(1, cbfb_smem::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  return (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
}

size_t cbf_backward_tc_scratch_bytes(int B, int N, int K);
int cbf_backward_tc(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, double,
                    double, float*, float*, void*, size_t, cudaStream_t);
// MACBF_CBF_BWD_FFMA=1 keeps this FP32-FFMA kernel (A/B reference of the tensor-core path in cbf_bwd_tc.cu)
static bool cbf_bwd_force_ffma() {      // read per call, so a test can switch paths inside one process
  const char* e = getenv("MACBF_CBF_BWD_FFMA");
  return e != nullptr && e[0] == '1';
}

size_t cbf_backward_scratch_bytes(int B, int N, int K) {
  const size_t ffma = (size_t)cbf_bwd_grid(B, N, K) * MACBF_CBF_PARAMS * sizeof(float);
  const size_t tc = cbf_backward_tc_scratch_bytes(B, N, K);
  return ffma > tc ? ffma : tc;
}

int cbf_backward(const float* states, const float* obstacles, const int32_t* idx, const float* params,
                 const float* grad_h, int B, int N, int M, int K, double dist_min, double obs_radius,
                 float* grad_params, float* grad_states, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "cbf_backward: B*N must be < 2^31");
  if (!cbf_bwd_force_ffma())
    return cbf_backward_tc(states, obstacles, idx, params, grad_h, B, N, M, K, dist_min, obs_radius, grad_params,
                           grad_states, scratch, scratch_bytes, st);
  const int grid = cbf_bwd_grid(B, N, K);
  const size_t need = (size_t)grid * MACBF_CBF_PARAMS * sizeof(float);
  MACBF_CHECK_ARG(scratch_bytes >= need, "cbf_backward: scratch too small (%zu < %zu)", scratch_bytes, need);
  const int TA = max(1, cbfb_smem::R / K);
  const size_t smem = cbfb_smem::total * sizeof(float);
  MACBF_CHECK_CUDA(cudaMemsetAsync(scratch, 0, need, st));
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cbf_backward_kernel<<<grid, MLP_THREADS, smem, st>>>(states, obstacles, idx, params, grad_h, N, M, K, agents, TA,
                                                      (float)dist_min, (float)obs_radius, (float*)scratch,
                                                      grad_states);
  MACBF_CHECK_LAUNCH("cbf_backward_kernel");
  reduce_partials_kernel<<<(MACBF_CBF_PARAMS + 255) / 256, 256, 0, st>>>((const float*)scratch, grid,
                                                                         MACBF_CBF_PARAMS, grad_params);
  MACBF_CHECK_LAUNCH("reduce_partials_kernel");
  return 0;
}

}  // namespace macbf
