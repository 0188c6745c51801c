// K3b - backward of the controller (autograd of controller.py:31-63 + the Euler step of
// train.py:70 / core.py:133-134, REPAIRED semantics).  Recompute-based, like K2b:
//   1. ctrl_edge_kernel (forward) rebuilds the pooled features,
//   2. ctrl_node_backward_kernel: node MLP forward + backward per tile of 64 agents; emits
//      d/dpool in place of pool, d/dstates and d/dgoals of the (p-g, v) inputs and of the action,
//   3. ctrl_edge_backward_kernel: edge MLP forward, routes d/dpool to the arg-max slot of every
//      (agent, channel) (torch.max backward; first maximum), backward through 5->64->128,
//   4. reduce_partials_kernel: per-CTA weight-gradient slices -> grad_params, in CTA order.
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace macbf {

__global__ void reduce_partials_kernel(const float* __restrict__ partial, int nslices, int P,
                                       float* __restrict__ out);
int controller_forward_pool_only(const float* states, const float* obstacles, const int32_t* idx,
                                 const float* params, int B, int N, int M, int K, double obs_radius, float* pool,
                                 cudaStream_t st);

// ------------------------------------------------------------------------------------------
namespace cnb_smem {
constexpr int R = 64;
constexpr int RS = 68;
constexpr int KP1 = 144;   // 132 inputs padded to a multiple of 16 (grad_w splits k over 16 threads)
constexpr int WS1 = 148, WS2 = 68, WS3 = 132, WS4 = 68;
constexpr int oW1 = 0;
constexpr int oB1 = oW1 + 64 * WS1;
constexpr int oW2 = oB1 + 64;
constexpr int oB2 = oW2 + 128 * WS2;
constexpr int oW3 = oB2 + 128;
constexpr int oB3 = oW3 + 64 * WS3;
constexpr int oW4 = oB3 + 64;
constexpr int oB4 = oW4 + 4 * WS4;
constexpr int oD4 = oB4 + 4;            // [4][RS]
constexpr int oA0 = oD4 + 4 * RS;       // [144][RS]
constexpr int oA1 = oA0 + KP1 * RS;     // [64][RS]
constexpr int oA2 = oA1 + 64 * RS;      // [128][RS]
constexpr int oA3 = oA2 + 128 * RS;     // [64][RS]
constexpr int total = oA3 + 64 * RS;
}  // namespace cnb_smem

__global__ void __launch_bounds__(MLP_THREADS, 1)
ctrl_node_backward_kernel(const float* __restrict__ states, const float* __restrict__ goals,
                          const float* __restrict__ params, const float* __restrict__ grad_u,
                          const float* __restrict__ grad_s_next, long long agents_total, float dt,
                          float* __restrict__ pool_dz, float* __restrict__ partial,
                          float* __restrict__ grad_states, float* __restrict__ grad_goals) {
  using namespace cnb_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* W3 = smem + oW3; float* B3 = smem + oB3;
  float* W4 = smem + oW4; float* B4 = smem + oB4;
  float* D4 = smem + oD4;
  float* A0 = smem + oA0; float* A1 = smem + oA1; float* A2 = smem + oA2; float* A3 = smem + oA3;
  float* gp = partial + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;

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
    // ---- A0 = [pool, p-g, v, 0...]^T ----
    float ex = 0.f, ey = 0.f, vx = 0.f, vy = 0.f;   // row-owner registers (threads 0..63)
    {
      const int rr = lane & 7, cc = lane >> 3;
      for (int it = warp; it < 8 * 8; it += MLP_THREADS / 32) {
        const int r = (it >> 3) * 8 + rr, c4 = (it & 7) * 4 + cc;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < rows) v = __ldg(reinterpret_cast<const float4*>(pool_dz + (a0 + r) * 128) + c4);
        A0[(4 * c4 + 0) * RS + r] = v.x; A0[(4 * c4 + 1) * RS + r] = v.y;
        A0[(4 * c4 + 2) * RS + r] = v.z; A0[(4 * c4 + 3) * RS + r] = v.w;
      }
      if (threadIdx.x < R) {
        const int r = threadIdx.x;
        if (r < rows) {
          const float4 s = __ldg(reinterpret_cast<const float4*>(states) + a0 + r);
          const float2 g = __ldg(reinterpret_cast<const float2*>(goals) + a0 + r);
          ex = xsub(s.x, g.x); ey = xsub(s.y, g.y); vx = s.z; vy = s.w;
        }
        A0[128 * RS + r] = ex; A0[129 * RS + r] = ey; A0[130 * RS + r] = vx; A0[131 * RS + r] = vy;
#pragma unroll
        for (int c = 132; c < KP1; ++c) A0[c * RS + r] = 0.f;
      }
    }
    __syncthreads();
    layer_nk<R, 136, 64, RS, RS, WS1, true>(A0, W1, B1, A1);
    __syncthreads();
    layer_nk<R, 64, 128, RS, RS, WS2, true>(A1, W2, B2, A2);
    __syncthreads();
    layer_nk<R, 128, 64, RS, RS, WS3, true>(A2, W3, B3, A3);
    __syncthreads();
    // ---- output layer + gains + action (controller.py:50-61): d/d(pre-sigmoid) per row ----
    float dex = 0.f, dey = 0.f, dvx = 0.f, dvy = 0.f;   // direct d/d(e), d/d(v) through the action
    float4 gsn = make_float4(0.f, 0.f, 0.f, 0.f);
    if (threadIdx.x < R) {
      const int r = threadIdx.x;
      float d_o[4] = {0.f, 0.f, 0.f, 0.f};
      if (r < rows) {
        float o[4] = {B4[0], B4[1], B4[2], B4[3]};
#pragma unroll 4
        for (int c = 0; c < 64; ++c) {
          const float a = A3[c * RS + r];
#pragma unroll
          for (int q = 0; q < 4; ++q) o[q] = fmaf(W4[q * WS4 + c], a, o[q]);
        }
        float gux = 0.f, guy = 0.f;
        if (grad_u != nullptr) {
          const float2 g = __ldg(reinterpret_cast<const float2*>(grad_u) + a0 + r);
          gux = g.x; guy = g.y;
        }
        if (grad_s_next != nullptr) {   // s' = s + [v, u] * dt
          gsn = __ldg(reinterpret_cast<const float4*>(grad_s_next) + a0 + r);
          gux = fmaf(dt, gsn.z, gux); guy = fmaf(dt, gsn.w, guy);
        }
        float kk[4], sg[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          sg[q] = 1.0f / (1.0f + expf(-o[q]));
          kk[q] = 2.0f * sg[q] + 0.2f;
        }
        // a_x = -(k1 e_x + k2 v_x), a_y = -(k3 e_y + k4 v_y)
        const float dk[4] = {-ex * gux, -vx * gux, -ey * guy, -vy * guy};
        dex = -kk[0] * gux; dvx = -kk[1] * gux; dey = -kk[2] * guy; dvy = -kk[3] * guy;
#pragma unroll
        for (int q = 0; q < 4; ++q) d_o[q] = dk[q] * 2.0f * sg[q] * (1.0f - sg[q]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) D4[q * RS + r] = d_o[q];
    }
    __syncthreads();
    {  // dW4 [4][64], db4
      const int q = threadIdx.x >> 6, c = threadIdx.x & 63;
      float s = 0.f;
      for (int r = 0; r < R; ++r) s = fmaf(D4[q * RS + r], A3[c * RS + r], s);
      gp[CT_NW4 + q * 64 + c] += s;
      if (threadIdx.x < 4) {
        float b = 0.f;
        for (int r = 0; r < R; ++r) b += D4[threadIdx.x * RS + r];
        gp[CT_NB4 + threadIdx.x] += b;
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 64 * R; e += MLP_THREADS) {   // delta3 over a3
      const int c = e >> 6, r = e & (R - 1);
      float s = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) s = fmaf(W4[q * WS4 + c], D4[q * RS + r], s);
      A3[c * RS + r] = (A3[c * RS + r] > 0.f) ? s : 0.f;
    }
    __syncthreads();
    grad_w<R, 128, 64, RS, RS, 128, 128>(A3, A2, gp + CT_NW3, gp + CT_NB3);
    __syncthreads();
    layer_kn<R, 64, 128, RS, RS, WS3, true>(A3, W3, A2, A2);
    __syncthreads();
    grad_w<R, 64, 128, RS, RS, 64, 64>(A2, A1, gp + CT_NW2, gp + CT_NB2);
    __syncthreads();
    layer_kn<R, 128, 64, RS, RS, WS2, true>(A2, W2, A1, A1);
    __syncthreads();
    grad_w<R, KP1, 64, RS, RS, 132, 132>(A1, A0, gp + CT_NW1, gp + CT_NB1);
    layer_kn<R, 64, 128, RS, RS, WS1, false>(A1, W1, nullptr, A2);   // d/dpool -> A2[c][r]
    {  // d/d(p-g), d/dv through the first node layer -> D4[q][r]
      const int q = threadIdx.x >> 6, r = threadIdx.x & 63;
      float s = 0.f;
#pragma unroll 8
      for (int n = 0; n < 64; ++n) s = fmaf(W1[n * WS1 + 128 + q], A1[n * RS + r], s);
      D4[q * RS + r] = s;
    }
    __syncthreads();
    {  // d/dpool replaces pool in global memory (this tile's rows were fully read above)
      const int rr = lane & 7, cc = lane >> 3;
      for (int it = warp; it < 8 * 8; it += MLP_THREADS / 32) {
        const int r = (it >> 3) * 8 + rr, c4 = (it & 7) * 4 + cc;
        if (r < rows) {
          float4 v;
          v.x = A2[(4 * c4 + 0) * RS + r]; v.y = A2[(4 * c4 + 1) * RS + r];
          v.z = A2[(4 * c4 + 2) * RS + r]; v.w = A2[(4 * c4 + 3) * RS + r];
          reinterpret_cast<float4*>(pool_dz + (a0 + r) * 128)[c4] = v;
        }
      }
    }
    if (threadIdx.x < rows) {
      const int r = threadIdx.x;
      const float gex = dex + D4[0 * RS + r], gey = dey + D4[1 * RS + r];
      const float gvx = dvx + D4[2 * RS + r], gvy = dvy + D4[3 * RS + r];
      if (grad_states != nullptr) {
        float4* p = reinterpret_cast<float4*>(grad_states) + a0 + r;
        float4 g = *p;
        // e = p - g ; s'_p = p + dt v ; s'_v = v + dt u (u handled above)
        g.x += gex + gsn.x; g.y += gey + gsn.y;
        g.z += gvx + gsn.z + dt * gsn.x; g.w += gvy + gsn.w + dt * gsn.y;
        *p = g;
      }
      if (grad_goals != nullptr) {
        float2* p = reinterpret_cast<float2*>(grad_goals) + a0 + r;
        float2 g = *p;
        g.x -= gex; g.y -= gey;
        *p = g;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
namespace ceb_smem {
constexpr int R = 128;
constexpr int RS = 132;
constexpr int WS1 = 8, WS2 = 68;
constexpr int oW1 = 0;
constexpr int oB1 = oW1 + 64 * WS1;
constexpr int oW2 = oB1 + 64;
constexpr int oB2 = oW2 + 128 * WS2;
constexpr int oMask = oB2 + 128;
constexpr int oIJ = oMask + R;
constexpr int oA0 = oIJ + 2 * R;
constexpr int oD0 = oA0 + 8 * RS;
constexpr int oA1 = oD0 + 8 * RS;
constexpr int oA2 = oA1 + 64 * RS;
constexpr int total = oA2 + 128 * RS;
}  // namespace ceb_smem

__global__ void __launch_bounds__(MLP_THREADS, 1)
ctrl_edge_backward_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                          const int32_t* __restrict__ idx, const float* __restrict__ params,
                          const float* __restrict__ dz, int N, int M, int K, long long agents_total, int TA,
                          float obs_radius, float* __restrict__ partial, float* __restrict__ grad_states) {
  using namespace ceb_smem;
  extern __shared__ __align__(16) float smem[];
  float* W1 = smem + oW1; float* B1 = smem + oB1;
  float* W2 = smem + oW2; float* B2 = smem + oB2;
  float* mask = smem + oMask;
  int* ij = reinterpret_cast<int*>(smem + oIJ);
  float* A0 = smem + oA0; float* D0 = smem + oD0; float* A1 = smem + oA1; float* A2 = smem + oA2;
  float* gp = partial + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;

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
      int gi = -1, gj = -1;
      if (r < rows) {
        const long long ga = a0 + r / K;
        const int b = (int)(ga / N), i = (int)(ga % N);
        const int j = __ldg(idx + a0 * K + r);
        const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        const float4 sj = load_candidate(states, obstacles, b, j, N, M);
        const EdgeGeom g = edge_geom(si, sj, i, j);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        mk = (ctrl_dist(g.dx, g.dy) < obs_radius) ? 1.f : 0.f;
        gi = (int)ga;
        gj = (j < N) ? (int)((long long)b * N + j) : -1;
      }
      mask[r] = mk;
      ij[r] = gi; ij[R + r] = gj;
#pragma unroll
      for (int c = 0; c < 5; ++c) A0[c * RS + r] = f[c];
      A0[5 * RS + r] = 0.f; A0[6 * RS + r] = 0.f; A0[7 * RS + r] = 0.f;
    }
    __syncthreads();
    layer_nk<R, 8, 64, RS, RS, WS1, true>(A0, W1, B1, A1);
    __syncthreads();
    layer_nk<R, 64, 128, RS, RS, WS2, true>(A1, W2, B2, A2);
    __syncthreads();
    // ---- max-pool backward (controller.py:46): delta2 over a2 ----
    for (int item = threadIdx.x; item < na * 128; item += MLP_THREADS) {
      const int c = item & 127, a = item >> 7;
      float* row = A2 + c * RS + a * K;
      const float* mrow = mask + a * K;
      float best = row[0] * mrow[0];
      int bi = 0;
      for (int k = 1; k < K; ++k) {
        const float v = row[k] * mrow[k];
        if (v > best) { best = v; bi = k; }
      }
      // through "* mask" and relu': only a strictly positive winner carries gradient
      const float g = (best > 0.f) ? __ldg(dz + (a0 + a) * 128 + c) : 0.f;
      for (int k = 0; k < K; ++k) row[k] = (k == bi) ? g : 0.f;
    }
    for (int e = threadIdx.x; e < 128 * (R - rows); e += MLP_THREADS) {   // padded rows carry nothing
      const int c = e / (R - rows), r = rows + e % (R - rows);
      A2[c * RS + r] = 0.f;
    }
    __syncthreads();
    grad_w<R, 64, 128, RS, RS, 64, 64>(A2, A1, gp + CT_EW2, gp + CT_EB2);
    __syncthreads();
    layer_kn<R, 128, 64, RS, RS, WS2, true>(A2, W2, A1, A1);   // delta1 over a1
    __syncthreads();
    for (int o = threadIdx.x; o < 64 * 6; o += MLP_THREADS) {   // dW1 [64][5], db1
      const int n = o / 6, k = o % 6;
      float s = 0.f;
      if (k < 5) {
        for (int r = 0; r < R; ++r) s = fmaf(A1[n * RS + r], A0[k * RS + r], s);
        gp[CT_EW1 + n * 5 + k] += s;
      } else {
        for (int r = 0; r < R; ++r) s += A1[n * RS + r];
        gp[CT_EB1 + n] += s;
      }
    }
    if (grad_states != nullptr) {
      for (int o = threadIdx.x; o < 4 * R; o += MLP_THREADS) {
        const int k = o >> 7, r = o & (R - 1);
        float s = 0.f;
#pragma unroll 8
        for (int n = 0; n < 64; ++n) s = fmaf(W1[n * WS1 + k], A1[n * RS + r], s);
        D0[k * RS + r] = s;
      }
      __syncthreads();
      if (threadIdx.x < rows) {
        const int r = threadIdx.x;
        const int gi = ij[r], gj = ij[R + r];
        if (gi != gj) {
          const float gx = D0[r], gy = D0[RS + r], gvx = D0[2 * RS + r], gvy = D0[3 * RS + r];
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

static inline int grid_for(long long ntiles) { return (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS); }

static inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

bool ctrl_edge_bwd_tc_applicable(int K);
int ctrl_edge_bwd_tc_slices();
size_t ctrl_edge_bwd_tc_scratch_bytes(int B, int N, int K);
uint8_t* ctrl_edge_bwd_tc_argk(void* scratch);
int ctrl_edge_bwd_tc(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, float*,
                     float*, void*, cudaStream_t, const uint8_t*);
int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk);
size_t ctrl_node_bwd_tc_scratch_bytes(int B, int N);
int ctrl_node_bwd_tc(const float*, const float*, const float*, const float*, const float*, int, int, double, float*, float*,
                     float*, float*, void*, cudaStream_t);
constexpr int CTRL_SLICES = 2 * NUM_SMS;   // the tensor-core edge backward runs two CTAs per SM
// MACBF_CTRL_BWD_FFMA=1 keeps the FP32-FFMA edge backward (A/B reference of ctrl_edge_bwd_tc.cu)
static bool ctrl_bwd_force_ffma() {     // read per call, so a test can switch paths inside one process
  const char* e = getenv("MACBF_CTRL_BWD_FFMA");
  return e != nullptr && e[0] == '1';
}
static inline size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

// scratch: [pool / dz : B*N*128 f32][partial : CTRL_SLICES * 34052 f32][winners + tile records of the tensor-core edge backward]
size_t controller_backward_scratch_bytes(int B, int N, int K) {
  return a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)) + a256((size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float)) +
         max_sz(ctrl_edge_bwd_tc_applicable(K) ? ctrl_edge_bwd_tc_scratch_bytes(B, N, K) : 0,
                ((size_t)B * N * 128 + 4096) + ctrl_node_bwd_tc_scratch_bytes(B, N));   // node records sit behind the winners
}

int controller_backward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                        const float* params, const float* grad_u, const float* grad_s_next, int B, int N, int M,
                        int K, double obs_radius, double dt, float* grad_params, float* grad_states,
                        float* grad_goals, void* scratch, size_t scratch_bytes, cudaStream_t st, float* pool_saved,
                        const uint8_t* argk_saved) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "controller_backward: B*N must be < 2^31");
  const size_t need = controller_backward_scratch_bytes(B, N, K);
  MACBF_CHECK_ARG(scratch_bytes >= need, "controller_backward: scratch too small (%zu < %zu)", scratch_bytes, need);
  // pooled features and arg-max slots kept by the training forward (macbf_step_forward_ex with
  // MACBF_STEP_SAVE_WINNERS) make the edge-net recompute unnecessary; pool_saved is consumed (becomes d/dpool)
  const bool saved = pool_saved != nullptr && argk_saved != nullptr && ctrl_edge_bwd_tc_applicable(K) && !ctrl_bwd_force_ffma();
  float* pool = saved ? pool_saved : (float*)scratch;
  float* partial = (float*)((char*)scratch + a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)));
  MACBF_CHECK_CUDA(cudaMemsetAsync(partial, 0, (size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float), st));
  const bool edge_tc = ctrl_edge_bwd_tc_applicable(K) && !ctrl_bwd_force_ffma();
  void* edge_scratch = (char*)partial + a256((size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float));
  int rc = 0;
  if (!saved)
    rc = edge_tc ? ctrl_edge_tc(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st, ctrl_edge_bwd_tc_argk(edge_scratch))
                 : controller_forward_pool_only(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st);
  if (rc) return rc;
  if (!ctrl_bwd_force_ffma()) {
    // the node records live where the edge backward's winners / records go afterwards - except the winners, which the
    // recomputed forward has already written: keep them (first B*N*128 bytes of edge_scratch, 1024-aligned)
    char* node_scratch = (char*)edge_scratch + (edge_tc ? (((size_t)agents * 128 + 2047) & ~(size_t)1023) : 0);
    rc = ctrl_node_bwd_tc(states, goals, params, grad_u, grad_s_next, B, N, dt, pool, partial, grad_states, grad_goals,
                          node_scratch, st);
    if (rc) return rc;
  } else {
    const long long ntiles = (agents + cnb_smem::R - 1) / cnb_smem::R;
    const size_t smem = cnb_smem::total * sizeof(float);
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_node_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctrl_node_backward_kernel<<<grid_for(ntiles), MLP_THREADS, smem, st>>>(
        states, goals, params, grad_u, grad_s_next, agents, (float)dt, pool, partial, grad_states, grad_goals);
    MACBF_CHECK_LAUNCH("ctrl_node_backward_kernel");
  }
  if (edge_tc) {
    rc = ctrl_edge_bwd_tc(states, obstacles, idx, params, pool, B, N, M, K, partial, grad_states, edge_scratch, st,
                          saved ? argk_saved : nullptr);
    if (rc) return rc;
  } else {
    const int TA = max(1, ceb_smem::R / K);
    const long long ntiles = (agents + TA - 1) / TA;
    const size_t smem = ceb_smem::total * sizeof(float);
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctrl_edge_backward_kernel<<<grid_for(ntiles), MLP_THREADS, smem, st>>>(
        states, obstacles, idx, params, pool, N, M, K, agents, TA, (float)obs_radius, partial, grad_states);
    MACBF_CHECK_LAUNCH("ctrl_edge_backward_kernel");
  }
  reduce_partials_kernel<<<(MACBF_CTRL_PARAMS + 255) / 256, 256, 0, st>>>(partial, CTRL_SLICES, MACBF_CTRL_PARAMS,
                                                                          grad_params);
  MACBF_CHECK_LAUNCH("reduce_partials_kernel");
  return 0;
}

}  // namespace macbf
