// K3a (tensor-core path) - controller edge net + masked max pooling (controller.py:33-46,
// REPAIRED) on tcgen05 / TMEM.  Same structure as cbf_tc.cu with one tensor-core layer:
//
//   features + layer 1 (5->64, FP32 FFMA) -> split hi/lo -> tcgen05.st  A2 [128 x 64] (TMEM)
//   layer 2: D2[128x128] = A2 * W2^T      3xTF32 tcgen05.mma (A from TMEM, W2 hi/lo in smem)
//   bias + ReLU + * mask (tcgen05.ld)     -> shared-memory stage [128 ch][128 rows]
//   max over the K slots of every agent   -> pool [agents,128] (global)
//
// TMEM columns: [0,64) A2_hi, [64,128) A2_lo, [128,256) D2.
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

namespace cetc {
constexpr int THREADS = 512;
constexpr int R = 128;
constexpr int RS = 132;   // stage row stride: lanes along channels hit distinct banks (see pooling)
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_D2 = 128;
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW1 = oW2LO + 128 * 64;     // [64][8]: 5 weights, 0, bias, 0
constexpr int oB2 = oW1 + 64 * 8;
constexpr int oMask = oB2 + 128;
constexpr int oStage = oMask + R;
constexpr int total_floats = oStage + 128 * RS;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cetc

__global__ void __launch_bounds__(cetc::THREADS, 1)
ctrl_edge_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                    const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int K,
                    long long agents_total, int TA, float obs_radius, float* __restrict__ pool) {
  using namespace cetc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO;
  float* W1 = sm + oW1; float* B2 = sm + oB2; float* mask = sm + oMask; float* stage = sm + oStage;
  __shared__ uint64_t bar_d2;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cgrp = warp >> 2;
  const int r = q * 32 + lane;

  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  if (threadIdx.x == 32) {
    mbar_init(&bar_d2, 1);
    fence_mbar_init();
  }
  fill_weight_sw128<128, 64>(params + CT_EW2, 64, W2hi, W2lo);
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) {
    const int n = t >> 3, k = t & 7;
    W1[t] = (k < 5) ? __ldg(params + CT_EW1 + n * 5 + k) : (k == 6 ? __ldg(params + CT_EB1 + n) : 0.f);
  }
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
  const uint32_t w2hi_a = smem_u32(W2hi), w2lo_a = smem_u32(W2lo);
  constexpr uint32_t IDESC2 = idesc_tf32(128, 128);

  const long long ntiles = (agents_total + TA - 1) / TA;
  uint32_t parity = 0;
  int pf_j = 0, pf_b = 0, pf_i = 0;
  bool pf_valid = false;
  float4 pf_si = make_float4(0.f, 0.f, 0.f, 0.f), pf_sj = pf_si;
  auto prefetch_ij = [&](long long t) {
    pf_valid = false;
    if (t < ntiles) {
      const long long a0n = t * TA;
      const int rowsn = (int)min((long long)TA, agents_total - a0n) * K;
      if (r < rowsn) {
        const long long ga = a0n + r / K;
        pf_b = (int)(ga / N); pf_i = (int)(ga % N);
        pf_j = __ldg(idx + a0n * K + r);
        pf_si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        pf_valid = true;
      }
    }
  };
  auto prefetch_sj = [&]() {
    if (pf_valid) pf_sj = load_candidate(states, obstacles, pf_b, pf_j, N, M);
  };
  prefetch_ij(blockIdx.x);
  prefetch_sj();
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, parity ^= 1) {
    const long long a0 = tile * TA;
    const int na = (int)min((long long)TA, agents_total - a0);
    const int rows = na * K;
    // ---- features (controller.py:33-40) and layer 1 for this thread's 16 output channels ----
    {
      float f[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (r < rows) {
        const EdgeGeom g = edge_geom(pf_si, pf_sj, pf_i, pf_j);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        mk = (ctrl_dist(g.dx, g.dy) < obs_radius) ? 1.f : 0.f;
      }
      if (cgrp == 0) mask[r] = mk;
      float hi[16], lo[16];
#pragma unroll
      for (int jn = 0; jn < 16; ++jn) {
        const float4 wa = *reinterpret_cast<const float4*>(W1 + (cgrp * 16 + jn) * 8);
        const float4 wb = *reinterpret_cast<const float4*>(W1 + (cgrp * 16 + jn) * 8 + 4);
        float v = wa.x * f[0];
        v = fmaf(wa.y, f[1], v); v = fmaf(wa.z, f[2], v); v = fmaf(wa.w, f[3], v);
        v = fmaf(wb.x, f[4], v);
        v = fmaxf(v + wb.z, 0.f);
        tf32_split(v, hi[jn], lo[jn]);
      }
      tmem_st16(tlane + C_A2HI + cgrp * 16, hi);
      tmem_st16(tlane + C_A2LO + cgrp * 16, lo);
      wait_st();
    }
    fence_before_sync();
    __syncthreads();   // also: the previous tile's pooling has finished reading `stage`
    if (threadIdx.x == 0) {
      fence_after_sync();
      uint32_t acc = 0;
#pragma unroll 1
      for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi
        const uint32_t a_col = tbase + (term == 0 ? C_A2LO : C_A2HI);
        const uint32_t b_base = (term == 1) ? w2lo_a : w2hi_a;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t baddr = b_base + (ks >> 2) * (128 * 128) + (ks & 3) * 32;
          mma_tf32_ts(tbase + C_D2, a_col + ks * 8, smem_desc_sw128(baddr), IDESC2, acc);
          acc = 1;
        }
      }
      mma_commit(&bar_d2);
    }
    prefetch_ij(tile + gridDim.x);
    const float mk = mask[r];
    mbar_wait(&bar_d2, parity);
    fence_after_sync();
    prefetch_sj();
    // ---- bias + ReLU + mask on this thread's 32 columns -> stage[c][r] ----
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int c0 = cgrp * 32 + half * 16;
      float v[16];
      tmem_ld16(tlane + C_D2 + c0, v);
      wait_ld();
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 bb = *reinterpret_cast<const float4*>(B2 + c0 + 4 * j4);
        stage[(c0 + 4 * j4 + 0) * RS + r] = fmaxf(v[4 * j4 + 0] + bb.x, 0.f) * mk;
        stage[(c0 + 4 * j4 + 1) * RS + r] = fmaxf(v[4 * j4 + 1] + bb.y, 0.f) * mk;
        stage[(c0 + 4 * j4 + 2) * RS + r] = fmaxf(v[4 * j4 + 2] + bb.z, 0.f) * mk;
        stage[(c0 + 4 * j4 + 3) * RS + r] = fmaxf(v[4 * j4 + 3] + bb.w, 0.f) * mk;
      }
    }
    fence_before_sync();
    __syncthreads();   // stage complete; all D2 reads done before the next tile's layer 2 rewrites it
    fence_after_sync();
    // ---- max over the K slots (controller.py:46, repair F2: values) ----
    for (int item = threadIdx.x; item < na * 128; item += THREADS) {
      const int c = item & 127, a = item >> 7;
      const float* row = stage + c * RS + a * K;
      const int rot = (c >> 3) & 3;   // lanes c and c+8 share a bank: start them on different slots
      float best = 0.f;               // relu(.) * mask >= 0 and slot 0 (self) is always unmasked
      for (int k = 0; k < K; ++k) {
        int kk = k + rot; kk = kk >= K ? kk - K : kk;
        best = fmaxf(best, row[kk]);
      }
      pool[(a0 + a) * 128 + c] = best;
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st) {
  const long long agents = (long long)B * N;
  const int TA = max(1, cetc::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cetc::smem_bytes));
  ctrl_edge_tc_kernel<<<grid, cetc::THREADS, cetc::smem_bytes, st>>>(states, obstacles, idx, params, N, M, K, agents, TA,
                                                                    (float)obs_radius, pool);
  MACBF_CHECK_LAUNCH("ctrl_edge_tc_kernel");
  return 0;
}

}  // namespace macbf
