// K2 (tensor-core path) - fused CBF head (cbf.py:21-45, REPAIRED) on tcgen05 / TMEM.
//
// One persistent CTA per SM, 16 warps.  A tile is 128 edge rows (floor(128/K) agents); row r is
// TMEM lane r.  Warp w works on lanes 32*(w%4).. (the only lanes tcgen05.ld/st lets it touch)
// and on column group w/4 of whatever is being produced.
//
//   features + layer 1 (6->64, FP32 FFMA)  -> split hi/lo -> tcgen05.st  A2 [128 x 64] (TMEM)
//   layer 2: D2[128x128] = A2 * W2^T       3xTF32 tcgen05.mma, A from TMEM, B (weights) from smem
//   bias+ReLU+split on D2 (tcgen05.ld)     -> tcgen05.st  A3 [128 x 128] (TMEM)
//   layer 3: D3[128x64]  = A3 * W3^T       3xTF32 tcgen05.mma
//   bias+ReLU, 64->1 dot, * mask           -> h (global)
//
// Hidden activations never leave the SM: they go registers -> TMEM -> tensor core -> TMEM ->
// registers.  Shared memory holds only the weights (hi and lo tf32 halves of W2, W3 in the
// canonical SWIZZLE_128B K-major layout, 128 KB) and a few hundred bytes of vectors.
// TMEM columns: [0,64) A2_hi / later D3, [64,128) A2_lo, [128,256) D2, [256,384) A3_hi,
// [384,512) A3_lo.
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

namespace cbftc {
constexpr int THREADS = 512;
constexpr int R = 128;
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_D3 = 0, C_D2 = 128, C_A3HI = 256, C_A3LO = 384;
// dynamic shared memory (floats, after 1024-B alignment)
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW3HI = oW2LO + 128 * 64;
constexpr int oW3LO = oW3HI + 64 * 128;
constexpr int oW1 = oW3LO + 64 * 128;      // [64][8]: 6 weights, bias, 0
constexpr int oB2 = oW1 + 64 * 8;
constexpr int oB3 = oB2 + 128;
constexpr int oW4 = oB3 + 64;
constexpr int oB4 = oW4 + 64;
constexpr int oMask = oB4 + 4;
constexpr int oHpart = oMask + R;          // [4][128]
constexpr int total_floats = oHpart + 4 * R;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cbftc

__global__ void __launch_bounds__(cbftc::THREADS, 1)
cbf_forward_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                      const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int K,
                      long long agents_total, int TA, float dist_min, float obs_radius, float* __restrict__ h) {
  using namespace cbftc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO;
  float* W3hi = sm + oW3HI; float* W3lo = sm + oW3LO;
  float* W1 = sm + oW1; float* B2 = sm + oB2; float* B3 = sm + oB3; float* W4 = sm + oW4; float* B4 = sm + oB4;
  float* mask = sm + oMask; float* hpart = sm + oHpart;
  __shared__ uint64_t bar_d2, bar_d3;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cgrp = warp >> 2;
  const int r = q * 32 + lane;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&bar_d2, 1);
    mbar_init(&bar_d3, 1);
    fence_mbar_init();
  }
  fill_weight_sw128<128, 64>(params + CBF_W2, 64, W2hi, W2lo);
  fill_weight_sw128<64, 128>(params + CBF_W3, 128, W3hi, W3lo);
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) {
    const int n = t >> 3, k = t & 7;
    W1[t] = (k < 6) ? __ldg(params + CBF_W1 + n * 6 + k) : (k == 6 ? __ldg(params + CBF_B1 + n) : 0.f);
  }
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CBF_B2 + t);
  for (int t = threadIdx.x; t < 64; t += THREADS) {
    B3[t] = __ldg(params + CBF_B3 + t);
    W4[t] = __ldg(params + CBF_W4 + t);
  }
  if (threadIdx.x == 0) B4[0] = __ldg(params + CBF_B4);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // weights: generic proxy -> tensor-core proxy
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
  const uint32_t w2hi_a = smem_u32(W2hi), w2lo_a = smem_u32(W2lo), w3hi_a = smem_u32(W3hi), w3lo_a = smem_u32(W3lo);
  constexpr uint32_t IDESC2 = idesc_tf32(128, 128), IDESC3 = idesc_tf32(128, 64);

  const long long ntiles = (agents_total + TA - 1) / TA;
  uint32_t parity = 0;
  // The edge endpoints of the NEXT tile are fetched while the tensor core works on this one:
  // idx and s_i right after layer 2 is issued, s_j (which needs idx) after layer 3 is issued.
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
    // ---- features (cbf.py:23-37) and layer 1 for this thread's 16 output channels ----
    {
      float f[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (r < rows) {
        const EdgeGeom g = edge_geom(pf_si, pf_sj, pf_i, pf_j);
        const float d = cbf_dist(g.dx, g.dy);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye; f[5] = xsub(d, dist_min);
        mk = (d <= obs_radius) ? 1.f : 0.f;
      }
      if (cgrp == 0) mask[r] = mk;
      float hi[16], lo[16];
#pragma unroll
      for (int jn = 0; jn < 16; ++jn) {
        const float4 wa = *reinterpret_cast<const float4*>(W1 + (cgrp * 16 + jn) * 8);
        const float4 wb = *reinterpret_cast<const float4*>(W1 + (cgrp * 16 + jn) * 8 + 4);
        float v = wa.x * f[0];
        v = fmaf(wa.y, f[1], v); v = fmaf(wa.z, f[2], v); v = fmaf(wa.w, f[3], v);
        v = fmaf(wb.x, f[4], v); v = fmaf(wb.y, f[5], v);
        v = fmaxf(v + wb.z, 0.f);
        tf32_split(v, hi[jn], lo[jn]);
      }
      tmem_st16(tlane + C_A2HI + cgrp * 16, hi);
      tmem_st16(tlane + C_A2LO + cgrp * 16, lo);
      wait_st();
    }
    fence_before_sync();
    __syncthreads();
    // ---- layer 2 on the tensor core ----
    if (threadIdx.x == 0) {
      fence_after_sync();
      uint32_t acc = 0;
#pragma unroll 1
      for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi (small terms first)
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
    mbar_wait(&bar_d2, parity);
    fence_after_sync();
    // ---- bias + ReLU + split on this thread's 32 columns of D2 -> A3 ----
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int c0 = cgrp * 32 + half * 16;
      float v[16], hi[16], lo[16];
      tmem_ld16(tlane + C_D2 + c0, v);
      wait_ld();
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 bb = *reinterpret_cast<const float4*>(B2 + c0 + 4 * j4);
        tf32_split(fmaxf(v[4 * j4 + 0] + bb.x, 0.f), hi[4 * j4 + 0], lo[4 * j4 + 0]);
        tf32_split(fmaxf(v[4 * j4 + 1] + bb.y, 0.f), hi[4 * j4 + 1], lo[4 * j4 + 1]);
        tf32_split(fmaxf(v[4 * j4 + 2] + bb.z, 0.f), hi[4 * j4 + 2], lo[4 * j4 + 2]);
        tf32_split(fmaxf(v[4 * j4 + 3] + bb.w, 0.f), hi[4 * j4 + 3], lo[4 * j4 + 3]);
      }
      tmem_st16(tlane + C_A3HI + c0, hi);
      tmem_st16(tlane + C_A3LO + c0, lo);
    }
    wait_st();
    fence_before_sync();
    __syncthreads();
    // ---- layer 3 on the tensor core (D3 reuses the A2 columns: layer 2 has completed) ----
    if (threadIdx.x == 0) {
      fence_after_sync();
      uint32_t acc = 0;
#pragma unroll 1
      for (int term = 0; term < 3; ++term) {
        const uint32_t a_col = tbase + (term == 0 ? C_A3LO : C_A3HI);
        const uint32_t b_base = (term == 1) ? w3lo_a : w3hi_a;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) {
          const uint32_t baddr = b_base + (ks >> 2) * (64 * 128) + (ks & 3) * 32;
          mma_tf32_ts(tbase + C_D3, a_col + ks * 8, smem_desc_sw128(baddr), IDESC3, acc);
          acc = 1;
        }
      }
      mma_commit(&bar_d3);
    }
    prefetch_sj();
    mbar_wait(&bar_d3, parity);
    fence_after_sync();
    // ---- bias + ReLU, 64 -> 1 (this thread: 16 of the 64 channels), * mask (cbf.py:43) ----
    {
      const int c0 = cgrp * 16;
      float v[16];
      tmem_ld16(tlane + C_D3 + c0, v);
      wait_ld();
      float s = 0.f;
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 bb = *reinterpret_cast<const float4*>(B3 + c0 + 4 * j4);
        const float4 ww = *reinterpret_cast<const float4*>(W4 + c0 + 4 * j4);
        s = fmaf(ww.x, fmaxf(v[4 * j4 + 0] + bb.x, 0.f), s);
        s = fmaf(ww.y, fmaxf(v[4 * j4 + 1] + bb.y, 0.f), s);
        s = fmaf(ww.z, fmaxf(v[4 * j4 + 2] + bb.z, 0.f), s);
        s = fmaf(ww.w, fmaxf(v[4 * j4 + 3] + bb.w, 0.f), s);
      }
      hpart[cgrp * R + r] = s;
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    if (threadIdx.x < rows) {
      const int t = threadIdx.x;
      const float s = (hpart[t] + hpart[R + t]) + (hpart[2 * R + t] + hpart[3 * R + t]);
      h[a0 * K + t] = (s + B4[0]) * mask[t];
    }
    // next tile: threads 0..127 rewrite mask[]/read hpart[] in program order; everyone else first
    // meets a __syncthreads before hpart is written again
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int cbf_forward_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                   int M, int K, double dist_min, double obs_radius, float* h, cudaStream_t st) {
  const long long agents = (long long)B * N;
  const int TA = max(1, cbftc::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_forward_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cbftc::smem_bytes));
  cbf_forward_tc_kernel<<<grid, cbftc::THREADS, cbftc::smem_bytes, st>>>(states, obstacles, idx, params, N, M, K, agents,
                                                                        TA, (float)dist_min, (float)obs_radius, h);
  MACBF_CHECK_LAUNCH("cbf_forward_tc_kernel");
  return 0;
}

}  // namespace macbf
