// K3a (tensor-core path) - controller edge net + masked max pooling (controller.py:33-46,
// REPAIRED) on tcgen05 / TMEM.
//
// One persistent CTA per SM: 16 worker warps + 1 MMA-issue warp, two tiles in flight (TMEM and
// the pooling stage are double-buffered), so the tensor core computes tile i+1 while the CUDA
// cores post-process tile i and prepare tile i+2.  A tile is 128 edge rows = floor(128/K) agents.
//
//   workers   features + layer 1 (5->64, FFMA) -> tf32 hi/lo -> tcgen05.st   A2[b] (TMEM)
//   MMA warp  layer 2: D2[b][128x128] = A2[b] * W2^T, 3xTF32 (cross terms first)
//   workers   tcgen05.ld, bias + ReLU + * mask -> stage[b] (smem, [channel][row])
//             max over the K slots of every agent -> pool [agents,128] (global)
//
// TMEM columns of buffer b: 256 b + [0,64) A2_hi, [64,128) A2_lo, [128,256) D2.
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

namespace cetc {
constexpr int WORKERS = 512;
constexpr int THREADS = WORKERS + 32;
constexpr int R = 128;
constexpr int RS = 132;   // stage row stride: 4 (mod 32) words -> conflict-free LDS.128 across channels
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_D2 = 128, C_BUF = 256;
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW1 = oW2LO + 128 * 64;     // [8][64]: rows 0..4 = W1[:,k], row 5 = 0, row 6 = bias, row 7 = 0
constexpr int oB2 = oW1 + 64 * 8;
constexpr int oStage = oB2 + 128;          // [2][128][RS]
constexpr int total_floats = oStage + 2 * 128 * RS;
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
  float* W1 = sm + oW1; float* B2 = sm + oB2; float* stage_all = sm + oStage;
  __shared__ uint64_t a2_full[2], d2_full[2];   // workers -> MMA (16 arrivals) / MMA -> workers (commit)
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&a2_full[0], 16); mbar_init(&a2_full[1], 16);
    mbar_init(&d2_full[0], 1); mbar_init(&d2_full[1], 1);
    fence_mbar_init();
  }
  fill_weight_sw128<128, 64>(params + CT_EW2, 64, W2hi, W2lo);
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) {
    const int k = t >> 6, n = t & 63;
    W1[t] = (k < 5) ? __ldg(params + CT_EW1 + n * 5 + k) : (k == 6 ? __ldg(params + CT_EB1 + n) : 0.f);
  }
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long ntiles = (agents_total + TA - 1) / TA;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC = idesc_tf32(128, 128);
    int it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      mbar_wait(&a2_full[b], (it >> 1) & 1);
      fence_after_sync();
      const uint32_t t0 = tb + b * C_BUF;
#pragma unroll
      for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi
        const uint32_t a_col = t0 + (term == 0 ? C_A2LO : C_A2HI);
        const uint64_t b_desc = (term == 1) ? d_w2lo : d_w2hi;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (128 * 128) + (ks & 3) * 32);
          if (elect_one()) mma_tf32_ts(t0 + C_D2, a_col + ks * 8, bd, IDESC, (term | ks) ? 1u : 0u);
        }
      }
      if (elect_one()) mma_commit(&d2_full[b]);
    }
  } else {
    // ================================= worker warps =================================
    const int q = warp & 3, cgrp = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int a_loc = r / K;

    int pf_j = 0, pf_b = 0, pf_i = 0;
    bool pf_valid = false;
    float4 pf_si = make_float4(0.f, 0.f, 0.f, 0.f), pf_sj = pf_si;
    auto prefetch_ij = [&](long long t) {
      pf_valid = false;
      if (t < ntiles) {
        const int ga = (int)(t * TA) + a_loc;      // B*N < 2^31 is checked by the host
        if (a_loc < TA && ga < (int)agents_total) {
          pf_b = ga / N; pf_i = ga - pf_b * N;
          pf_j = __ldg(idx + (size_t)(t * TA) * K + r);
          pf_si = __ldg(reinterpret_cast<const float4*>(states) + ga);
          pf_valid = true;
        }
      }
    };
    auto prefetch_sj = [&]() {
      if (pf_valid) pf_sj = load_candidate(states, obstacles, pf_b, pf_j, N, M);
    };
    // features (controller.py:33-40) + layer 1 for this thread's 16 channels -> A2[b]; returns the mask
    auto layer1 = [&](int b) -> float {
      float f[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (pf_valid) {
        const EdgeGeom g = edge_geom(pf_si, pf_sj, pf_i, pf_j);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        mk = (ctrl_dist(g.dx, g.dy) < obs_radius) ? 1.f : 0.f;   // controller.py:38-39
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const float* wb = W1 + cgrp * 16 + half * 8;
        float2 acc[4];
        {
          const float4 b0 = *reinterpret_cast<const float4*>(wb + 6 * 64), b1 = *reinterpret_cast<const float4*>(wb + 6 * 64 + 4);
          acc[0] = make_float2(b0.x, b0.y); acc[1] = make_float2(b0.z, b0.w);
          acc[2] = make_float2(b1.x, b1.y); acc[3] = make_float2(b1.z, b1.w);
        }
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const float4 w0 = *reinterpret_cast<const float4*>(wb + k * 64), w1 = *reinterpret_cast<const float4*>(wb + k * 64 + 4);
          const float2 fk = make_float2(f[k], f[k]);
          acc[0] = __ffma2_rn(make_float2(w0.x, w0.y), fk, acc[0]);
          acc[1] = __ffma2_rn(make_float2(w0.z, w0.w), fk, acc[1]);
          acc[2] = __ffma2_rn(make_float2(w1.x, w1.y), fk, acc[2]);
          acc[3] = __ffma2_rn(make_float2(w1.z, w1.w), fk, acc[3]);
        }
        float hi[8], lo[8];
#pragma unroll
        for (int p2 = 0; p2 < 4; ++p2) {
          float2 h2, l2;
          split2_trunc(relu2(acc[p2]), h2, l2);
          hi[2 * p2] = h2.x; hi[2 * p2 + 1] = h2.y; lo[2 * p2] = l2.x; lo[2 * p2 + 1] = l2.y;
        }
        tmem_st8(tlane + b * C_BUF + C_A2HI + cgrp * 16 + half * 8, hi);
        tmem_st8(tlane + b * C_BUF + C_A2LO + cgrp * 16 + half * 8, lo);
      }
      wait_st();
      warp_arrive_after_tmem(&a2_full[b]);
      return mk;
    };

    const long long g = gridDim.x;
    prefetch_ij(blockIdx.x);
    prefetch_sj();
    float mk0 = layer1(0), mk1 = 0.f;
    prefetch_ij(blockIdx.x + g);
    prefetch_sj();
    if (blockIdx.x + g < ntiles) mk1 = layer1(1);
    prefetch_ij(blockIdx.x + 2 * g);

    int it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += g, ++it) {
      const int b = it & 1;
      const long long a0 = tile * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      float* stage = stage_all + b * (128 * RS);
      prefetch_sj();                                   // endpoints of tile + 2g (idx arrived last iteration)
      mbar_wait(&d2_full[b], (it >> 1) & 1);
      fence_after_sync();
      // ---- bias + ReLU + mask on this thread's 32 columns -> stage[c][r] ----
#pragma unroll
      for (int part = 0; part < 4; ++part) {
        const int c0 = cgrp * 32 + part * 8;
        float v[8];
        tmem_ld8(tlane + b * C_BUF + C_D2 + c0, v);
        wait_ld();
        const float2 mk2 = make_float2(mk0, mk0);
#pragma unroll
        for (int j4 = 0; j4 < 2; ++j4) {
          const float4 bb = *reinterpret_cast<const float4*>(B2 + c0 + 4 * j4);
          const float2 y0 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 0], v[4 * j4 + 1]), make_float2(bb.x, bb.y))), mk2);
          const float2 y1 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 2], v[4 * j4 + 3]), make_float2(bb.z, bb.w))), mk2);
          stage[(c0 + 4 * j4 + 0) * RS + r] = y0.x;
          stage[(c0 + 4 * j4 + 1) * RS + r] = y0.y;
          stage[(c0 + 4 * j4 + 2) * RS + r] = y1.x;
          stage[(c0 + 4 * j4 + 3) * RS + r] = y1.y;
        }
      }
      fence_before_sync();
      named_bar_sync(1, WORKERS);                      // stage[b] complete
      // ---- max over the K slots (controller.py:46, repair F2: values) ----
      for (int item = threadIdx.x; item < na * 128; item += WORKERS) {
        const int c = item & 127, a = item >> 7;
        const float* row = stage + c * RS + a * K;
        float best = 0.f;                              // relu(.) * mask >= 0; slot 0 (self) is never masked
        if ((K & 3) == 0) {
          for (int k = 0; k < K; k += 4) {
            const float4 x = *reinterpret_cast<const float4*>(row + k);
            best = fmaxf(fmaxf(best, fmaxf(x.x, x.y)), fmaxf(x.z, x.w));
          }
        } else {
          for (int k = 0; k < K; ++k) best = fmaxf(best, row[k]);
        }
        pool[(a0 + a) * 128 + c] = best;
      }
      // ---- D2[b] and A2[b] are free: features + layer 1 of tile + 2g, then fetch idx of tile + 3g ----
      mk0 = mk1;
      if (tile + 2 * g < ntiles) mk1 = layer1(b);
      prefetch_ij(tile + 3 * g);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "controller_forward: B*N must be < 2^31");
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
