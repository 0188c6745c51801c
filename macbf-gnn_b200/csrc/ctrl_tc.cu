// K3a (tensor-core path) - controller edge net + masked max pooling (controller.py:33-46,
// REPAIRED) on tcgen05 / TMEM.
//
// One persistent CTA per SM: 16 worker warps + 1 MMA-issue warp, two tiles in flight (the TMEM
// tile is double-buffered), so the tensor core computes tile i+1 while the CUDA cores post-process
// tile i and prepare tile i+2.  A tile is 128 edge rows = floor(128/K) agents.  Both dense layers
// run on the tensor core as 3xTF32 MMAs (cross terms first); the CUDA cores build the 5 edge
// features and run the epilogues:
//
//   workers(4 warps)  edge features (controller.py:33-40) -> tf32 hi/lo -> smem A1 [128 x 8]
//   MMA warp          layer 1: D1[b][128x64]  = A1 * W1^T    (A and B from smem, K = 8)
//   workers           bias + ReLU, hi/lo; hi overwrites D1 in place              -> A2[b] (TMEM)
//   MMA warp          layer 2: D2[b][128x128] = A2[b] * W2^T (A from TMEM)
//   workers           bias + ReLU + * mask -> stage (smem, [channel][row])
//                     max over the K slots of every agent                         -> pool (global)
//
// MMA issue order: L2(i), L1(i+2), L2(i+1), L1(i+3), ...
// TMEM columns of buffer b: 256 b + [0,64) D1 -> A2_hi, [64,128) A2_lo, [128,256) D2.
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

#ifdef MACBF_TRACE
__device__ unsigned long long g_ce_trace[3][4096];
#define CE_TRACE(who, ev)                                                                          \
  do {                                                                                             \
    if (blockIdx.x == 0 && lane == 0 && trace_n < 4096)                                            \
      g_ce_trace[who][trace_n++] = ((unsigned long long)(ev) << 56) | (clock64() & 0x00ffffffffffffffull); \
  } while (0)
extern "C" int macbf_debug_ce_trace(unsigned long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_ce_trace, sizeof(g_ce_trace));
}
#else
#define CE_TRACE(who, ev) do { } while (0)
#endif

namespace cetc {
constexpr int WORKERS = 512;
constexpr int THREADS = WORKERS + 32;
constexpr int R = 128;
constexpr int RS = 132;   // stage row stride: 4 (mod 32) words -> conflict-free LDS.128 across channels
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_D2 = 128, C_BUF = 256;
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW1HI = oW2LO + 128 * 64;    // [64 rows][32]: k = 0..4 used
constexpr int oW1LO = oW1HI + 64 * 32;
constexpr int oA1HI = oW1LO + 64 * 32;     // [128 rows][32]
constexpr int oA1LO = oA1HI + 128 * 32;
constexpr int oB1 = oA1LO + 128 * 32;
constexpr int oB2 = oB1 + 64;
constexpr int oMask = oB2 + 128;           // [4][128] radius masks of the tiles in flight (tile ordinal & 3)
constexpr int oStage = oMask + 4 * 128;    // [128][RS]
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
  float* W1hi = sm + oW1HI; float* W1lo = sm + oW1LO;
  float* A1hi = sm + oA1HI; float* A1lo = sm + oA1LO;
  float* B1 = sm + oB1; float* B2 = sm + oB2; float* mask = sm + oMask; float* stage = sm + oStage;
  // workers -> MMA: a1_full (4 arrivals: feature warps), a2_full[2] (16);  MMA -> workers: d1_full[2], d2_full[2]
  // (per-buffer barriers advance every second tile, so a lagging warp can never be lapped by two phases)
  __shared__ uint64_t a1_full, a2_full[2], d1_full[2], d2_full[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&a1_full, 4); mbar_init(&a2_full[0], 16); mbar_init(&a2_full[1], 16);
    mbar_init(&d1_full[0], 1); mbar_init(&d1_full[1], 1); mbar_init(&d2_full[0], 1); mbar_init(&d2_full[1], 1);
    fence_mbar_init();
  }
  fill_weight_sw128<128, 64>(params + CT_EW2, 64, W2hi, W2lo);
  for (int t = threadIdx.x; t < 64 * 32; t += THREADS) {   // W1 (64 x 5): chunk c of row n at position c ^ (n & 7)
    const int n = t >> 5, pos = (t >> 2) & 7, e = t & 3;
    const int k = ((pos ^ (n & 7)) << 2) + e;
    float hi = 0.f, lo = 0.f;
    if (k < 5) tf32_split(__ldg(params + CT_EW1 + n * 5 + k), hi, lo);
    W1hi[t] = hi; W1lo[t] = lo;
  }
  for (int t = threadIdx.x; t < 128 * 32; t += THREADS) { A1hi[t] = 0.f; A1lo[t] = 0.f; }
  for (int t = threadIdx.x; t < 64; t += THREADS) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long ntiles = (agents_total + TA - 1) / TA;
  const long long g = gridDim.x;
  [[maybe_unused]] int trace_n = 0;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1hi)), d_w1lo = smem_desc_sw128(smem_u32(W1lo));
    const uint64_t d_a1hi = smem_desc_sw128(smem_u32(A1hi)), d_a1lo = smem_desc_sw128(smem_u32(A1lo));
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC1 = idesc_tf32(128, 64), IDESC2 = idesc_tf32(128, 128);
    // layer 1 of local tile ordinal j (TMEM buffer j & 1); a1_full / d1_full complete once per tile
    auto issue_layer1 = [&](int j) {
      mbar_wait(&a1_full, j & 1);
      fence_after_sync();
      const uint32_t d1 = tb + (j & 1) * C_BUF + C_A2HI;
      if (elect_one()) mma_tf32_ss(d1, d_a1lo, d_w1hi, IDESC1, 0u);
      if (elect_one()) mma_tf32_ss(d1, d_a1hi, d_w1lo, IDESC1, 1u);
      if (elect_one()) mma_tf32_ss(d1, d_a1hi, d_w1hi, IDESC1, 1u);
      if (elect_one()) mma_commit(&d1_full[j & 1]);
    };
    issue_layer1(0);
    if (blockIdx.x + g < ntiles) issue_layer1(1);
    int it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += g, ++it) {
      const int b = it & 1;
      mbar_wait(&a2_full[b], (it >> 1) & 1);
      fence_after_sync();
      CE_TRACE(1, 1);
      const uint32_t t0 = tb + b * C_BUF;
#pragma unroll
      for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi
        const uint32_t a_col = t0 + (term == 0 ? C_A2LO : C_A2HI);
        const uint64_t b_desc = (term == 1) ? d_w2lo : d_w2hi;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (128 * 128) + (ks & 3) * 32);
          if (elect_one()) mma_tf32_ts(t0 + C_D2, a_col + ks * 8, bd, IDESC2, (term | ks) ? 1u : 0u);
        }
      }
      if (elect_one()) mma_commit(&d2_full[b]);
      CE_TRACE(1, 2);
      // layer 1 of tile + 2g goes into this buffer's A2 columns (layer 2 above has consumed them in order)
      if (tile + 2 * g < ntiles) issue_layer1(it + 2);
      CE_TRACE(1, 3);
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
      if (cgrp == 0 && t < ntiles) {
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
    // features (controller.py:33-40) of this thread's row -> A1 (smem, swizzled); returns the radius mask
    auto features = [&](int j) {
      float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (pf_valid) {
        const EdgeGeom e = edge_geom(pf_si, pf_sj, pf_i, pf_j);
        f[0] = e.dx; f[1] = e.dy; f[2] = e.dvx; f[3] = e.dvy; f[4] = e.eye;
        mk = (ctrl_dist(e.dx, e.dy) < obs_radius) ? 1.f : 0.f;   // controller.py:38-39
      }
      float hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) tf32_split_trunc(f[e], hi[e], lo[e]);
      const int p0 = (0 ^ (r & 7)) << 2, p1 = (1 ^ (r & 7)) << 2;
      *reinterpret_cast<float4*>(A1hi + r * 32 + p0) = make_float4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<float4*>(A1hi + r * 32 + p1) = make_float4(hi[4], hi[5], hi[6], hi[7]);
      *reinterpret_cast<float4*>(A1lo + r * 32 + p0) = make_float4(lo[0], lo[1], lo[2], lo[3]);
      *reinterpret_cast<float4*>(A1lo + r * 32 + p1) = make_float4(lo[4], lo[5], lo[6], lo[7]);
      mask[(j & 3) * 128 + r] = mk;                     // read by every warp's layer-2 epilogue of tile j
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a1_full);
    };
    // layer-1 epilogue of local tile ordinal j: D1 -> bias + ReLU + split -> A2[j & 1]
    auto epilogue1 = [&](int j) {
      mbar_wait(&d1_full[j & 1], (j >> 1) & 1);
      fence_after_sync();
      const uint32_t tl = tlane + (j & 1) * C_BUF;
      const int c0 = cgrp * 16;
      float v[16];
      tmem_ld16(tl + C_A2HI + c0, v);
      wait_ld();
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float hi[8], lo[8];
#pragma unroll
        for (int j4 = 0; j4 < 2; ++j4) {
          const float4 bb = *reinterpret_cast<const float4*>(B1 + c0 + half * 8 + 4 * j4);
          float2 h2, l2;
          split2_trunc(relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 0], v[half * 8 + 4 * j4 + 1]), make_float2(bb.x, bb.y))), h2, l2);
          hi[4 * j4 + 0] = h2.x; hi[4 * j4 + 1] = h2.y; lo[4 * j4 + 0] = l2.x; lo[4 * j4 + 1] = l2.y;
          split2_trunc(relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 2], v[half * 8 + 4 * j4 + 3]), make_float2(bb.z, bb.w))), h2, l2);
          hi[4 * j4 + 2] = h2.x; hi[4 * j4 + 3] = h2.y; lo[4 * j4 + 2] = l2.x; lo[4 * j4 + 3] = l2.y;
        }
        tmem_st8(tl + C_A2HI + c0 + half * 8, hi);
        tmem_st8(tl + C_A2LO + c0 + half * 8, lo);
      }
      wait_st();
      warp_arrive_after_tmem(&a2_full[j & 1]);
    };

    // prologue: tiles 0 and 1 through layer 1, features of tile 2 staged
    prefetch_ij(blockIdx.x);
    prefetch_sj();
    if (cgrp == 0) features(0);
    epilogue1(0);                                     // also proves layer 1 of tile 0 has read A1
    prefetch_ij(blockIdx.x + g);
    prefetch_sj();
    if (blockIdx.x + g < ntiles) {
      if (cgrp == 0) features(1);
      epilogue1(1);
    }
    prefetch_ij(blockIdx.x + 2 * g);
    prefetch_sj();
    if (cgrp == 0 && blockIdx.x + 2 * g < ntiles) features(2);
    prefetch_ij(blockIdx.x + 3 * g);

    int it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += g, ++it) {
      const int b = it & 1;
      const long long a0 = tile * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      prefetch_sj();                                   // endpoints of tile + 3g (idx arrived last iteration)
      mbar_wait(&d2_full[b], (it >> 1) & 1);
      fence_after_sync();
      if (warp == 0) CE_TRACE(0, 1); else if (warp == 15) CE_TRACE(2, 1);
      named_bar_sync(2, WORKERS);
      if (warp == 0) CE_TRACE(0, 2); else if (warp == 15) CE_TRACE(2, 2);                      // previous tile's pooling has finished reading `stage`
      // ---- bias + ReLU + mask on this thread's 32 columns -> stage[c][r] ----
      {
        const float mk = mask[(it & 3) * 128 + r];
        const float2 mk2 = make_float2(mk, mk);
        float v[32];
#pragma unroll
        for (int part = 0; part < 4; ++part) tmem_ld8(tlane + b * C_BUF + C_D2 + cgrp * 32 + part * 8, v + part * 8);
        wait_ld();
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const int c0 = cgrp * 32 + 4 * j4;
          const float4 bb = *reinterpret_cast<const float4*>(B2 + c0);
          const float2 y0 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 0], v[4 * j4 + 1]), make_float2(bb.x, bb.y))), mk2);
          const float2 y1 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 2], v[4 * j4 + 3]), make_float2(bb.z, bb.w))), mk2);
          stage[(c0 + 0) * RS + r] = y0.x;
          stage[(c0 + 1) * RS + r] = y0.y;
          stage[(c0 + 2) * RS + r] = y1.x;
          stage[(c0 + 3) * RS + r] = y1.y;
        }
      }
      fence_before_sync();
      if (warp == 0) CE_TRACE(0, 3); else if (warp == 15) CE_TRACE(2, 3);
      named_bar_sync(1, WORKERS);                      // stage complete; D2[b] fully read
      if (warp == 0) CE_TRACE(0, 4); else if (warp == 15) CE_TRACE(2, 4);
      // ---- layer-1 epilogue of tile + 2g into this buffer (its layer 1 was issued right after layer 2 above) ----
      if (tile + 2 * g < ntiles) epilogue1(it + 2);
      if (warp == 0) CE_TRACE(0, 5); else if (warp == 15) CE_TRACE(2, 5);
      if (cgrp != 0) {
        // ---- masked max over the K slots (controller.py:46, repair F2: values): 12 of the 16 warps ----
        for (int item = threadIdx.x - 128; item < na * 128; item += WORKERS - 128) {
          const int c = item & 127, a = item >> 7;
          const float* row = stage + c * RS + a * K;
          float best = 0.f;                            // relu(.) * mask >= 0; slot 0 (self) is never masked
          if (K == 12) {                               // the reference's TOP_K: three independent 16-byte loads
            const float4 x0 = *reinterpret_cast<const float4*>(row), x1 = *reinterpret_cast<const float4*>(row + 4),
                         x2 = *reinterpret_cast<const float4*>(row + 8);
            best = fmaxf(fmaxf(fmaxf(x0.x, x0.y), fmaxf(x0.z, x0.w)),
                         fmaxf(fmaxf(fmaxf(x1.x, x1.y), fmaxf(x1.z, x1.w)), fmaxf(fmaxf(x2.x, x2.y), fmaxf(x2.z, x2.w))));
          } else if ((K & 3) == 0) {
            for (int k = 0; k < K; k += 4) {
              const float4 x = *reinterpret_cast<const float4*>(row + k);
              best = fmaxf(fmaxf(best, fmaxf(x.x, x.y)), fmaxf(x.z, x.w));
            }
          } else {
            for (int k = 0; k < K; ++k) best = fmaxf(best, row[k]);
          }
          pool[(a0 + a) * 128 + c] = best;
        }
      }
      // ---- ... while the 4 feature warps build A1 of tile + 3g (layer 1 of tile + 2g has retired:
      //      d1_full was observed in epilogue1 above) ----
      if (cgrp == 0 && tile + 3 * g < ntiles) features(it + 3);
      prefetch_ij(tile + 4 * g);
      if (warp == 0) CE_TRACE(0, 6); else if (warp == 15) CE_TRACE(2, 6);
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
