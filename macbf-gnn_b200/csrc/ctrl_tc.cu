// K3a (tensor-core path) - controller edge net + masked max pooling (controller.py:33-46,
// REPAIRED) on tcgen05 / TMEM.
//
// One persistent CTA per SM running TWO independent tile pipelines ("groups") that share the
// weights in shared memory and the tensor core: group g = 8 worker warps + 1 MMA-issue warp + half
// of TMEM, and takes every second tile of the CTA.  Inside a group a tile's phases are serial
// (MMA -> epilogue -> MMA -> epilogue -> pooling); the two groups sit at different phases, so
// while one waits for the tensor core or a barrier the other keeps the issue slots busy.
// A tile is 128 edge rows = floor(128/K) agents (TMEM lane = row).  Both dense layers are 3xTF32
// MMAs (cross terms first):
//
//   feature warps (4/group)  edge features (controller.py:33-40) -> tf32 hi/lo -> smem A1 [128 x 8]
//   MMA warp                 layer 1: D1[128x64]  = A1 * W1^T   (A and B from smem, K = 8)
//   workers                  bias + ReLU, hi/lo; hi overwrites D1 in place           -> A2 (TMEM)
//   MMA warp                 layer 2: D2[128x128] = A2 * W2^T   (A from TMEM)
//   workers                  bias + ReLU + * mask -> stage (smem, 64 channels at a time),
//                            max over the K slots of every agent (controller.py:46)  -> pool (global)
//
// TMEM columns of group g: 256 g + [0,64) D1 -> A2_hi, [64,128) A2_lo, [128,256) D2.
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
constexpr int GROUP_THREADS = 256;           // 8 worker warps per group
constexpr int WORKERS = 2 * GROUP_THREADS;
constexpr int THREADS = WORKERS + 64;        // + one MMA-issue warp per group
constexpr int R = 128;
constexpr int RS = 132;   // stage row stride: 4 (mod 32) words -> conflict-free LDS.128 across channels
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_D2 = 128, C_GRP = 256;
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW1HI = oW2LO + 128 * 64;      // [64 rows][32]: k = 0..4 used
constexpr int oW1LO = oW1HI + 64 * 32;
constexpr int oA1 = oW1LO + 64 * 32;         // [group][hi, lo][128 rows][32]
constexpr int oB1 = oA1 + 2 * 2 * 128 * 32;
constexpr int oB2 = oB1 + 64;
constexpr int oMask = oB2 + 128;             // [group][4][128] radius masks of the tiles in flight (tile ordinal & 3)
constexpr int oStage = oMask + 2 * 4 * 128;  // [group][64 channels][RS]
constexpr int total_floats = oStage + 2 * 64 * RS;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cetc

// WIN = true (backward pass): also writes the first arg-max slot of every (agent, channel) to `argk`.
template <bool WIN>
__global__ void __launch_bounds__(cetc::THREADS, 1)
ctrl_edge_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                    const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int K,
                    long long agents_total, int TA, float obs_radius, float* __restrict__ pool,
                    const int* __restrict__ skip, uint8_t* __restrict__ argk) {
  MACBF_SKIP_IF_SET(skip);
  using namespace cetc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO;
  float* W1hi = sm + oW1HI; float* W1lo = sm + oW1LO;
  float* B1 = sm + oB1; float* B2 = sm + oB2;
  // per group: workers -> MMA: a1_full (4 arrivals: feature warps), a2_full (8); MMA -> workers: d1_full, d2_full
  __shared__ uint64_t a1_full[2], a2_full[2], d1_full[2], d2_full[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    for (int gi = 0; gi < 2; ++gi) {
      mbar_init(&a1_full[gi], 4); mbar_init(&a2_full[gi], 8);
      mbar_init(&d1_full[gi], 1); mbar_init(&d2_full[gi], 1);
    }
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
  for (int t = threadIdx.x; t < 2 * 2 * 128 * 32; t += THREADS) sm[oA1 + t] = 0.f;
  for (int t = threadIdx.x; t < 64; t += THREADS) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long ntiles = (agents_total + TA - 1) / TA;
  const long long gdim = gridDim.x;
  [[maybe_unused]] int trace_n = 0;

  if (warp >= WORKERS / 32) {
    // ========================= MMA-issue warp of group gi =========================
    const int gi = warp - WORKERS / 32;
    float* A1hi = sm + oA1 + gi * (2 * 128 * 32);
    float* A1lo = A1hi + 128 * 32;
    const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1hi)), d_w1lo = smem_desc_sw128(smem_u32(W1lo));
    const uint64_t d_a1hi = smem_desc_sw128(smem_u32(A1hi)), d_a1lo = smem_desc_sw128(smem_u32(A1lo));
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
    const uint32_t tg = __shfl_sync(0xffffffffu, tbase, 0) + gi * C_GRP;
    constexpr uint32_t IDESC1 = idesc_tf32(128, 64), IDESC2 = idesc_tf32(128, 128);
    auto issue_layer1 = [&](int j) {                 // layer 1 of the group's j-th tile
      mbar_wait(&a1_full[gi], j & 1);
      fence_after_sync();
      if (elect_one()) mma_tf32_ss(tg + C_A2HI, d_a1lo, d_w1hi, IDESC1, 0u);
      if (elect_one()) mma_tf32_ss(tg + C_A2HI, d_a1hi, d_w1lo, IDESC1, 1u);
      if (elect_one()) mma_tf32_ss(tg + C_A2HI, d_a1hi, d_w1hi, IDESC1, 1u);
      if (elect_one()) mma_commit(&d1_full[gi]);
    };
    const long long first = blockIdx.x + gi * gdim;
    if (first < ntiles) issue_layer1(0);
    int j = 0;
    for (long long tile = first; tile < ntiles; tile += 2 * gdim, ++j) {
      mbar_wait(&a2_full[gi], j & 1);
      fence_after_sync();
      if (gi == 0) CE_TRACE(1, 1);
#pragma unroll
      for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi
        const uint32_t a_col = tg + (term == 0 ? C_A2LO : C_A2HI);
        const uint64_t b_desc = (term == 1) ? d_w2lo : d_w2hi;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (128 * 128) + (ks & 3) * 32);
          if (elect_one()) mma_tf32_ts(tg + C_D2, a_col + ks * 8, bd, IDESC2, (term | ks) ? 1u : 0u);
        }
      }
      if (elect_one()) mma_commit(&d2_full[gi]);
      if (gi == 0) CE_TRACE(1, 2);
      // layer 1 of the group's next tile: D1 lands in the A2 columns layer 2 above has consumed (in order)
      if (tile + 2 * gdim < ntiles) issue_layer1(j + 1);
      if (gi == 0) CE_TRACE(1, 3);
    }
  } else {
    // ============================ worker warps of group gi ============================
    const int gi = warp >> 3, wl = warp & 7;
    const int q = wl & 3, cg = wl >> 2;              // TMEM lane quarter, column half
    const int gtid = threadIdx.x & (GROUP_THREADS - 1);
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + gi * C_GRP + ((uint32_t)(q * 32) << 16);
    const int a_loc = r / K;
    float* A1hi = sm + oA1 + gi * (2 * 128 * 32);
    float* A1lo = A1hi + 128 * 32;
    float* mask = sm + oMask + gi * (4 * 128);
    float* stage = sm + oStage + gi * (64 * RS);
    const int bar_id = 1 + gi;
    const long long first = blockIdx.x + gi * gdim, step = 2 * gdim;

    int pf_j = 0, pf_b = 0, pf_i = 0;
    bool pf_valid = false;
    float4 pf_si = make_float4(0.f, 0.f, 0.f, 0.f), pf_sj = pf_si;
    auto prefetch_ij = [&](long long t) {
      pf_valid = false;
      if (cg == 0 && t < ntiles) {
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
    // features (controller.py:33-40) of this thread's row of the group's j-th tile -> A1 (smem, swizzled)
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
      mask[(j & 3) * 128 + r] = mk;                     // read by every warp's layer-2 epilogue of that tile
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a1_full[gi]);
    };
    // layer-1 epilogue of the group's j-th tile: D1 -> bias + ReLU + split -> A2 (32 columns per thread)
    auto epilogue1 = [&](int j) {
      mbar_wait(&d1_full[gi], j & 1);
      fence_after_sync();
#pragma unroll
      for (int h16 = 0; h16 < 2; ++h16) {
        const int c0 = cg * 32 + h16 * 16;
        float v[16];
        tmem_ld16(tlane + C_A2HI + c0, v);
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
          tmem_st8(tlane + C_A2HI + c0 + half * 8, hi);
          tmem_st8(tlane + C_A2LO + c0 + half * 8, lo);
        }
      }
      wait_st();
      warp_arrive_after_tmem(&a2_full[gi]);
    };

    if (first < ntiles) {
      // prologue: the group's tile 0 through layer 1, features of tile 1 staged
      prefetch_ij(first);
      prefetch_sj();
      if (cg == 0) features(0);
      epilogue1(0);                                   // also proves layer 1 of tile 0 has read A1
      prefetch_ij(first + step);
      prefetch_sj();
      if (cg == 0 && first + step < ntiles) features(1);
      prefetch_ij(first + 2 * step);
    }

    int j = 0;
    for (long long tile = first; tile < ntiles; tile += step, ++j) {
      const long long a0 = tile * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      const bool has_next = tile + step < ntiles;
      prefetch_sj();                                   // endpoints of the tile after next (idx arrived last iteration)
      mbar_wait(&d2_full[gi], j & 1);
      fence_after_sync();
      if (warp == 0) CE_TRACE(0, 1); else if (warp == 4) CE_TRACE(2, 1);
      const float mk = mask[(j & 3) * 128 + r];
      const float2 mk2 = make_float2(mk, mk);
#pragma unroll 1
      for (int hf = 0; hf < 2; ++hf) {                 // 64 output channels at a time through the stage
        named_bar_sync(bar_id, GROUP_THREADS);         // the previous pooling pass has finished reading `stage`
        {
          float v[32];
          float4 bias[8];                              // issued before the TMEM wait so their latency is hidden
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) bias[j4] = *reinterpret_cast<const float4*>(B2 + hf * 64 + cg * 32 + 4 * j4);
#pragma unroll
          for (int part = 0; part < 4; ++part) tmem_ld8(tlane + C_D2 + hf * 64 + cg * 32 + part * 8, v + part * 8);
          wait_ld();
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const int cl = cg * 32 + 4 * j4;           // channel within this half
            const float4 bb = bias[j4];
            const float2 y0 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 0], v[4 * j4 + 1]), make_float2(bb.x, bb.y))), mk2);
            const float2 y1 = __fmul2_rn(relu2(__fadd2_rn(make_float2(v[4 * j4 + 2], v[4 * j4 + 3]), make_float2(bb.z, bb.w))), mk2);
            stage[(cl + 0) * RS + r] = y0.x;
            stage[(cl + 1) * RS + r] = y0.y;
            stage[(cl + 2) * RS + r] = y1.x;
            stage[(cl + 3) * RS + r] = y1.y;
          }
        }
        fence_before_sync();
        if (warp == 0) CE_TRACE(0, 2 + 4 * hf); else if (warp == 4) CE_TRACE(2, 2 + 4 * hf);
        named_bar_sync(bar_id, GROUP_THREADS);         // stage complete (and, for hf = 1, D2 fully read)
        if (warp == 0) CE_TRACE(0, 3 + 4 * hf); else if (warp == 4) CE_TRACE(2, 3 + 4 * hf);
        // layer-1 epilogue of the next tile first, so its layer 2 can start while we pool
        if (hf == 1 && has_next) epilogue1(j + 1);
        if (warp == 0) CE_TRACE(0, 4 + 4 * hf); else if (warp == 4) CE_TRACE(2, 4 + 4 * hf);
        // ---- masked max over the K slots (controller.py:46, repair F2: values).  First half: all 8
        //      warps.  Second half: the 4 non-feature warps, while the feature warps build the A1 tile
        //      of the tile after next (layer 1 of the next tile has retired: d1_full observed above) ----
        const bool feat_now = (hf == 1) && (tile + 2 * step < ntiles);
        if (feat_now && cg == 0) features(j + 2);
        const int p_first = feat_now ? gtid - GROUP_THREADS / 2 : gtid;
        const int p_step = feat_now ? GROUP_THREADS / 2 : GROUP_THREADS;
        for (int item = p_first; item < na * 64 && item >= 0; item += p_step) {
          const int c = item & 63, a = item >> 6;
          const float* row = stage + c * RS + a * K;
          float best = 0.f;                            // relu(.) * mask >= 0; slot 0 (self) is never masked
          if constexpr (WIN) {
            // backward pass: also the FIRST arg-max slot (torch.max backward); 255 = no gradient (winner is 0)
            int bi = 0;
            best = row[0];
            for (int k = 1; k < K; ++k)
              if (row[k] > best) { best = row[k]; bi = k; }
            argk[(a0 + a) * 128 + hf * 64 + c] = (best > 0.f) ? (uint8_t)bi : (uint8_t)255;
          } else if (K == 12) {                        // the reference's TOP_K: three independent 16-byte loads
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
          pool[(a0 + a) * 128 + hf * 64 + c] = best;
        }
        if (warp == 0) CE_TRACE(0, 5 + 4 * hf); else if (warp == 4) CE_TRACE(2, 5 + 4 * hf);
      }
      prefetch_ij(tile + 3 * step);
      if (warp == 0) CE_TRACE(0, 10);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int ctrl_edge_tc_v1(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                    int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "controller_forward: B*N must be < 2^31");
  const int TA = max(1, cetc::R / K);
  const long long ntiles = (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  if (argk != nullptr) {
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)cetc::smem_bytes));
    ctrl_edge_tc_kernel<true><<<grid, cetc::THREADS, cetc::smem_bytes, st>>>(states, obstacles, idx, params, N, M, K, agents,
                                                                            TA, (float)obs_radius, pool, tl_skip_flag, argk);
  } else {
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)cetc::smem_bytes));
    ctrl_edge_tc_kernel<false><<<grid, cetc::THREADS, cetc::smem_bytes, st>>>(states, obstacles, idx, params, N, M, K, agents,
                                                                             TA, (float)obs_radius, pool, tl_skip_flag, argk);
  }
  MACBF_CHECK_LAUNCH("ctrl_edge_tc_kernel");
  return 0;
}

}  // namespace macbf
