// K3b (tensor-core path) - controller node net, gains, action and Euler step
// (controller.py:47-61, core.py:74-86, train.py:70) on tcgen05 / TMEM.
//
// A tile is 128 agents (TMEM lane = agent).  All three hidden layers run as 3xTF32 MMAs with the
// activations in tensor memory and the hi/lo halves of the weights resident in shared memory:
//
//   A0 = [pool(128), p-g (2), v (2), 0-pad (4)]  -> tcgen05.st (hi, lo)         K = 136
//   layer 1: D1[128x64]  = A0 * W1^T   (17 k-steps x 3 terms)
//   bias + ReLU + split                            -> A1 (reuses the A0 columns)
//   layer 2: D2[128x128] = A1 * W2^T;  bias + ReLU + split, hi in place -> A2
//   layer 3: D3[128x64]  = A2 * W3^T
//   bias + ReLU, 64 -> 4 (FFMA), gains 2*sigmoid+0.2, action, s' = s + [v,u] dt  -> u, s_next
//
// TMEM columns: A0_hi [0,136) A0_lo [136,272) D1 [272,336)   (layer 1)
//               A1_hi [0,64)  A1_lo [64,128)  D2/A2_hi [128,256)  A2_lo [256,384)  D3 [384,448)
// Each CTA handles only a few tiles (agents/128/148), so the phases of a tile run back to back
// (block barriers) instead of the decoupled pipeline of the edge kernels.
#include "common.cuh"
#include "tc_common.cuh"
#include "ctrl_node_bwd_layout.cuh"

namespace macbf {

// Optional phase timeline (build with -DMACBF_TRACE; tools/node_trace.py reads it back): CTA 0, warp 0.
#ifdef MACBF_TRACE
__device__ unsigned long long g_node_trace[4096];
#define MACBF_TRACE_MARK(who, ev)                                                                          \
  do {                                                                                                     \
    if (blockIdx.x == 0 && lane == 0 && trace_n < 4096)                                                    \
      g_node_trace[trace_n++] = ((unsigned long long)(ev) << 56) | (clock64() & 0x00ffffffffffffffull);    \
  } while (0)
extern "C" int macbf_debug_node_trace(unsigned long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_node_trace, sizeof(g_node_trace));
}
#else
#define MACBF_TRACE_MARK(who, ev) do { } while (0)
#endif

namespace cntc {
constexpr int THREADS = 512;
constexpr int R = 128;
constexpr int K1 = 136;                    // 132 inputs + 4 zero columns (k-step granularity 8)
constexpr int K1P = 160;                   // shared-memory K extent of W1 (5 blocks of 32)
constexpr uint32_t C_A0HI = 0, C_A0LO = 136, C_D1 = 272;
constexpr uint32_t C_A1HI = 0, C_A1LO = 64, C_Y = 128, C_A2LO = 256, C_D3 = 384;
constexpr int oW1HI = 0;
constexpr int oW1LO = oW1HI + 64 * K1P;
constexpr int oW2HI = oW1LO + 64 * K1P;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW3HI = oW2LO + 128 * 64;
constexpr int oW3LO = oW3HI + 64 * 128;
constexpr int oB1 = oW3LO + 64 * 128;
constexpr int oB2 = oB1 + 64;
constexpr int oB3 = oB2 + 128;
constexpr int oW4 = oB3 + 64;              // [4][64]
constexpr int oB4 = oW4 + 256;
constexpr int oPart = oB4 + 4;             // [4 cgrp][4 out][128]; DUMP: then [16 warps][64] partial dW4 sums
constexpr int oD4 = oPart + 16 * R;        // DUMP: d/d(pre-sigmoid) [4][128]
constexpr int total_floats = oD4 + 4 * R;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cntc

// DUMP = true is the first kernel of the tensor-core node backward (ctrl_node_bwd_tc.cu): the same forward, additionally
// writing per tile (ctrl_node_bwd_layout.cuh) the bf16 hi/lo images of pool, a1, a2, the ReLU masks, and per agent
// d/d(pre-sigmoid outputs) from the incoming gradients (controller.py:50-61 backwards) and the direct d/d(e, v) through
// the action; accumulates dW4, db4 into this CTA's gradient slice.  u / s_next are not written.
template <bool DUMP>
__global__ void __launch_bounds__(cntc::THREADS, 1)
ctrl_node_tc_kernel(const float* __restrict__ states, const float* __restrict__ goals,
                    const float* __restrict__ pool, const float* __restrict__ params, long long agents_total,
                    int agent_begin, float dt, float* __restrict__ u, float* __restrict__ s_next, const int* __restrict__ skip,
                    const float* __restrict__ grad_u, const float* __restrict__ grad_s_next,
                    unsigned char* __restrict__ rec, float* __restrict__ slices) {
  MACBF_SKIP_IF_SET(skip);
  using namespace cntc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W1hi = sm + oW1HI; float* W1lo = sm + oW1LO;
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO;
  float* W3hi = sm + oW3HI; float* W3lo = sm + oW3LO;
  float* B1 = sm + oB1; float* B2 = sm + oB2; float* B3 = sm + oB3; float* W4 = sm + oW4; float* B4 = sm + oB4;
  float* part = sm + oPart;
  [[maybe_unused]] float* d4s = sm + oD4;
  [[maybe_unused]] float acc4 = 0.f;         // DUMP: thread t < 256 owns dW4[t >> 6][t & 63]; 256..259 own db4
  __shared__ uint64_t bar[3];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cgrp = warp >> 2;
  const int r = q * 32 + lane;
  [[maybe_unused]] int trace_n = 0;
  if (warp == 0) MACBF_TRACE_MARK(0, 1);

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_init(&bar[2], 1);
    fence_mbar_init();
  }
  {   // all three weight matrices: every load in flight before the first split / store.  W1 (64 x 132) is zero-padded
      // to K = 160 in the SWIZZLE_128B K-major layout
    WeightFill<64, K1P, THREADS> f1;
    WeightFill<128, 64, THREADS> f2;
    WeightFill<64, 128, THREADS> f3;
    f1.template load<false>(params + CT_NW1, 132, 1, 132);
    f2.template load<false>(params + CT_NW2, 64, 1);
    f3.template load<false>(params + CT_NW3, 128, 1);
    f1.template store<false>(W1hi, W1lo);
    f2.template store<false>(W2hi, W2lo);
    f3.template store<false>(W3hi, W3lo);
  }
  for (int t = threadIdx.x; t < 64; t += THREADS) {
    B1[t] = __ldg(params + CT_NB1 + t);
    B3[t] = __ldg(params + CT_NB3 + t);
  }
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_NB2 + t);
  for (int t = threadIdx.x; t < 256; t += THREADS) W4[t] = __ldg(params + CT_NW4 + t);
  if (threadIdx.x < 4) B4[threadIdx.x] = __ldg(params + CT_NB4 + threadIdx.x);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
  const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
  const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1hi)), d_w1lo = smem_desc_sw128(smem_u32(W1lo));
  const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
  const uint64_t d_w3hi = smem_desc_sw128(smem_u32(W3hi)), d_w3lo = smem_desc_sw128(smem_u32(W3lo));
  constexpr uint32_t IDESC64 = idesc_tf32(128, 64), IDESC128 = idesc_tf32(128, 128);

  // tiles cover the agents [agent_begin, agents_total) (agent_begin > 0: a row range; never with DUMP)
  const long long ntiles = (agents_total - agent_begin + R - 1) / R;
  // This thread's inputs of a tile (32 pooled channels of agent row r; column group 3 also the state and the goal) are
  // requested one tile ahead, right after the previous tile's have been written to tensor memory: the 64 KB of a tile
  // then arrive under that tile's three GEMMs instead of in front of them.
  float4 pfp[8];
  float4 pf_s = make_float4(0.f, 0.f, 0.f, 0.f);
  float2 pf_g = make_float2(0.f, 0.f);
  auto load_tile = [&](long long tile) {
    if (tile >= ntiles) return;                     // warp-uniform
    // rows past the end read the last agent's row (never used: `valid` zeroes them when the tile is built) - an
    // unconditional load has no select hanging on it, so nothing waits for the data before the next tile needs it
    const long long ga = min(agent_begin + tile * R + r, agents_total - 1);
    const float4* prow = reinterpret_cast<const float4*>(pool + ga * 128 + cgrp * 32);
#pragma unroll
    for (int i = 0; i < 8; ++i) pfp[i] = __ldg(prow + i);
    if (cgrp == 3) {
      pf_s = __ldg(reinterpret_cast<const float4*>(states) + ga);
      pf_g = __ldg(reinterpret_cast<const float2*>(goals) + ga);
    }
  };
  if (warp == 0) MACBF_TRACE_MARK(0, 2);
  load_tile(blockIdx.x);
  uint32_t par = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, par ^= 1) {
    const long long a0 = agent_begin + tile * R;
    const bool valid = a0 + r < agents_total;
    [[maybe_unused]] unsigned char* tr = DUMP ? rec + (size_t)tile * cnrec::BYTES : nullptr;
    if (warp == 0) MACBF_TRACE_MARK(0, 3);
    // state and goal of this row for the head at the end of the tile (column group 0): requested now, so that the
    // head does not start with an L2 round trip while every other warp waits at the barrier behind it
    float4 hs = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 hg = make_float2(0.f, 0.f);
    if (cgrp == 0 && valid) {
      hs = __ldg(reinterpret_cast<const float4*>(states) + a0 + r);
      hg = __ldg(reinterpret_cast<const float2*>(goals) + a0 + r);
    }
    // ---- A0 = [pool, p - g, v, 0]: this thread's 32 pooled channels (+ the 8 tail columns) ----
    {
      [[maybe_unused]] uint4 pbh[2], pbl[2];
#pragma unroll
      for (int part8 = 0; part8 < 4; ++part8) {
        const float4 x0 = pfp[2 * part8], x1 = pfp[2 * part8 + 1];
        const float vz = valid ? 1.f : 0.f;           // rows past the end are exact zeros
        const float v[8] = {x0.x * vz, x0.y * vz, x0.z * vz, x0.w * vz, x1.x * vz, x1.y * vz, x1.z * vz, x1.w * vz};
        float hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) tf32_split_trunc(v[e], hi[e], lo[e]);
        tmem_st8(tlane + C_A0HI + cgrp * 32 + part8 * 8, hi);
        tmem_st8(tlane + C_A0LO + cgrp * 32 + part8 * 8, lo);
        if constexpr (DUMP) {      // this thread's 32 channels = pairs (cgrp & 1) * 2 + {0, 1} of 64-channel block cgrp >> 1
          bf16_split8(v, pbh[part8 & 1], pbl[part8 & 1]);
          if (part8 & 1) {
            img_store16(tr + cnrec::P_HI + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + (part8 >> 1), pbh[0], pbh[1]);
            img_store16(tr + cnrec::P_LO + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + (part8 >> 1), pbl[0], pbl[1]);
          }
        }
      }
      if (cgrp == 3) {
        float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (valid) {
          const float4 s = pf_s;
          const float2 g = pf_g;
          v[0] = xsub(s.x, g.x); v[1] = xsub(s.y, g.y); v[2] = s.z; v[3] = s.w;   // controller.py:47
        }
        float hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) tf32_split_trunc(v[e], hi[e], lo[e]);
        tmem_st8(tlane + C_A0HI + 128, hi);
        tmem_st8(tlane + C_A0LO + 128, lo);
      }
      wait_st();
    }
    if (warp == 0) MACBF_TRACE_MARK(0, 4);
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- layer 1 ----
      fence_after_sync();
      mma_3xtf32_ts_commit<K1 / 8>(tb + C_D1, tb + C_A0HI, tb + C_A0LO, d_w1hi, d_w1lo, (64 * 128), IDESC64, &bar[0]);
    }
    load_tile(tile + gridDim.x);                   // next tile's inputs, requested once layer 1 is on its way
    mbar_wait(&bar[0], par);
    fence_after_sync();
    if (warp == 0) MACBF_TRACE_MARK(0, 5);
    {  // bias + ReLU + split on this thread's 16 channels of D1 -> A1
      const int c0 = cgrp * 16;
      float v[16];
      tmem_ld16(tlane + C_D1 + c0, v);
      wait_ld();
      [[maybe_unused]] uint32_t bits = 0;
      [[maybe_unused]] uint4 bh[2], bl[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float hi[8], lo[8];
        [[maybe_unused]] float x[8];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          float2 h2, l2;
          const float2 xa = relu2(__fadd2_rn(make_float2(v[half * 8 + e], v[half * 8 + e + 1]),
                                             make_float2(B1[c0 + half * 8 + e], B1[c0 + half * 8 + e + 1])));
          split2_trunc(xa, h2, l2);
          hi[e] = h2.x; hi[e + 1] = h2.y; lo[e] = l2.x; lo[e + 1] = l2.y;
          if constexpr (DUMP) { x[e] = xa.x; x[e + 1] = xa.y; }
        }
        tmem_st8(tlane + C_A1HI + c0 + half * 8, hi);
        tmem_st8(tlane + C_A1LO + c0 + half * 8, lo);
        if constexpr (DUMP) {
#pragma unroll
          for (int e = 0; e < 8; ++e) bits |= (x[e] > 0.f ? 1u : 0u) << (half * 8 + e);
          bf16_split8(x, bh[half], bl[half]);
        }
      }
      if constexpr (DUMP) {
        img_store16(tr + cnrec::A1_HI, r, cgrp, bh[0], bh[1]);
        img_store16(tr + cnrec::A1_LO, r, cgrp, bl[0], bl[1]);
        reinterpret_cast<uint32_t*>(tr + cnrec::MASKS)[(0 * 4 + cgrp) * R + r] = bits;
      }
      wait_st();
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- layer 2 ----
      fence_after_sync();
      mma_3xtf32_ts_commit<8>(tb + C_Y, tb + C_A1HI, tb + C_A1LO, d_w2hi, d_w2lo, (128 * 128), IDESC128, &bar[1]);
    }
    if (warp == 0) MACBF_TRACE_MARK(0, 6);
    mbar_wait(&bar[1], par);
    fence_after_sync();
    if (warp == 0) MACBF_TRACE_MARK(0, 7);
    [[maybe_unused]] uint32_t bits2 = 0;
#pragma unroll
    for (int half16 = 0; half16 < 2; ++half16) {   // bias + ReLU + split on 32 channels of D2: hi in place, lo aside
      const int c0 = cgrp * 32 + half16 * 16;
      float v[16];
      tmem_ld16(tlane + C_Y + c0, v);
      wait_ld();
      [[maybe_unused]] uint4 bh[2], bl[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float hi[8], lo[8];
        [[maybe_unused]] float x[8];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          float2 h2, l2;
          const float2 xa = relu2(__fadd2_rn(make_float2(v[half * 8 + e], v[half * 8 + e + 1]),
                                             make_float2(B2[c0 + half * 8 + e], B2[c0 + half * 8 + e + 1])));
          split2_trunc(xa, h2, l2);
          hi[e] = h2.x; hi[e + 1] = h2.y; lo[e] = l2.x; lo[e + 1] = l2.y;
          if constexpr (DUMP) { x[e] = xa.x; x[e + 1] = xa.y; }
        }
        tmem_st8(tlane + C_Y + c0 + half * 8, hi);
        tmem_st8(tlane + C_A2LO + c0 + half * 8, lo);
        if constexpr (DUMP) {
#pragma unroll
          for (int e = 0; e < 8; ++e) bits2 |= (x[e] > 0.f ? 1u : 0u) << (half16 * 16 + half * 8 + e);
          bf16_split8(x, bh[half], bl[half]);
        }
      }
      if constexpr (DUMP) {
        img_store16(tr + cnrec::A2_HI + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + half16, bh[0], bh[1]);
        img_store16(tr + cnrec::A2_LO + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + half16, bl[0], bl[1]);
      }
    }
    if constexpr (DUMP) reinterpret_cast<uint32_t*>(tr + cnrec::MASKS)[(1 * 4 + cgrp) * R + r] = bits2;
    wait_st();
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- layer 3 ----
      fence_after_sync();
      mma_3xtf32_ts_commit<16>(tb + C_D3, tb + C_Y, tb + C_A2LO, d_w3hi, d_w3lo, (64 * 128), IDESC64, &bar[2]);
    }
    if (warp == 0) MACBF_TRACE_MARK(0, 8);
    mbar_wait(&bar[2], par);
    fence_after_sync();
    if (warp == 0) MACBF_TRACE_MARK(0, 9);
    {  // bias + ReLU and the 64 -> 4 layer over this thread's 16 channels
      const int c0 = cgrp * 16;
      float v[16];
      tmem_ld16(tlane + C_D3 + c0, v);
      wait_ld();
      float2 o01 = make_float2(0.f, 0.f), o23 = o01;   // the four outputs as two packed f32x2 accumulators (same roundings)
      [[maybe_unused]] uint32_t bits = 0;
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const float a = fmaxf(v[e] + B3[c0 + e], 0.f);
        if constexpr (DUMP) bits |= (a > 0.f ? 1u : 0u) << e;
        const float2 a2 = make_float2(a, a);
        o01 = __ffma2_rn(make_float2(W4[0 * 64 + c0 + e], W4[1 * 64 + c0 + e]), a2, o01);
        o23 = __ffma2_rn(make_float2(W4[2 * 64 + c0 + e], W4[3 * 64 + c0 + e]), a2, o23);
      }
      part[(cgrp * 4 + 0) * R + r] = o01.x; part[(cgrp * 4 + 1) * R + r] = o01.y;
      part[(cgrp * 4 + 2) * R + r] = o23.x; part[(cgrp * 4 + 3) * R + r] = o23.y;
      if constexpr (DUMP) reinterpret_cast<uint32_t*>(tr + cnrec::MASKS)[(2 * 4 + cgrp) * R + r] = bits;
    }
    fence_before_sync();
    __syncthreads();   // partial sums visible; every TMEM read of this tile is done
    fence_after_sync();
    if constexpr (DUMP) {
      // ---- head backward (controller.py:50-61): d/d(pre-sigmoid), direct d/d(e, v); then dW4, db4 ----
      if (cgrp == 0) {
        float d_o[4] = {0.f, 0.f, 0.f, 0.f}, dir[4] = {0.f, 0.f, 0.f, 0.f}, ev[4] = {0.f, 0.f, 0.f, 0.f};
        if (valid) {
          float kk[4], sg[4];
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            const float o = B4[qq] + ((part[qq * R + r] + part[(4 + qq) * R + r]) + (part[(8 + qq) * R + r] + part[(12 + qq) * R + r]));
            sg[qq] = 1.0f / (1.0f + expf(-o));
            kk[qq] = 2.0f * sg[qq] + 0.2f;
          }
          const float4 s = hs;
          const float2 g = hg;
          ev[0] = xsub(s.x, g.x); ev[1] = xsub(s.y, g.y); ev[2] = s.z; ev[3] = s.w;
          float gux = 0.f, guy = 0.f;
          if (grad_u != nullptr) {
            const float2 gu = __ldg(reinterpret_cast<const float2*>(grad_u) + a0 + r);
            gux = gu.x; guy = gu.y;
          }
          if (grad_s_next != nullptr) {   // s' = s + [v, u] * dt
            const float4 gsn = __ldg(reinterpret_cast<const float4*>(grad_s_next) + a0 + r);
            gux = fmaf(dt, gsn.z, gux); guy = fmaf(dt, gsn.w, guy);
          }
          // a_x = -(k1 e_x + k2 v_x), a_y = -(k3 e_y + k4 v_y)
          const float dk[4] = {-ev[0] * gux, -ev[2] * gux, -ev[1] * guy, -ev[3] * guy};
          dir[0] = -kk[0] * gux; dir[2] = -kk[1] * gux; dir[1] = -kk[2] * guy; dir[3] = -kk[3] * guy;
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) d_o[qq] = dk[qq] * 2.0f * sg[qq] * (1.0f - sg[qq]);
        }
        float4* aux = reinterpret_cast<float4*>(tr + cnrec::AUX) + 3 * r;
        aux[0] = make_float4(d_o[0], d_o[1], d_o[2], d_o[3]);
        aux[1] = make_float4(dir[0], dir[1], dir[2], dir[3]);
        aux[2] = make_float4(ev[0], ev[1], ev[2], ev[3]);
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) d4s[qq * R + r] = d_o[qq];
      }
      __syncthreads();
      {  // per-warp sums over its 32 agents of d_o[q] * a3[c] for this thread's 16 channels (D3 is still in TMEM)
        const int c0 = cgrp * 16;
        float v[16];
        tmem_ld16(tlane + C_D3 + c0, v);
        wait_ld();
        const float dq[4] = {d4s[r], d4s[R + r], d4s[2 * R + r], d4s[3 * R + r]};
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float a = fmaxf(v[e] + B3[c0 + e], 0.f);
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            const float sw = warp_sum(dq[qq] * a);
            if (lane == 0) part[warp * 64 + qq * 16 + e] = sw;
          }
        }
      }
      fence_before_sync();
      __syncthreads();
      if (threadIdx.x < 256) {
        const int qq = threadIdx.x >> 6, c = threadIdx.x & 63, cg = c >> 4, e = c & 15;
        acc4 += (part[(cg * 4 + 0) * 64 + qq * 16 + e] + part[(cg * 4 + 1) * 64 + qq * 16 + e]) +
                (part[(cg * 4 + 2) * 64 + qq * 16 + e] + part[(cg * 4 + 3) * 64 + qq * 16 + e]);
      } else if (threadIdx.x < 260) {
        float sb = 0.f;
        for (int rr = 0; rr < R; ++rr) sb += d4s[(threadIdx.x - 256) * R + rr];
        acc4 += sb;
      }
    } else if (cgrp == 0 && valid) {
      float kk[4];
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) {
        const float o = B4[qq] + ((part[qq * R + r] + part[(4 + qq) * R + r]) + (part[(8 + qq) * R + r] + part[(12 + qq) * R + r]));
        kk[qq] = 2.0f * (1.0f / (1.0f + expf(-o))) + 0.2f;          // controller.py:53
      }
      const float4 s = hs;
      const float2 g = hg;
      const float ex = xsub(s.x, g.x), ey = xsub(s.y, g.y);
      const float ax = xadd(xmul(ex, -kk[0]), xmul(s.z, -kk[1]));   // controller.py:54-61
      const float ay = xadd(xmul(ey, -kk[2]), xmul(s.w, -kk[3]));
      reinterpret_cast<float2*>(u)[a0 + r] = make_float2(ax, ay);
      if (s_next != nullptr) {
        float4 n;
        n.x = xadd(s.x, xmul(s.z, dt)); n.y = xadd(s.y, xmul(s.w, dt));
        n.z = xadd(s.z, xmul(ax, dt)); n.w = xadd(s.w, xmul(ay, dt));
        reinterpret_cast<float4*>(s_next)[a0 + r] = n;
      }
    }
    __syncthreads();   // `part` is rewritten by the next tile
  }
  if constexpr (DUMP) {
    float* gp = slices + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;
    if (threadIdx.x < 256) gp[CT_NW4 + threadIdx.x] += acc4;          // dW4 [4][64] is contiguous
    else if (threadIdx.x < 260) gp[CT_NB4 + threadIdx.x - 256] += acc4;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) MACBF_TRACE_MARK(0, 10);
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int ctrl_node_tc(const float* states, const float* goals, const float* pool, const float* params, long long agents,
                 double dt, float* u, float* s_next, cudaStream_t st) {
  long long a_end = agents;
  int a_begin = 0;
  if (rows_active()) {
    a_begin = tl_row_range.begin;
    a_end = tl_row_range.end < agents ? tl_row_range.end : agents;
    if (a_end <= a_begin) return 0;
  }
  const long long ntiles = (a_end - a_begin + cntc::R - 1) / cntc::R;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_node_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cntc::smem_bytes));
  ctrl_node_tc_kernel<false><<<grid, cntc::THREADS, cntc::smem_bytes, st>>>(states, goals, pool, params, a_end, a_begin, (float)dt, u,
                                                                           s_next, tl_skip_flag, nullptr, nullptr, nullptr,
                                                                           nullptr);
  MACBF_CHECK_LAUNCH("ctrl_node_tc_kernel");
  return 0;
}

// first kernel of the tensor-core node backward: records for tiles 0 .. ceil(agents/128) and dW4/db4 slices
int ctrl_node_tc_dump(const float* states, const float* goals, const float* pool, const float* params, long long agents,
                      double dt, const float* grad_u, const float* grad_s_next, unsigned char* rec, float* slices,
                      cudaStream_t st) {
  const long long ntiles = (agents + cntc::R - 1) / cntc::R;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_node_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cntc::smem_bytes));
  ctrl_node_tc_kernel<true><<<grid, cntc::THREADS, cntc::smem_bytes, st>>>(states, goals, pool, params, agents, 0, (float)dt,
                                                                          nullptr, nullptr, nullptr, grad_u, grad_s_next, rec,
                                                                          slices);
  MACBF_CHECK_LAUNCH("ctrl_node_tc_kernel<dump>");
  return 0;
}

}  // namespace macbf
