// K2 (tensor-core path) - fused CBF head (cbf.py:21-45, REPAIRED) on tcgen05 / TMEM.
//
// One persistent CTA per SM: 16 worker warps + 1 MMA-issue warp, decoupled by mbarriers so the
// tensor core is never idle in steady state.  A tile is 128 edge rows (floor(128/K) agents);
// row r is TMEM lane r.  Worker warp w owns lanes 32*(w%4).. (the lanes tcgen05.ld/st lets it
// touch) and column group w/4 of whatever is being produced.  ALL dense layers of the MLP
// except the final 64->1 dot run on the tensor core; the CUDA cores only build the 6 edge
// features and run the epilogues:
//
//   workers(4 warps)  edge features (cbf.py:23-37) -> tf32 hi/lo -> shared memory A1 [128 x 8]
//   MMA warp          layer 1: D1[128x64]  = A1 * W1^T      (A and B from smem, K = 8)
//   workers           tcgen05.ld, bias + ReLU, hi/lo; hi overwrites D1 in place   -> A2   (TMEM)
//   MMA warp          layer 2: D2[128x128] = A2 * W2^T in two 64-column chunks    -> D2   (TMEM)
//   workers           per chunk: bias + ReLU, hi/lo; hi overwrites D2 in place    -> A3   (TMEM)
//   MMA warp          layer 3: cross terms per 64-wide K chunk, then hi*hi        -> D3   (TMEM)
//   workers           bias + ReLU, 64 -> 1 dot, * mask                            -> h    (global)
//
// Issue order of the MMA warp: L2(i), L1(i+1), L3(i), L2(i+1), ... so the workers turn D1(i+1)
// into A2(i+1) while the tensor core runs L3(i), and L2(i+1) starts the moment L3(i) retires.
// Every GEMM is a 3xTF32 split (hi*hi + lo*hi + hi*lo) with fp32 accumulation in TMEM; the small
// cross terms always go first because the tensor core's accumulate is not round-to-nearest
// (tools/tc_probe.cu).
// Hidden activations never leave the SM.  Shared memory: hi/lo halves of W1, W2, W3 in the
// canonical SWIZZLE_128B K-major layout (144 KB), the A1 staging tiles (32 KB), small vectors.
//
// TMEM columns: [0,64) D1 -> A2_hi (in place)  [64,128) A2_lo  [128,256) D2 -> A3_hi (in place)
//               [256,384) A3_lo  [384,448) D3
#include "common.cuh"
#include "tc_common.cuh"
#include "cbf_bwd_layout.cuh"

namespace macbf {

// Optional phase timeline (build with -DMACBF_TRACE; tools/tc_trace.py reads it back): CTA 0 logs
// (event, SM clock) pairs for worker warp 0 and for the MMA warp.
#ifdef MACBF_TRACE
__device__ unsigned long long g_tc_trace[2][4096];
#define TC_TRACE(who, ev)                                                                          \
  do {                                                                                             \
    if (blockIdx.x == 0 && lane == 0 && trace_n < 4096)                                            \
      g_tc_trace[who][trace_n++] = ((unsigned long long)(ev) << 56) | (clock64() & 0x00ffffffffffffffull); \
  } while (0)
extern "C" int macbf_debug_tc_trace(unsigned long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_tc_trace, sizeof(g_tc_trace));
}
#else
#define TC_TRACE(who, ev) do { } while (0)
#endif

namespace cbftc {
constexpr int WORKERS = 512;
constexpr int THREADS = WORKERS + 32;
constexpr int R = 128;
constexpr uint32_t C_A2HI = 0, C_A2LO = 64, C_Y = 128, C_A3LO = 256, C_D3 = 384;
// dynamic shared memory (floats, after 1024-B alignment); every MMA operand slab is 1024-B aligned
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW3HI = oW2LO + 128 * 64;
constexpr int oW3LO = oW3HI + 64 * 128;
constexpr int oW1HI = oW3LO + 64 * 128;    // [64 rows][32]: k = 0..5 used (one 128-B swizzle row per channel)
constexpr int oW1LO = oW1HI + 64 * 32;
constexpr int oA1HI = oW1LO + 64 * 32;     // [128 rows][32]: the 6 edge features of the next tile
constexpr int oA1LO = oA1HI + 128 * 32;
constexpr int oB1 = oA1LO + 128 * 32;
constexpr int oB2 = oB1 + 64;
constexpr int oB3 = oB2 + 128;
constexpr int oW4 = oB3 + 64;
constexpr int oB4 = oW4 + 64;
constexpr int oHpart = oB4 + 4;            // [2][4][128]
constexpr int oGms = oHpart + 2 * 4 * R;    // [4][128] ring: dL/dh * mask of the tiles in flight (DUMP mode)
constexpr int total_floats = oGms + 4 * R;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cbftc

// DUMP = true is the first kernel of the tensor-core backward (cbf_bwd_tc.cu): the same forward, over
// tiles [tile_begin, tile_end), that also writes per tile (record `tile - tile_begin`, cbf_bwd_layout.cuh)
// the bf16 hi/lo images of the layer-1/2 activations, the ReLU bit masks of all three layers, the edge
// features and gm = dL/dh * radius mask, and accumulates dW4 / db4 into this CTA's gradient slice.
// DUMP with grad_h == NULL is the RECORDING FORWARD of a training step (macbf_cbf_forward_record): dL/dh does not
// exist yet, so the record's GM slot receives the radius mask alone, h is written as by the plain forward, the
// layer-3 activations go to `a3buf` ([tile][column group][4][row][4] fp32; dW4 needs their values, not only their
// signs) and no gradient slice is touched - cbf_dgrad_tc_kernel finishes gm, dW4 and db4 when dL/dh arrives.
template <int MODE>       // 0: plain forward, 1: DUMP (first kernel of the backward), 2: DUMP as the recording forward
__global__ void __launch_bounds__(cbftc::THREADS, 1)
cbf_forward_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                      const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int K,
                      long long agents_total, int agent_begin, int TA, float dist_min, float obs_radius, float* __restrict__ h,
                      long long tile_begin, long long tile_end, const float* __restrict__ grad_h,
                      unsigned char* __restrict__ rec, float* __restrict__ slices, float* __restrict__ a3buf) {
  using namespace cbftc;
  using namespace tc;
  constexpr bool DUMP = MODE != 0, RECORD = MODE == 2;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO;
  float* W3hi = sm + oW3HI; float* W3lo = sm + oW3LO;
  float* W1hi = sm + oW1HI; float* W1lo = sm + oW1LO;
  float* A1hi = sm + oA1HI; float* A1lo = sm + oA1LO;
  float* B1 = sm + oB1; float* B2 = sm + oB2; float* B3 = sm + oB3; float* W4 = sm + oW4; float* B4 = sm + oB4;
  float* hpart = sm + oHpart;
  [[maybe_unused]] float* gms = sm + oGms;
  // workers -> MMA: a1_full (4 arrivals: the feature warps), a2_full, a3_full[2], d3_free (16: one per worker warp)
  // MMA -> workers (tcgen05.commit): d1_full, d2_full[2], d3_full
  __shared__ uint64_t a1_full, a2_full, a3_full[2], d3_free, d1_full, d2_full[2], d3_full;
  // DUMP: the feature warps publish gm of tile ordinal j on gm_full[j & 3] (4 arrivals); every worker acquires it in
  // its layer-3 epilogue.  The producers run 2 ordinals ahead of the consumers, so with one barrier per slot of the
  // 4-deep gm ring a barrier is never more than one phase ahead of a waiter (a parity wait would alias otherwise).
  __shared__ uint64_t gm_full[4];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef MACBF_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0) {       // entry: SM clock and wall clock (ns), to derive the clock rate
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    g_tc_trace[1][4090] = clock64(); g_tc_trace[1][4091] = gt;
  }
#endif

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&a1_full, 4); mbar_init(&a2_full, 16); mbar_init(&a3_full[0], 16); mbar_init(&a3_full[1], 16);
    mbar_init(&d3_free, 16);
    mbar_init(&d1_full, 1); mbar_init(&d2_full[0], 1); mbar_init(&d2_full[1], 1); mbar_init(&d3_full, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&gm_full[i], 4);
    fence_mbar_init();
  }
  {   // both weight matrices: all loads in flight, then the splits and stores
    WeightFill<128, 64, THREADS> f2;
    WeightFill<64, 128, THREADS> f3;
    f2.template load<false>(params + CBF_W2, 64, 1);
    f3.template load<false>(params + CBF_W3, 128, 1);
    f2.template store<false>(W2hi, W2lo);
    f3.template store<false>(W3hi, W3lo);
  }
  // W1 (64 x 6): channel n is one 128-B row; 16-B chunk c sits at position c ^ (n & 7); A1 likewise
  {
    float w1 = 0.f;                                   // the 384 real entries: one load each, in flight with the rest
    if (threadIdx.x < 64 * 6) w1 = __ldg(params + CBF_W1 + threadIdx.x);
    for (int t = threadIdx.x; t < 64 * 32; t += THREADS) {
      const int n = t >> 5, pos = (t >> 2) & 7, e = t & 3;
      const int k = ((pos ^ (n & 7)) << 2) + e;
      if (k >= 6) { W1hi[t] = 0.f; W1lo[t] = 0.f; }
    }
    if (threadIdx.x < 64 * 6) {
      const int n = threadIdx.x / 6, k = threadIdx.x - n * 6;
      float hi, lo;
      tf32_split(w1, hi, lo);
      const int t = n * 32 + ((((k >> 2) ^ (n & 7)) << 2) | (k & 3));
      W1hi[t] = hi; W1lo[t] = lo;
    }
  }
  for (int t = threadIdx.x; t < 128 * 32; t += THREADS) { A1hi[t] = 0.f; A1lo[t] = 0.f; }
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CBF_B2 + t);
  for (int t = threadIdx.x; t < 64; t += THREADS) {
    B1[t] = __ldg(params + CBF_B1 + t);
    B3[t] = __ldg(params + CBF_B3 + t);
    W4[t] = __ldg(params + CBF_W4 + t);
  }
  if (threadIdx.x == 0) B4[0] = __ldg(params + CBF_B4);
  fence_proxy_async_smem();   // weights and zeroed A1 tiles: generic proxy -> tensor-core proxy
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long ntiles = tile_end;
  const long long first = tile_begin + blockIdx.x;
  [[maybe_unused]] int trace_n = 0;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    // Everything here is warp-uniform; only the tcgen05 instructions sit under elect_one().
    const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1hi)), d_w1lo = smem_desc_sw128(smem_u32(W1lo));
    const uint64_t d_a1hi = smem_desc_sw128(smem_u32(A1hi)), d_a1lo = smem_desc_sw128(smem_u32(A1lo));
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
    const uint64_t d_w3hi = smem_desc_sw128(smem_u32(W3hi)), d_w3lo = smem_desc_sw128(smem_u32(W3lo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC = idesc_tf32(128, 64);
    // A whole batch of MMAs sits under ONE elect_one(), and stepping a weight descriptor along K or to another row block
    // only changes its low word (the 14-bit address field): one 32-bit add per MMA.  With an elect and 64-bit
    // descriptor arithmetic per MMA the issuing warp needed about as long per 128x64x8 MMA (35 cycles on the tensor
    // core) as the tensor core itself, and starved it.
    const uint32_t w2hi_lo = (uint32_t)d_w2hi, w2lo_lo = (uint32_t)d_w2lo, w3hi_lo = (uint32_t)d_w3hi, w3lo_lo = (uint32_t)d_w3lo;
    const uint32_t w_hi32 = (uint32_t)(d_w2hi >> 32);      // identical for every SWIZZLE_128B K-major operand
    auto issue_layer1 = [&](uint32_t parity) {       // D1 = A1 * W1^T, one k-step of 8 per term
      mbar_wait(&a1_full, parity);
      fence_after_sync();
      if (elect_one()) {
        mma_tf32_ss(tb + C_A2HI, d_a1lo, d_w1hi, IDESC, 0u);
        mma_tf32_ss(tb + C_A2HI, d_a1hi, d_w1lo, IDESC, 1u);
        mma_tf32_ss(tb + C_A2HI, d_a1hi, d_w1hi, IDESC, 1u);
        mma_commit(&d1_full);
      }
      __syncwarp();
    };
    issue_layer1(0);
    uint32_t par = 0;
    int it = 0;
    for (long long tile = first; tile < ntiles; tile += gridDim.x, par ^= 1, ++it) {
      mbar_wait(&a2_full, par);
      fence_after_sync();
      TC_TRACE(1, 1);
      // layer 2, output columns [64c, 64c+64): B rows 64c.. of every 32-wide K block
      if (elect_one()) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
#pragma unroll
          for (int term = 0; term < 3; ++term) {         // lo*hi, hi*lo, hi*hi (small terms first)
            const uint32_t a_col = tb + (term == 0 ? C_A2LO : C_A2HI);
            const uint32_t b_lo = ((term == 1) ? w2lo_lo : w2hi_lo) + ((c * (64 * 128)) >> 4);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              mma_tf32_ts_split(tb + C_Y + c * 64, a_col + ks * 8, b_lo + (((ks >> 2) * (128 * 128) + (ks & 3) * 32) >> 4),
                                w_hi32, IDESC, (term | ks) ? 1u : 0u);
          }
          mma_commit(&d2_full[c]);
        }
      }
      __syncwarp();
      TC_TRACE(1, 2);
      // layer 1 of the next tile: its D1 lands in the A2 columns this tile's layer 2 has just finished reading
      if (tile + gridDim.x < ntiles) issue_layer1(par ^ 1);
      // layer 3 accumulates into ONE tile D3: the cross terms of both 64-wide K chunks first (each as
      // soon as its A3 chunk exists), the two hi*hi passes last - small terms still meet a small sum,
      // and the epilogue reads 64 TMEM columns instead of 128.
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&a3_full[c], par);
        if (c == 0 && it > 0) mbar_wait(&d3_free, par ^ 1);   // previous tile's D3 has been read
        fence_after_sync();
        TC_TRACE(1, 3 + c);
        if (elect_one()) {
#pragma unroll
          for (int term = 0; term < 2; ++term) {         // lo*hi, hi*lo of K chunk [64c, 64c+64)
            const uint32_t a_col = tb + (term == 0 ? C_A3LO : C_Y) + c * 64;
            const uint32_t b_lo = ((term == 1) ? w3lo_lo : w3hi_lo) + (((2 * c) * (64 * 128)) >> 4);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              mma_tf32_ts_split(tb + C_D3, a_col + ks * 8, b_lo + (((ks >> 2) * (64 * 128) + (ks & 3) * 32) >> 4), w_hi32, IDESC,
                                (c | term | ks) ? 1u : 0u);
          }
        }
        __syncwarp();
      }
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 16; ++ks)                  // hi*hi over the whole K = 128
          mma_tf32_ts_split(tb + C_D3, tb + C_Y + ks * 8, w3hi_lo + (((ks >> 2) * (64 * 128) + (ks & 3) * 32) >> 4), w_hi32, IDESC, 1u);
        mma_commit(&d3_full);
      }
      __syncwarp();
      TC_TRACE(1, 5);
    }
  } else {
    // ================================= worker warps =================================
    const int q = warp & 3, cgrp = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int a_loc = r / K;                       // agent of this row within a tile (tile-invariant)

    // Edge endpoints of upcoming tiles are fetched one phase ahead of their use (feature warps only).
    int pf_j = 0, pf_b = 0, pf_i = 0;
    [[maybe_unused]] long long pf_tile = 0;
    [[maybe_unused]] int pf_slot = 0;
    bool pf_valid = false;
    float4 pf_si = make_float4(0.f, 0.f, 0.f, 0.f), pf_sj = pf_si;
    auto prefetch_ij = [&](long long t) {
      pf_valid = false;
      pf_tile = t;
      if (cgrp == 0 && t < ntiles) {
        if (!DUMP && TA == 0) {
          // edge tiles (plain forward): 128 consecutive edges, whatever agents they belong to - h is per edge, so a
          // tile need not hold whole agents and no row is padding (128 / K agents per tile would waste 128 % K rows)
          const long long e = (long long)agent_begin * K + t * R + r;   // B*N*K < 2^31 is checked by the host
          if (e < agents_total * K) {
            const int ga = (int)e / K;
            pf_b = ga / N; pf_i = ga - pf_b * N;
            pf_j = __ldg(idx + e);
            pf_si = __ldg(reinterpret_cast<const float4*>(states) + ga);
            pf_valid = true;
          }
        } else {
          const int ga = (int)(t * TA) + a_loc;    // B*N < 2^31 is checked by the host
          if (a_loc < TA && ga < (int)agents_total) {
            pf_b = ga / N; pf_i = ga - pf_b * N;
            pf_j = __ldg(idx + (size_t)(t * TA) * K + r);
            pf_si = __ldg(reinterpret_cast<const float4*>(states) + ga);
            pf_valid = true;
          }
        }
      }
    };
    auto prefetch_sj = [&]() {
      if (pf_valid) pf_sj = load_candidate(states, obstacles, pf_b, pf_j, N, M);
    };
    // features (cbf.py:23-37) of this thread's row -> A1 (smem, swizzled row r); returns the radius mask
    auto features = [&]() -> float {
      float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      float mk = 0.f;
      if (pf_valid) {
        const EdgeGeom g = edge_geom(pf_si, pf_sj, pf_i, pf_j);
        const float d = cbf_dist(g.dx, g.dy);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye; f[5] = xsub(d, dist_min);
        mk = (d <= obs_radius) ? 1.f : 0.f;
      }
      if constexpr (DUMP) {
        unsigned char* tr = rec + (size_t)(pf_tile - tile_begin) * cbfrec::BYTES;
        float gm = 0.f, dd = 1.f;
        if (pf_valid) {
          gm = RECORD ? mk : __ldg(grad_h + (size_t)(pf_tile * TA) * K + r) * mk;      // d h / d y4 = mask (cbf.py:43)
          dd = xadd(f[5], dist_min);
        }
        float4* xr = reinterpret_cast<float4*>(tr + cbfrec::XREC) + 2 * r;
        xr[0] = make_float4(f[0], f[1], f[2], f[3]);
        xr[1] = make_float4(f[4], f[5], dd, 0.f);
        reinterpret_cast<float*>(tr + cbfrec::GM)[r] = gm;
        if constexpr (RECORD) reinterpret_cast<float*>(tr + cbfrec::GMASK)[r] = gm;    // recording forward: the mask alone
        if constexpr (!RECORD) {
          gms[(pf_slot & 3) * R + r] = gm;
          __syncwarp();
          if (lane == 0) mbar_arrive(&gm_full[pf_slot & 3]);
          ++pf_slot;
        }
      }
      float hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) tf32_split(f[e], hi[e], lo[e]);
      const int p0 = (0 ^ (r & 7)) << 2, p1 = (1 ^ (r & 7)) << 2;      // swizzled positions of chunks 0 and 1
      *reinterpret_cast<float4*>(A1hi + r * 32 + p0) = make_float4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<float4*>(A1hi + r * 32 + p1) = make_float4(hi[4], hi[5], hi[6], hi[7]);
      *reinterpret_cast<float4*>(A1lo + r * 32 + p0) = make_float4(lo[0], lo[1], lo[2], lo[3]);
      *reinterpret_cast<float4*>(A1lo + r * 32 + p1) = make_float4(lo[4], lo[5], lo[6], lo[7]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a1_full);
      return mk;
    };
    // layer-1 epilogue: D1 -> bias + ReLU + split -> A2 (hi in place over D1, lo aside)
    auto epilogue1 = [&](uint32_t parity, [[maybe_unused]] long long tile_e) {
      mbar_wait(&d1_full, parity);
      fence_after_sync();
      const int c0 = cgrp * 16;
      float v[16];
      tmem_ld16(tlane + C_A2HI + c0, v);
      wait_ld();
      [[maybe_unused]] uint32_t bits = 0;
      [[maybe_unused]] unsigned char* tr = DUMP ? rec + (size_t)(tile_e - tile_begin) * cbfrec::BYTES : nullptr;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float hi[8], lo[8];
        [[maybe_unused]] float x[8];
#pragma unroll
        for (int j4 = 0; j4 < 2; ++j4) {
          const float4 bb = *reinterpret_cast<const float4*>(B1 + c0 + half * 8 + 4 * j4);
          float2 h2, l2;
          const float2 xa = relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 0], v[half * 8 + 4 * j4 + 1]), make_float2(bb.x, bb.y)));
          split2_trunc(xa, h2, l2);
          hi[4 * j4 + 0] = h2.x; hi[4 * j4 + 1] = h2.y; lo[4 * j4 + 0] = l2.x; lo[4 * j4 + 1] = l2.y;
          const float2 xb = relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 2], v[half * 8 + 4 * j4 + 3]), make_float2(bb.z, bb.w)));
          split2_trunc(xb, h2, l2);
          hi[4 * j4 + 2] = h2.x; hi[4 * j4 + 3] = h2.y; lo[4 * j4 + 2] = l2.x; lo[4 * j4 + 3] = l2.y;
          if constexpr (DUMP) { x[4 * j4 + 0] = xa.x; x[4 * j4 + 1] = xa.y; x[4 * j4 + 2] = xb.x; x[4 * j4 + 3] = xb.y; }
        }
        tmem_st8(tlane + C_A2HI + c0 + half * 8, hi);
        tmem_st8(tlane + C_A2LO + c0 + half * 8, lo);
        if constexpr (DUMP) {
#pragma unroll
          for (int e = 0; e < 8; ++e) bits |= (x[e] > 0.f ? 1u : 0u) << (half * 8 + e);
        }
      }
      // (the layer-1 activations themselves are not dumped: the wgrad kernel rebuilds them from the edge features)
      if constexpr (DUMP) reinterpret_cast<uint32_t*>(tr + cbfrec::MASKS)[(0 * 4 + cgrp) * R + r] = bits;
      wait_st();
      warp_arrive_after_tmem(&a2_full);
    };

    // Software pipeline of the feature warps: the features of tile t+2 are built during tile t
    // (in the slack before its layer-3 epilogue), so layer 1 of the next tile never waits for them.
    const long long g = gridDim.x;
    float m0 = 0.f, m1 = 0.f, m2 = 0.f;              // radius masks of tiles t, t+g, t+2g (feature warps)
    prefetch_ij(first);
    prefetch_sj();
    if (cgrp == 0) m0 = features();
    epilogue1(0, first);
    prefetch_ij(first + g);
    prefetch_sj();
    if (cgrp == 0 && first + g < ntiles) m1 = features();
    prefetch_ij(first + 2 * g);
    [[maybe_unused]] float acc4[16];
    [[maybe_unused]] float accb4 = 0.f;
    [[maybe_unused]] int gslot = 0;
    if constexpr (DUMP && !RECORD) {
#pragma unroll
      for (int q = 0; q < 16; ++q) acc4[q] = 0.f;
    }

    uint32_t par = 0;
    for (long long tile = first; tile < ntiles; tile += g, par ^= 1) {
      [[maybe_unused]] unsigned char* tr = DUMP ? rec + (size_t)(tile - tile_begin) * cbfrec::BYTES : nullptr;
      const long long row0 = (!DUMP && TA == 0) ? (long long)agent_begin * K + tile * R : tile * TA * K;      // first edge of the tile
      const int rows = (!DUMP && TA == 0) ? (int)min((long long)R, agents_total * K - row0)
                                 : (int)min((long long)TA, agents_total - tile * TA) * K;
      const float mk = m0;
      const bool has_next = tile + g < ntiles;
      prefetch_sj();                                   // endpoints of tile + 2g (their idx arrived last iteration)
      // ---- layer-2 epilogue per 64-column chunk: bias + ReLU + split; hi in place, lo aside ----
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&d2_full[c], par);
        fence_after_sync();
        if (warp == 0) TC_TRACE(0, 1 + 2 * c);
        const int c0 = c * 64 + cgrp * 16;
        float v[16];
        tmem_ld16(tlane + C_Y + c0, v);
        wait_ld();
        [[maybe_unused]] uint32_t bits = 0;
        [[maybe_unused]] uint4 bh[2], bl[2];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float hi[8], lo[8];
          [[maybe_unused]] float x[8];
#pragma unroll
          for (int j4 = 0; j4 < 2; ++j4) {
            const float4 bb = *reinterpret_cast<const float4*>(B2 + c0 + half * 8 + 4 * j4);
            float2 h2, l2;
            const float2 xa = relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 0], v[half * 8 + 4 * j4 + 1]), make_float2(bb.x, bb.y)));
            split2_trunc(xa, h2, l2);
            hi[4 * j4 + 0] = h2.x; hi[4 * j4 + 1] = h2.y; lo[4 * j4 + 0] = l2.x; lo[4 * j4 + 1] = l2.y;
            const float2 xb = relu2(__fadd2_rn(make_float2(v[half * 8 + 4 * j4 + 2], v[half * 8 + 4 * j4 + 3]), make_float2(bb.z, bb.w)));
            split2_trunc(xb, h2, l2);
            hi[4 * j4 + 2] = h2.x; hi[4 * j4 + 3] = h2.y; lo[4 * j4 + 2] = l2.x; lo[4 * j4 + 3] = l2.y;
            if constexpr (DUMP) { x[4 * j4 + 0] = xa.x; x[4 * j4 + 1] = xa.y; x[4 * j4 + 2] = xb.x; x[4 * j4 + 3] = xb.y; }
          }
          tmem_st8(tlane + C_Y + c0 + half * 8, hi);
          tmem_st8(tlane + C_A3LO + c0 + half * 8, lo);
          if constexpr (DUMP) {
#pragma unroll
            for (int e = 0; e < 8; ++e) bits |= (x[e] > 0.f ? 1u : 0u) << (half * 8 + e);
            bf16_split8(x, bh[half], bl[half]);
          }
        }
        if constexpr (DUMP) {
          img_store16(tr + cbfrec::A2_HI + c * cbfrec::BLK, r, cgrp, bh[0], bh[1]);
          img_store16(tr + cbfrec::A2_LO + c * cbfrec::BLK, r, cgrp, bl[0], bl[1]);
          reinterpret_cast<uint32_t*>(tr + cbfrec::MASKS)[((1 + c) * 4 + cgrp) * R + r] = bits;
        }
        wait_st();
        warp_arrive_after_tmem(&a3_full[c]);
        if (warp == 0) TC_TRACE(0, 2 + 2 * c);
      }
      // ---- layer-1 epilogue of the NEXT tile (overlaps this tile's layer 3 on the tensor core) ----
      if (has_next) epilogue1(par ^ 1, tile + g);
      // ---- features of tile + 2g -> A1 (layer 1 of tile + g has retired: d1_full was just observed) ----
      if (cgrp == 0 && tile + 2 * g < ntiles) {
        m2 = features();
        prefetch_ij(tile + 3 * g);
      }
      if (warp == 0) TC_TRACE(0, 5);
      // ---- layer-3 epilogue: bias + ReLU, 64 -> 1 over this thread's 16 channels ----
      mbar_wait(&d3_full, par);
      fence_after_sync();
      if (warp == 0) TC_TRACE(0, 6);
      {
        const int c0 = cgrp * 16;
        float v[16];
        tmem_ld16(tlane + C_D3 + c0, v);
        wait_ld();
        warp_arrive_after_tmem(&d3_free);
        if constexpr (DUMP) {
          uint32_t bits = 0;
          if constexpr (RECORD) {                       // recording forward: the activations themselves, for dW4 later
            float4* dst = reinterpret_cast<float4*>(a3buf) + ((size_t)(tile - tile_begin) * 4 + cgrp) * 4 * R + r;   // [j4][row]: a warp's store is 512 B
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              float z[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                z[q] = fmaxf(v[4 * j4 + q] + B3[c0 + 4 * j4 + q], 0.f);
                bits |= (z[q] > 0.f ? 1u : 0u) << (4 * j4 + q);
              }
              dst[j4 * R] = make_float4(z[0], z[1], z[2], z[3]);
            }
          } else {
            mbar_wait(&gm_full[gslot & 3], (gslot >> 2) & 1);   // published >= 1 tile ago: an acquire, not a wait
            const float gm = gms[(gslot & 3) * R + r];
            ++gslot;
#pragma unroll
            for (int q = 0; q < 16; ++q) {
              const float z = fmaxf(v[q] + B3[c0 + q], 0.f);
              bits |= (z > 0.f ? 1u : 0u) << q;
              acc4[q] = fmaf(gm, z, acc4[q]);              // dW4 (cbf.py:19)
            }
            if (cgrp == 0) accb4 += gm;
          }
          reinterpret_cast<uint32_t*>(tr + cbfrec::MASKS)[(3 * 4 + cgrp) * R + r] = bits;
        }
        float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 bb = *reinterpret_cast<const float4*>(B3 + c0 + 4 * j4);
          const float4 ww = *reinterpret_cast<const float4*>(W4 + c0 + 4 * j4);
          s2 = __ffma2_rn(make_float2(ww.x, ww.y), relu2(__fadd2_rn(make_float2(v[4 * j4 + 0], v[4 * j4 + 1]), make_float2(bb.x, bb.y))), s2);
          s2 = __ffma2_rn(make_float2(ww.z, ww.w), relu2(__fadd2_rn(make_float2(v[4 * j4 + 2], v[4 * j4 + 3]), make_float2(bb.z, bb.w))), s2);
        }
        if (h != nullptr) {
          const float s = s2.x + s2.y;
          float* hp = hpart + (par * 4) * R;
          hp[cgrp * R + r] = s;
          named_bar_sync(1, WORKERS);
          if (cgrp == 0 && r < rows) {
            const float t = (hp[r] + hp[R + r]) + (hp[2 * R + r] + hp[3 * R + r]);
            h[row0 + r] = (t + B4[0]) * mk;            // cbf.py:43
          }
        }
      }
      if (warp == 0) TC_TRACE(0, 7);
      m0 = m1; m1 = m2;
    }
    if constexpr (DUMP && !RECORD) {
      // dW4 / db4 of this CTA: lanes, then the 4 lane-quadrant warps of each column group
      named_bar_sync(1, WORKERS);
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        const float s = warp_sum(acc4[q]);
        if (lane == 0) hpart[warp * 16 + q] = s;
      }
      const float sb = warp_sum(accb4);
      if (lane == 0) hpart[256 + warp] = sb;
      named_bar_sync(1, WORKERS);
      float* gp = slices + (size_t)blockIdx.x * MACBF_CBF_PARAMS;
      if (threadIdx.x < 64) {
        const int cg = threadIdx.x >> 4, q = threadIdx.x & 15;
        gp[CBF_W4 + threadIdx.x] += (hpart[(cg * 4 + 0) * 16 + q] + hpart[(cg * 4 + 1) * 16 + q]) +
                                    (hpart[(cg * 4 + 2) * 16 + q] + hpart[(cg * 4 + 3) * 16 + q]);
      } else if (threadIdx.x == 64) {
        gp[CBF_B4] += (hpart[256 + 0] + hpart[256 + 1]) + (hpart[256 + 2] + hpart[256 + 3]);
      }
    }
  }
  fence_before_sync();
  __syncthreads();
#ifdef MACBF_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    g_tc_trace[1][4092] = clock64(); g_tc_trace[1][4093] = gt;
  }
#endif
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int cbf_forward_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                   int M, int K, double dist_min, double obs_radius, float* h, cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "cbf_forward: B*N must be < 2^31");
  // edge tiles (TA = 0) whenever the edge count fits 32-bit arithmetic, else tiles of whole agents
  const bool edge_tiles = agents * K < (1ll << 31);
  long long a_end = agents;
  int a_begin = 0;
  if (rows_active()) {                             // a row range of one env (macbf_step_forward_rows_*)
    MACBF_CHECK_ARG(B == 1 && edge_tiles, "cbf_forward: a row range needs B = 1 and N*K < 2^31");
    a_begin = tl_row_range.begin;
    a_end = tl_row_range.end < agents ? tl_row_range.end : agents;
    if (a_end <= a_begin) return 0;
  }
  const int TA = edge_tiles ? 0 : max(1, cbftc::R / K);
  const long long ntiles = edge_tiles ? ((a_end - a_begin) * K + cbftc::R - 1) / cbftc::R : (agents + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_forward_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cbftc::smem_bytes));
  cbf_forward_tc_kernel<0><<<grid, cbftc::THREADS, cbftc::smem_bytes, st>>>(
      states, obstacles, idx, params, N, M, K, a_end, a_begin, TA, (float)dist_min, (float)obs_radius, h, 0, ntiles, nullptr,
      nullptr, nullptr, nullptr);
  MACBF_CHECK_LAUNCH("cbf_forward_tc_kernel");
  return 0;
}

// First kernel of the tensor-core backward: tiles [tile_begin, tile_end) -> records + dW4/db4 slices.  With grad_h ==
// NULL it is the recording forward: h and a3buf are written, slices is not.
int cbf_forward_tc_dump(const float* states, const float* obstacles, const int32_t* idx, const float* params, int N, int M,
                        int K, long long agents, double dist_min, double obs_radius, long long tile_begin,
                        long long tile_end, const float* grad_h, unsigned char* rec, float* slices, cudaStream_t st,
                        float* h, float* a3buf) {
  MACBF_CHECK_ARG((grad_h != nullptr) != (a3buf != nullptr) && (grad_h != nullptr) == (slices != nullptr) &&
                      (grad_h != nullptr) == (h == nullptr),
                  "cbf_forward_tc_dump: either (grad_h, slices) or (h, a3buf)");
  const int TA = max(1, cbftc::R / K);
  const long long nt = tile_end - tile_begin;
  const int grid = (int)(nt < NUM_SMS ? nt : (long long)NUM_SMS);
  if (grad_h == nullptr) {
    MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_forward_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)cbftc::smem_bytes));
    cbf_forward_tc_kernel<2><<<grid, cbftc::THREADS, cbftc::smem_bytes, st>>>(
        states, obstacles, idx, params, N, M, K, agents, 0, TA, (float)dist_min, (float)obs_radius, h, tile_begin, tile_end,
        nullptr, rec, nullptr, a3buf);
    MACBF_CHECK_LAUNCH("cbf_forward_tc_kernel<record>");
    return 0;
  }
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_forward_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cbftc::smem_bytes));
  cbf_forward_tc_kernel<1><<<grid, cbftc::THREADS, cbftc::smem_bytes, st>>>(
      states, obstacles, idx, params, N, M, K, agents, 0, TA, (float)dist_min, (float)obs_radius, nullptr, tile_begin, tile_end,
      grad_h, rec, slices, nullptr);
  MACBF_CHECK_LAUNCH("cbf_forward_tc_kernel<dump>");
  return 0;
}

}  // namespace macbf
