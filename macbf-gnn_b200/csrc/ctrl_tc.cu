// K3a - controller edge net + masked max pooling (controller.py:33-46, REPAIRED) on tcgen05 / TMEM, with the
// second layer TRANSPOSED so that the pooling needs no shared-memory stage.
//
// A tile is 128 edge rows = floor(128/K) agents.  Per tile:
//
//   feature warps   edge features (controller.py:33-40) -> tf32 hi/lo -> X [128 edges x 8] in shared memory
//                   (SWIZZLE_32B rows of 32 bytes).  An edge outside the sensing radius (controller.py:38-39) is
//                   REPLACED by the agent's own self edge (0,0,0,0,1): its column then repeats slot 0, which is never
//                   masked, so the max over the K slots is untouched - exactly what multiplying by the 0/1 mask
//                   before the max does for non-negative (ReLU) values - and the mask costs nothing downstream.
//   MMA warp        layer 1:  D1[128 edges x 64]  = X * W1^T                 (A, B from shared memory, K = 8)
//   workers         bias + ReLU + hi/lo split; one thread = one edge row writes its 64 channels as one K-major
//                   row of the layer-2 B operand A2 [128 edges x 64] (SWIZZLE_128B, conflict-free 16-byte stores)
//   MMA warp        layer 2:  D2^T[128 channels x 128 edges] = W2 * A2^T     (A = W2 resident in tensor memory)
//   workers         channel c = TMEM lane, the K slots of an agent = K consecutive TMEM columns:
//                   max over the slots on registers, + bias, ReLU (monotone, so max first is exact), one coalesced
//                   128-byte store of pool[agent][32 channels] per warp                     (controller.py:46)
//
// Both dense layers are 3xTF32 (cross terms first).  Warp roles of the persistent CTA (one per SM):
//   8 "row" warps    thread = (edge row, 32 of its 64 channels): layer-1 epilogue of tile j
//   4 "pool" warps   thread = channel: pooling of tile j
//   features of tile j+2 (thread = edge row): the pool warps for even j, row warps 4-7 for odd j, so that no warp group
//   carries more per tile than the tensor core needs for its MMAs
//   2 MMA warps      even / odd tiles (one warp needs about as long to issue a tile's 27 MMAs as the tensor core to run them)
// The roles only meet through mbarriers, so the chain  layer 1 -> epilogue -> layer 2 -> pooling  of consecutive tiles
// overlaps freely (two D1, D2, X and A2 buffers); with every warp doing every stage in turn the latency of one
// stage delayed the next tile's MMAs and the tensor core idled 40 % of the time.
// W2 (the A operand of layer 2, hi and lo) lives in TENSOR MEMORY for the whole kernel: with both operands in shared
// memory every 64-cycle MMA would read 8 KB from it - the whole shared-memory bandwidth of the SM - and starve the
// workers' stores; with A in TMEM only the 4 KB of activations are read.
// TMEM columns: [0,128) D1 x 2, [128,384) D2 x 2, [384,448) W2 hi, [448,512) W2 lo.
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

// Optional phase timeline (build with -DMACBF_TRACE; tools/ce_trace.py reads it back): CTA 0 logs clock64 at the phase
// boundaries of one row warp (0), the two MMA warps (1, 3) and one pool warp (2).
#ifdef MACBF_TRACE
__device__ unsigned long long g_ce2_trace[4][4096];
#define CE2_TRACE(who, ev)                                                                          \
  do {                                                                                              \
    if (blockIdx.x == 0 && lane == 0 && trace_n < 4096)                                             \
      g_ce2_trace[who][trace_n++] = ((unsigned long long)(ev) << 56) | (clock64() & 0x00ffffffffffffffull); \
  } while (0)
extern "C" int macbf_debug_ce2_trace(unsigned long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_ce2_trace, sizeof(g_ce2_trace));
}
#else
#define CE2_TRACE(who, ev) do { } while (0)
#endif

namespace cet2 {
constexpr int ROW_WARPS = 8, POOL_WARPS = 4;
constexpr int WORKERS = 32 * (ROW_WARPS + POOL_WARPS);
constexpr int THREADS = WORKERS + 64;        // + two MMA-issue warps (even / odd tiles)
constexpr int R = 128;
constexpr uint32_t C_D1 = 0, C_D2 = 128, C_W2HI = 384, C_W2LO = 448;
constexpr int oA2 = 0;                       // [buffer][hi, lo][2 slabs][128 edges][32]  K-major SWIZZLE_128B
constexpr int oW1 = oA2 + 2 * 2 * 128 * 64;  // [hi, lo][64 rows][8]   SWIZZLE_32B
constexpr int oX = oW1 + 2 * 64 * 8;         // [buffer][hi, lo][128 rows][8] SWIZZLE_32B
constexpr int oB1 = oX + 2 * 2 * 128 * 8;
constexpr int oB2 = oB1 + 64;
constexpr int oMask = oB2 + 128;             // [8][128] sensing-radius masks of the tiles in flight (WIN variant only)
constexpr int total_floats = oMask + 8 * 128;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cet2

// WIN = true (training forward / backward pass): also writes the first arg-max slot of every (agent, channel) to `argk`
// (255 = the maximum is 0: no gradient).  KT = 12: the reference's TOP_K fully unrolled; KT = 0: any K.
template <bool WIN, int KT>
__global__ void __launch_bounds__(cet2::THREADS, 1)
ctrl_edge_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                     const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int M, int Krt,
                     long long agents_total, int agent_begin, int TA, float obs_radius, float* __restrict__ pool,
                     const int* __restrict__ skip, uint8_t* __restrict__ argk) {
  MACBF_SKIP_IF_SET(skip);
  using namespace cet2;
  using namespace tc;
  const int K = KT ? KT : Krt;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W1hi = sm + oW1; float* W1lo = W1hi + 64 * 8;
  float* B1 = sm + oB1; float* B2 = sm + oB2;
  // row warps -> MMA: a2_full (8 arrivals), x_full (4); pool warps -> MMA: d2_free (4); MMA -> workers: d1_full, d2_full
  __shared__ uint64_t x_full[2], a2_full[2], d1_full[2], d2_full[2], d2_free[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  [[maybe_unused]] int trace_n = 0;
  if (warp == 0) CE2_TRACE(0, 10);

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(&x_full[b], 4); mbar_init(&a2_full[b], ROW_WARPS); mbar_init(&d2_free[b], POOL_WARPS);
      mbar_init(&d1_full[b], 1); mbar_init(&d2_full[b], 1);
    }
    fence_mbar_init();
  }
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) {     // W1 (64 x 5, zero-padded to k = 8), SWIZZLE_32B
    const int n = t >> 3, pos = (t >> 2) & 1, e = t & 3;
    const int k = ((pos ^ ((n >> 2) & 1)) << 2) + e;
    float hi = 0.f, lo = 0.f;
    if (k < 5) tf32_split(__ldg(params + CT_EW1 + n * 5 + k), hi, lo);
    W1hi[t] = hi; W1lo[t] = lo;
  }
  for (int t = threadIdx.x; t < 64; t += THREADS) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  if (warp < ROW_WARPS) {
    // W2 (128 x 64, torch (out, in) layout) -> TMEM: lane = output channel, column = k; this warp's 32 lanes x 32 columns
    const int q = warp & 3, hh = warp >> 2, ch = q * 32 + lane;
    const uint32_t tl = tbase + ((uint32_t)(q * 32) << 16);
#pragma unroll
    for (int c0 = 0; c0 < 32; c0 += 16) {
      float hi[16], lo[16];
#pragma unroll
      for (int e = 0; e < 16; e += 4) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(params + CT_EW2 + ch * 64 + hh * 32 + c0 + e));
        tf32_split(w.x, hi[e + 0], lo[e + 0]); tf32_split(w.y, hi[e + 1], lo[e + 1]);
        tf32_split(w.z, hi[e + 2], lo[e + 2]); tf32_split(w.w, hi[e + 3], lo[e + 3]);
      }
      tmem_st16(tl + C_W2HI + hh * 32 + c0, hi);
      tmem_st16(tl + C_W2LO + hh * 32 + c0, lo);
    }
    wait_st();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  // tiles cover the agents [agent_begin, agents_total) (agent_begin > 0: a row range of one env, B = 1)
  const long long ntiles = (agents_total - agent_begin + TA - 1) / TA;
  const long long gdim = gridDim.x;
  const int nj = (ntiles > blockIdx.x) ? (int)((ntiles - blockIdx.x + gdim - 1) / gdim) : 0;   // tiles of this CTA
  if (warp == 0) CE2_TRACE(0, 11);

  // ---- features (thread = edge row r of a tile).  Two warp groups share the work: the pool warps build the X tiles of
  //      the CTA's even tiles, row warps 4-7 those of the odd tiles.  idx / s_i of a tile are requested two of the group's
  //      tiles ahead (`nxt`), its far endpoints s_j one ahead (`cur`, `cur_sj`), always right AFTER the group has
  //      published a tile: fence.proxy.async (a MEMBAR) also waits for the thread's outstanding global loads, so a
  //      load issued just before an epilogue's fence would put its DRAM latency on the path of the MMA hand-off ----
  const int r = (warp & 3) * 32 + lane;
  struct Pf { int j, b, i; bool valid; float4 si; };
  const int a_loc = r / K;
  auto load_ij = [&](int j) {
    Pf p;
    p.j = 0; p.b = 0; p.i = 0; p.valid = false; p.si = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < nj) {
      const long long t = blockIdx.x + (long long)j * gdim;
      const int ga = agent_begin + (int)(t * TA) + a_loc;        // B*N < 2^31 is checked by the host
      if (a_loc < TA && ga < (int)agents_total) {
        p.b = ga / N; p.i = ga - p.b * N;
        p.j = __ldg(idx + (size_t)(agent_begin + t * TA) * K + r);
        p.si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        p.valid = true;
      }
    }
    return p;
  };
  auto load_sj = [&](const Pf& p) {
    return p.valid ? load_candidate(states, obstacles, p.b, p.j, N, M) : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  auto features = [&](int j, const Pf& p, const float4 sj) {
    float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    [[maybe_unused]] float mk = 0.f;
    if (p.valid) {
      const EdgeGeom e = edge_geom(p.si, sj, p.i, p.j);
      const bool in_range = ctrl_dist(e.dx, e.dy) < obs_radius;      // controller.py:38-39
      // out of range -> the agent's self edge (see the header).  The WIN variant must know WHICH slot won, and a
      // copy of the self edge in an earlier slot would win a tie against the real one (N <= K keeps index order, so
      // slot 0 is not always the agent itself): it keeps the true features and multiplies by the mask instead.
      const bool keep = WIN || in_range;
      f[0] = keep ? e.dx : 0.f; f[1] = keep ? e.dy : 0.f;
      f[2] = keep ? e.dvx : 0.f; f[3] = keep ? e.dvy : 0.f;
      f[4] = keep ? e.eye : 1.f;
      mk = in_range ? 0.f : __int_as_float(0xff800000);   // additive: 0 inside the radius, -inf outside
    }
    // 8-deep ring: features(j + 8) follows d1_full(j + 6) -> layer 2 of tile j + 4 issued -> d2_free(j + 2) -> the
    // pooling of tile j (the reader) is over
    if constexpr (WIN) sm[oMask + (j & 7) * 128 + r] = mk;
    if (warp == 0) CE2_TRACE(0, 8);
    float hi[8], lo[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) tf32_split_trunc(f[e], hi[e], lo[e]);
    float* Xhi = sm + oX + (j & 1) * (2 * 128 * 8);
    float* Xlo = Xhi + 128 * 8;
    const int sw = (r >> 2) & 1;
    *reinterpret_cast<float4*>(Xhi + r * 8 + ((0 ^ sw) << 2)) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(Xhi + r * 8 + ((1 ^ sw) << 2)) = make_float4(hi[4], hi[5], hi[6], hi[7]);
    *reinterpret_cast<float4*>(Xlo + r * 8 + ((0 ^ sw) << 2)) = make_float4(lo[0], lo[1], lo[2], lo[3]);
    *reinterpret_cast<float4*>(Xlo + r * 8 + ((1 ^ sw) << 2)) = make_float4(lo[4], lo[5], lo[6], lo[7]);
    if (warp == 0) CE2_TRACE(0, 9);
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&x_full[j & 1]);
  };
  if (warp >= WORKERS / 32) {
    // ==================== MMA-issue warp p: the tiles j = p, p + 2, ... (buffers p) ====================
    const int p = warp - WORKERS / 32;
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    const uint64_t d_w1hi = smem_desc_sw32(smem_u32(W1hi)), d_w1lo = smem_desc_sw32(smem_u32(W1lo));
    const uint64_t d_xhi = smem_desc_sw32(smem_u32(sm + oX + p * (2 * 128 * 8)));
    const uint64_t d_xlo = smem_desc_sw32(smem_u32(sm + oX + p * (2 * 128 * 8) + 128 * 8));
    const uint64_t d_a2hi = smem_desc_sw128(smem_u32(sm + oA2 + p * (2 * 128 * 64)));
    const uint64_t d_a2lo = smem_desc_sw128(smem_u32(sm + oA2 + p * (2 * 128 * 64) + 128 * 64));
    const uint32_t a2hi_lo = (uint32_t)d_a2hi, a2lo_lo = (uint32_t)d_a2lo, a2_hi32 = (uint32_t)(d_a2hi >> 32);
    constexpr uint32_t IDESC1 = idesc_tf32(128, 64), IDESC2 = idesc_tf32(128, 128);
    const uint32_t d1 = tb + C_D1 + p * 64, d2 = tb + C_D2 + p * 128;
    auto issue_layer1 = [&](int j) {                  // D1[p] is free: a2_full of this buffer's previous tile was observed
      mbar_wait(&x_full[p], (j >> 1) & 1);
      fence_after_sync();
      if (elect_one()) {
        mma_tf32_ss(d1, d_xlo, d_w1hi, IDESC1, 0u);
        mma_tf32_ss(d1, d_xhi, d_w1lo, IDESC1, 1u);
        mma_tf32_ss(d1, d_xhi, d_w1hi, IDESC1, 1u);
        mma_commit(&d1_full[p]);
      }
      __syncwarp();
    };
    if (p < nj) issue_layer1(p);
    for (int j = p; j < nj; j += 2) {
      mbar_wait(&a2_full[p], (j >> 1) & 1);
      if (j >= 2) mbar_wait(&d2_free[p], ((j - 2) >> 1) & 1);     // the pool warps have read this buffer's previous tile
      fence_after_sync();
      CE2_TRACE(1 + 2 * p, 1);
      if (elect_one()) {
#pragma unroll
        for (int term = 0; term < 3; ++term) {           // W2hi * a_lo, W2lo * a_hi, W2hi * a_hi
          const uint32_t a_col = tb + ((term == 1) ? C_W2LO : C_W2HI);
          const uint32_t b_lo = (term == 0) ? a2lo_lo : a2hi_lo;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t adv = ((ks >> 2) * (128 * 128) + (ks & 3) * 32) >> 4;
            mma_tf32_ts_split(d2, a_col + ks * 8, b_lo + adv, a2_hi32, IDESC2, (term | ks) ? 1u : 0u);
          }
        }
        mma_commit(&d2_full[p]);
      }
      __syncwarp();
      CE2_TRACE(1 + 2 * p, 2);
      if (j + 2 < nj) issue_layer1(j + 2);
      CE2_TRACE(1 + 2 * p, 3);
    }
  } else if (warp < ROW_WARPS) {
    // ===== row warps: thread = (edge row r, channel half); layer-1 epilogue of tile j; half 0 also builds X of tile j + 2 =====
    const int q = warp & 3, half = warp >> 2;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    float b1r[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) b1r[c] = B1[half * 32 + c];
    // feature group of the tiles of parity `half`: row warps 4-7 always build the odd tiles; row warps 0-3 build the
    // even ones in the WIN variant, whose pooling is heavier (there the pool warps only pool)
    const bool fgrp = WIN || (half == 1);
    Pf cur, nxt;
    float4 cur_sj = make_float4(0.f, 0.f, 0.f, 0.f);
    cur.valid = nxt.valid = false;
    if (fgrp) {
      cur = load_ij(half);
      cur_sj = load_sj(cur);
      if (half < nj) features(half, cur, cur_sj);
      cur = load_ij(half + 2);
      cur_sj = load_sj(cur);
      nxt = load_ij(half + 4);
    }
    for (int j = 0; j < nj; ++j) {
      const int b = j & 1;
      if (warp == 0) CE2_TRACE(0, 1);
      // D1 -> bias + ReLU + split -> channels [32 half, 32 half + 32) of row r of A2 = one 128-byte slab row
      mbar_wait(&d1_full[b], (j >> 1) & 1);
      if (j >= 2) mbar_wait(&d2_full[b], ((j - 2) >> 1) & 1);     // layer 2 of tile j - 2 has finished reading A2[b]
      fence_after_sync();
      if (warp == 0) CE2_TRACE(0, 2);
      float v[32];
      tmem_ld16(tlane + C_D1 + b * 64 + half * 32, v);
      tmem_ld16(tlane + C_D1 + b * 64 + half * 32 + 16, v + 16);
      wait_ld();
      float* hi_row = sm + oA2 + b * (2 * 128 * 64) + half * (128 * 32) + r * 32;
#pragma unroll
      for (int cc = 0; cc < 8; ++cc) {
        float2 h0, l0, h1, l1;
        split2_trunc(relu2(__fadd2_rn(make_float2(v[4 * cc + 0], v[4 * cc + 1]), make_float2(b1r[4 * cc + 0], b1r[4 * cc + 1]))), h0, l0);
        split2_trunc(relu2(__fadd2_rn(make_float2(v[4 * cc + 2], v[4 * cc + 3]), make_float2(b1r[4 * cc + 2], b1r[4 * cc + 3]))), h1, l1);
        const int pos = (cc ^ (r & 7)) << 2;
        *reinterpret_cast<float4*>(hi_row + pos) = make_float4(h0.x, h0.y, h1.x, h1.y);
        *reinterpret_cast<float4*>(hi_row + 128 * 64 + pos) = make_float4(l0.x, l0.y, l1.x, l1.y);
      }
      if (warp == 0) CE2_TRACE(0, 3);
      fence_proxy_async_smem();
      warp_arrive_after_tmem(&a2_full[b]);
      if (warp == 0) CE2_TRACE(0, 4);
      if (fgrp && (j & 1) == half) {
        if (j + 2 < nj) features(j + 2, cur, cur_sj); // layer 1 of tile j has retired (d1_full above): X[b] is free
        cur = nxt;
        cur_sj = load_sj(cur);                        // s_j of tile j + 4 (its idx arrived two iterations ago)
        nxt = load_ij(j + 6);
        if (warp == 4) CE2_TRACE(0, 5);
      }
    }
  } else {
    // ===================== pool warps: thread = channel; pooling of every tile =====================
    const int q = warp & 3;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const float b2c = B2[r];
    // feature group of the even tiles (not in the WIN variant): X tile 0 now, tile 2 on its way
    Pf cur, nxt;
    float4 cur_sj = make_float4(0.f, 0.f, 0.f, 0.f);
    cur.valid = nxt.valid = false;
    if constexpr (!WIN) {
      cur = load_ij(0);
      cur_sj = load_sj(cur);
      if (0 < nj) features(0, cur, cur_sj);
      cur = load_ij(2);
      cur_sj = load_sj(cur);
      nxt = load_ij(4);
    }
    for (int j = 0; j < nj; ++j) {
      const int b = j & 1;
      if (q == 0) CE2_TRACE(2, 1);
      if (!WIN && !(j & 1)) {                         // while layer 2 of tile j runs: the X tile of tile j + 2
        if (j + 2 < nj) {
          mbar_wait(&d1_full[b], (j >> 1) & 1);       // layer 1 of tile j has retired: X[b] is free
          features(j + 2, cur, cur_sj);
        }
        cur = nxt;
        cur_sj = load_sj(cur);
        nxt = load_ij(j + 6);
      }
      const long long t = blockIdx.x + (long long)j * gdim;
      const long long a0 = agent_begin + t * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      mbar_wait(&d2_full[b], (j >> 1) & 1);
      fence_after_sync();
      if (q == 0) CE2_TRACE(2, 2);
      const uint32_t tcol = tlane + C_D2 + b * 128;
      // WIN: masks of tile j, written by features(j) (row warps); the ring's reuse distance is argued there
      [[maybe_unused]] const float* mask_t = sm + oMask + (j & 7) * 128;
      const int c = r;
      if constexpr (KT == 12) {
        // 10 agents x 12 slots = columns [0,120), in two halves of 5 agents (loads aligned to their width); the buffer is
        // handed back to the MMA warp as soon as the second half sits in registers
        float v[60];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          if (hf == 0) {
            tmem_ld16(tcol + 0, v); tmem_ld16(tcol + 16, v + 16); tmem_ld16(tcol + 32, v + 32);
            tmem_ld8(tcol + 48, v + 48); tmem_ld4(tcol + 56, v + 56);
            wait_ld();
          } else {
            tmem_ld4(tcol + 60, v); tmem_ld16(tcol + 64, v + 4); tmem_ld16(tcol + 80, v + 20); tmem_ld16(tcol + 96, v + 36);
            tmem_ld8(tcol + 112, v + 52);
            wait_ld();
            warp_arrive_after_tmem(&d2_free[b]);
            if (q == 0) CE2_TRACE(2, 3);
          }
#pragma unroll
          for (int al = 0; al < 5; ++al) {
            const int a = 5 * hf + al;
            const float* x = v + 12 * al;
            float y;
            int bi = 0;
            if constexpr (WIN) {
              // slots outside the radius drop out through an additive -inf (the reference multiplies relu(.) by 0: the
              // same maximum, since the agent's own slot is always inside); bias and ReLU are monotone, so the max is
              // taken first; torch.max's arg-max is the FIRST slot that attains it
              const float4* mrow = reinterpret_cast<const float4*>(mask_t + 12 * a);
              const float4 m0 = mrow[0], m1 = mrow[1], m2 = mrow[2];
              float xp[12];
              {
                const float2 p0 = __fadd2_rn(make_float2(x[0], x[1]), make_float2(m0.x, m0.y));
                const float2 p1 = __fadd2_rn(make_float2(x[2], x[3]), make_float2(m0.z, m0.w));
                const float2 p2 = __fadd2_rn(make_float2(x[4], x[5]), make_float2(m1.x, m1.y));
                const float2 p3 = __fadd2_rn(make_float2(x[6], x[7]), make_float2(m1.z, m1.w));
                const float2 p4 = __fadd2_rn(make_float2(x[8], x[9]), make_float2(m2.x, m2.y));
                const float2 p5 = __fadd2_rn(make_float2(x[10], x[11]), make_float2(m2.z, m2.w));
                xp[0] = p0.x; xp[1] = p0.y; xp[2] = p1.x; xp[3] = p1.y; xp[4] = p2.x; xp[5] = p2.y;
                xp[6] = p3.x; xp[7] = p3.y; xp[8] = p4.x; xp[9] = p4.y; xp[10] = p5.x; xp[11] = p5.y;
              }
              const float m = fmaxf(fmaxf(fmaxf(fmaxf(xp[0], xp[1]), fmaxf(xp[2], xp[3])), fmaxf(fmaxf(xp[4], xp[5]), fmaxf(xp[6], xp[7]))),
                                    fmaxf(fmaxf(xp[8], xp[9]), fmaxf(xp[10], xp[11])));
              y = fmaxf(m + b2c, 0.f);
#pragma unroll
              for (int k = 11; k >= 0; --k)
                if (xp[k] == m) bi = k;
            } else {
              const float m = fmaxf(fmaxf(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])), fmaxf(fmaxf(x[4], x[5]), fmaxf(x[6], x[7]))),
                                    fmaxf(fmaxf(x[8], x[9]), fmaxf(x[10], x[11])));
              y = fmaxf(m + b2c, 0.f);
            }
            if (a < na) {
              pool[(a0 + a) * 128 + c] = y;
              if constexpr (WIN) argk[(a0 + a) * 128 + c] = (y > 0.f) ? (uint8_t)bi : (uint8_t)255;
            }
          }
        }
      } else {
        for (int a = 0; a < TA; ++a) {
          float y = 0.f;
          int bi = 0;
          float m = __int_as_float(0xff800000);
          if ((K & 3) == 0) {
            for (int k = 0; k < K; k += 4) {
              float x[4];
              tmem_ld4(tcol + a * K + k, x);
              wait_ld();
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if constexpr (WIN) {
                  const float xp = x[e] + mask_t[a * K + k + e];
                  if (k + e == 0 || xp > m) { m = xp; bi = k + e; }
                } else {
                  m = fmaxf(m, x[e]);
                }
              }
            }
          } else {
            for (int k = 0; k < K; ++k) {
              float x;
              tmem_ld1(tcol + a * K + k, &x);
              wait_ld();
              if constexpr (WIN) {
                const float xp = x + mask_t[a * K + k];
                if (k == 0 || xp > m) { m = xp; bi = k; }
              } else {
                m = fmaxf(m, x);
              }
            }
          }
          y = fmaxf(m + b2c, 0.f);
          if (a < na) {
            pool[(a0 + a) * 128 + c] = y;
            if constexpr (WIN) argk[(a0 + a) * 128 + c] = (y > 0.f) ? (uint8_t)bi : (uint8_t)255;
          }
        }
        warp_arrive_after_tmem(&d2_free[b]);
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) CE2_TRACE(0, 12);
  if (warp == 0) tmem_dealloc(tbase, 512);
}

template <bool WIN, int KT>
static int launch_ce2(const float* states, const float* obstacles, const int32_t* idx, const float* params, int N, int M, int K,
                      long long agents, int agent_begin, int TA, float obs_radius, float* pool, cudaStream_t st, uint8_t* argk,
                      int grid) {
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_tc_kernel<WIN, KT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)cet2::smem_bytes));
  ctrl_edge_tc_kernel<WIN, KT><<<grid, cet2::THREADS, cet2::smem_bytes, st>>>(states, obstacles, idx, params, N, M, K, agents, agent_begin, TA,
                                                                             obs_radius, pool, tl_skip_flag, argk);
  return 0;
}

int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "controller_forward: B*N must be < 2^31");
  const int TA = max(1, cet2::R / K);
  long long a_end = agents;
  int a_begin = 0;
  if (rows_active()) {
    MACBF_CHECK_ARG(B == 1, "controller_forward: a row range needs B = 1");
    a_begin = tl_row_range.begin;
    a_end = tl_row_range.end < agents ? tl_row_range.end : agents;
    if (a_end <= a_begin) return 0;
  }
  const long long ntiles = (a_end - a_begin + TA - 1) / TA;
  const int grid = (int)(ntiles < NUM_SMS ? ntiles : (long long)NUM_SMS);
  const float rad = (float)obs_radius;
  int rc;
  if (K == 12) rc = argk ? launch_ce2<true, 12>(states, obstacles, idx, params, N, M, K, a_end, a_begin, TA, rad, pool, st, argk, grid)
                         : launch_ce2<false, 12>(states, obstacles, idx, params, N, M, K, a_end, a_begin, TA, rad, pool, st, argk, grid);
  else rc = argk ? launch_ce2<true, 0>(states, obstacles, idx, params, N, M, K, a_end, a_begin, TA, rad, pool, st, argk, grid)
                 : launch_ce2<false, 0>(states, obstacles, idx, params, N, M, K, a_end, a_begin, TA, rad, pool, st, argk, grid);
  if (rc) return rc;
  MACBF_CHECK_LAUNCH("ctrl_edge_tc_kernel");
  return 0;
}

}  // namespace macbf
