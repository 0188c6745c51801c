// K3b' (tensor-core path) - backward of the controller edge net + masked max pooling (autograd of
// controller.py:33-46, REPAIRED), given d/dpool from the node backward and, from the recomputed forward
// (ctrl_edge_tc with `argk`), the winning slot of every (agent, channel) (255 = the maximum is 0: no gradient).
//
// With the winners known, layer 2 needs no recomputation: dZ2[row][c] = d/dpool[agent][c] if slot(row) == winner
// else 0 (torch.max backward through "* mask" and ReLU).  ctrl_edge_bwd_tc_kernel, TWO co-resident CTAs per SM
// (8 worker warps + 1 MMA-issue warp, 256 TMEM columns each - they hide each other's latencies), per tile of
// 128 edge rows (floor(128/K) agents; thread = row r, 32 of every 64 channels):
//   workers   edge features (controller.py:33-40) -> record; ReLU bits of a1 = relu(W1 x + b1) (5 -> 64, CUDA cores)
//   workers   per 64-channel chunk: dZ2 from winners + d/dpool: tf32 hi/lo -> TMEM
//   MMA       dA1 += dZ2[:, chunk] W2[chunk, :]  (A from TMEM, B = W2^T in shared memory, 3xTF32)
//   workers   dZ1 = dA1 [z1 > 0]; dW1, db1; d/dstates through [dx, dy, dvx, dvy] (fp32 atomics)
// dW2 = dZ2^T a1 and db2 are GEMMs over the EDGES: wgrad_tc_kernel<1> (cbf_bwd_tc.cu) rebuilds both operand images on
// chip (a1 from the recorded edge features, dZ2 from d/dpool and the winners: bf16x3, fp32 accumulation in TMEM).
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

int wgrad_tc_ctrl(const unsigned char* rec, size_t rec_stride, long long ntiles, float* slices, int grid, cudaStream_t st,
                  const float* params, uint32_t xoff, const float* dpool, const uint8_t* argk, int K, int TA, long long agents,
                  long long tile0);

namespace cebtc {
constexpr int WORKERS = 256;
constexpr int THREADS = WORKERS + 32;
constexpr int R = 128;
// TMEM (256 columns): [0,64) dZ2_hi of the current chunk  [64,128) dZ2_lo  [128,192) dA1
constexpr uint32_t C_ZHI = 0, C_ZLO = 64, C_DA1 = 128;
// record: just the edge features float[128][8] - the wgrad kernel rebuilds both of its operand images (a1 from the
// features, dZ2 from d/dpool and the winners)
constexpr uint32_t XREC = 0, REC_BYTES = 128 * 8 * 4;
constexpr int DZS = 132;                       // 16-B aligned rows, = 4 (mod 32) words: conflict-free stores and LDS.128
constexpr int oW2THI = 0;                      // B[n = in (64)][k = out (128)] = W2[k][n]
constexpr int oW2TLO = oW2THI + 64 * 128;
constexpr int oDz1 = oW2TLO + 64 * 128;        // dZ1^T stage [64][129]
constexpr int oXs = oDz1 + 64 * DZS;           // [8][128]: per feature (f0..f4, then the bias column of ones) every row
constexpr int oXpart = oXs + 128 * 8;          // [2][4][128]
constexpr int oW1 = oXpart + 2 * 4 * 128;      // W1 plain [64][8] (5 used), b1 [64]
constexpr int oB1 = oW1 + 64 * 8;
constexpr int oW1T = oB1 + 64;                 // W1 feature-major [5][64]: four channels of one feature per LDS.128 (packed FMAs)
constexpr int total_floats = oW1T + 5 * 64;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
static_assert(smem_bytes <= 113 * 1024, "two CTAs per SM must fit");
}  // namespace cebtc

__global__ void __launch_bounds__(cebtc::THREADS, 2)
ctrl_edge_bwd_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                        const int32_t* __restrict__ idx, const float* __restrict__ params,
                        const float* __restrict__ dpool, const uint8_t* __restrict__ argk, int N, int M, int K,
                        long long agents_total, int TA, long long tile_begin, long long tile_end,
                        unsigned char* __restrict__ rec, float* __restrict__ slices, float* __restrict__ grad_states) {
  using namespace cebtc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2Thi = sm + oW2THI; float* W2Tlo = sm + oW2TLO;
  float* dz1s = sm + oDz1; float* xs = sm + oXs; float* xpart = sm + oXpart; float* W1 = sm + oW1; float* B1 = sm + oB1;
  float* W1T = sm + oW1T;
  // workers -> MMA: z_full (8 arrivals, once per chunk); MMA -> workers: c_done (chunk's MMAs retired, once per chunk)
  __shared__ uint64_t z_full, c_done;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  if (threadIdx.x == 32) {
    mbar_init(&z_full, 8); mbar_init(&c_done, 1);
    fence_mbar_init();
  }
  {
    WeightFill<64, 128, THREADS> f2;
    f2.template load<true>(params + CT_EW2, 1, 64);
    f2.template store<true>(W2Thi, W2Tlo);
  }
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) W1[t] = ((t & 7) < 5) ? __ldg(params + CT_EW1 + (t >> 3) * 5 + (t & 7)) : 0.f;
  for (int t = threadIdx.x; t < 64; t += THREADS) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 5 * 64; t += THREADS) W1T[t] = __ldg(params + CT_EW1 + (t & 63) * 5 + (t >> 6));
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long first = tile_begin + blockIdx.x;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    const uint64_t d_wthi = smem_desc_sw128(smem_u32(W2Thi)), d_wtlo = smem_desc_sw128(smem_u32(W2Tlo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC = idesc_tf32(128, 64);
    uint32_t n = 0;                                     // chunk counter: z_full / c_done complete once per chunk
    for (long long tile = first; tile < tile_end; tile += gridDim.x) {
#pragma unroll
      for (int c = 0; c < 2; ++c, ++n) {
        mbar_wait(&z_full, n & 1);
        fence_after_sync();
#pragma unroll
        for (int term = 0; term < 3; ++term) {          // lo*hi, hi*lo, hi*hi of K chunk [64c, 64c+64)
          const uint32_t a_col = tb + (term == 0 ? C_ZLO : C_ZHI);
          const uint64_t b_desc = desc_advance((term == 1) ? d_wtlo : d_wthi, (2 * c) * (64 * 128));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (64 * 128) + (ks & 3) * 32);
            if (elect_one()) mma_tf32_ts(tb + C_DA1, a_col + ks * 8, bd, IDESC, (c | term | ks) ? 1u : 0u);
          }
        }
        if (elect_one()) mma_commit(&c_done);
      }
    }
  } else {
    // ================================= worker warps =================================
    const int q = warp & 3, cgrp = warp >> 2;          // lane quadrant, half of every 64-channel chunk
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int c0 = cgrp * 32;
    const int a_loc = r / K, k_loc = r - a_loc * K;
    const int wc = threadIdx.x & 63, wf = threadIdx.x >> 6;     // dW1 / db1 outputs (wc, wf) and (wc, wf + 4); column 5 = bias
    float accw1a = 0.f, accw1b = 0.f;
    uint32_t n = 0;
    for (long long tile = first; tile < tile_end; tile += gridDim.x) {
      unsigned char* tr = rec + (size_t)(tile - tile_begin) * REC_BYTES;
      const long long a0 = tile * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      const bool live = a_loc < na;
      const long long ga = a0 + (live ? a_loc : 0);
      // ---- this row's d/dpool and winners (issued first: they are only needed after the layer-1 math) ----
      uint4 wk[2][2];
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          wk[c][h] = live ? __ldg(reinterpret_cast<const uint4*>(argk + ga * 128 + c * 64 + c0) + h) : make_uint4(~0u, ~0u, ~0u, ~0u);
      if (live) {   // this thread's two 128-byte lines of d/dpool: on their way to L1 while layer 1 is evaluated
        asm volatile("prefetch.global.L1 [%0];" ::"l"(dpool + ga * 128 + c0));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(dpool + ga * 128 + 64 + c0));
      }
      // ---- edge features (controller.py:33-40) ----
      float f[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      long long gj = -1;
      if (live) {
        const int b = (int)(ga / N), i = (int)(ga - (long long)b * N);
        const int j = __ldg(idx + (size_t)a0 * K + r);
        const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
        const float4 sj = load_candidate(states, obstacles, b, j, N, M);
        const EdgeGeom g = edge_geom(si, sj, i, j);
        f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
        gj = (j < N) ? (long long)b * N + j : -1;
      }
      if (cgrp == 0) {
#pragma unroll
        for (int ff = 0; ff < 5; ++ff) xs[ff * R + r] = f[ff];
        xs[5 * R + r] = 1.f;
      }
      if (cgrp == 0) {   // the wgrad kernel rebuilds the a1 images from these (same fmaf chain as below)
        float4* xr = reinterpret_cast<float4*>(tr + XREC) + 2 * r;
        xr[0] = make_float4(f[0], f[1], f[2], f[3]);
        xr[1] = make_float4(f[4], 0.f, 0.f, 0.f);
      }
      // ---- ReLU bits of a1 = relu(W1 x + b1), channels c0 .. c0+31 ----
      // (the forward's fmaf chain b1 + w0 f0 + .. + w4 f4 per channel, four channels at a time as packed f32x2 FMAs: the
      // same roundings at half the instructions)
      uint32_t mb1 = 0;
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        const float4 bb = *reinterpret_cast<const float4*>(B1 + c0 + e);
        float2 za = make_float2(bb.x, bb.y), zb = make_float2(bb.z, bb.w);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const float4 w = *reinterpret_cast<const float4*>(W1T + k * 64 + c0 + e);
          const float2 fk = make_float2(f[k], f[k]);
          za = __ffma2_rn(make_float2(w.x, w.y), fk, za);
          zb = __ffma2_rn(make_float2(w.z, w.w), fk, zb);
        }
        mb1 |= (za.x > 0.f ? 1u : 0u) << e;
        mb1 |= (za.y > 0.f ? 1u : 0u) << (e + 1);
        mb1 |= (zb.x > 0.f ? 1u : 0u) << (e + 2);
        mb1 |= (zb.y > 0.f ? 1u : 0u) << (e + 3);
      }
      // ---- per 64-channel chunk: dZ2 from the winners -> TMEM hi/lo + image; the MMA warp accumulates dA1 ----
#pragma unroll
      for (int c = 0; c < 2; ++c, ++n) {
        if (n > 0) {                                    // the previous chunk's MMAs have read the dZ2 columns
          mbar_wait(&c_done, (n - 1) & 1);
          fence_after_sync();
        }
#pragma unroll
        for (int pair = 0; pair < 2; ++pair) {
          float4 dp[4];
#pragma unroll
          for (int v4 = 0; v4 < 4; ++v4)
            dp[v4] = live ? __ldg(reinterpret_cast<const float4*>(dpool + ga * 128 + c * 64 + c0 + pair * 16) + v4)
                          : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            float x[8], hi[8], lo[8];
            const uint4 w4 = wk[c][pair];
            const uint32_t wlo = half ? w4.z : w4.x, whi = half ? w4.w : w4.y;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const uint32_t wbyte = ((e < 4 ? wlo : whi) >> (8 * (e & 3))) & 0xffu;
              const float4 d4 = dp[half * 2 + (e >> 2)];
              const float dv = (e & 3) == 0 ? d4.x : (e & 3) == 1 ? d4.y : (e & 3) == 2 ? d4.z : d4.w;
              x[e] = (wbyte == (uint32_t)k_loc) ? dv : 0.f;
            }
            tf32_split_trunc8(x, hi, lo);
            tmem_st8(tlane + C_ZHI + c0 + pair * 16 + half * 8, hi);
            tmem_st8(tlane + C_ZLO + c0 + pair * 16 + half * 8, lo);
          }
        }
        wait_st();
        warp_arrive_after_tmem(&z_full);
      }
      // ---- dZ1 = dA1 * [z1 > 0]; d/dfeatures partials; stage dZ1^T for dW1 ----
      mbar_wait(&c_done, (n - 1) & 1);
      fence_after_sync();
      {
        float2 p01 = make_float2(0.f, 0.f), p23 = p01;       // d/d(f0, f1), d/d(f2, f3): packed f32x2 FMAs
#pragma unroll
        for (int part = 0; part < 2; ++part) {
          float v[16];
          tmem_ld16(tlane + C_DA1 + c0 + part * 16, v);
          wait_ld();
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int ch = c0 + part * 16 + e;
            const float dz = ((mb1 >> (part * 16 + e)) & 1u) ? v[e] : 0.f;
            dz1s[ch * DZS + r] = dz;
            const float4 wa = *reinterpret_cast<const float4*>(W1 + ch * 8);
            const float2 dz2 = make_float2(dz, dz);
            p01 = __ffma2_rn(dz2, make_float2(wa.x, wa.y), p01); p23 = __ffma2_rn(dz2, make_float2(wa.z, wa.w), p23);
          }
        }
#pragma unroll
        for (int ff = 0; ff < 4; ++ff) xpart[(cgrp * 4 + ff) * R + r] = ff == 0 ? p01.x : ff == 1 ? p01.y : ff == 2 ? p23.x : p23.y;
      }
      fence_before_sync();
      named_bar_sync(1, WORKERS);
      {
        const float4* dzr = reinterpret_cast<const float4*>(dz1s + wc * DZS);
        const float4* xr = reinterpret_cast<const float4*>(xs + wf * R);
        const int wf2 = wf + 4;                         // 4, 5 for wf = 0, 1; idle otherwise
        const float4* xr2 = reinterpret_cast<const float4*>(xs + (wf2 < 6 ? wf2 : wf) * R);
        float2 s01 = make_float2(0.f, 0.f), t01 = s01;      // packed f32x2 FMAs: (even rows, odd rows) partial sums
#pragma unroll 8
        for (int e4 = 0; e4 < R / 4; ++e4) {
          const float4 d = dzr[e4], x = xr[e4];
          s01 = __ffma2_rn(make_float2(d.x, d.y), make_float2(x.x, x.y), s01);
          s01 = __ffma2_rn(make_float2(d.z, d.w), make_float2(x.z, x.w), s01);
          if (wf2 < 6) {
            const float4 y = xr2[e4];
            t01 = __ffma2_rn(make_float2(d.x, d.y), make_float2(y.x, y.y), t01);
            t01 = __ffma2_rn(make_float2(d.z, d.w), make_float2(y.z, y.w), t01);
          }
        }
        accw1a += s01.x + s01.y;
        accw1b += t01.x + t01.y;
      }
      if (grad_states != nullptr && cgrp == 0 && live && gj != ga) {   // the self edge contributes +g - g = 0
        float d0[4];
#pragma unroll
        for (int ff = 0; ff < 4; ++ff) d0[ff] = xpart[(0 * 4 + ff) * R + r] + xpart[(1 * 4 + ff) * R + r];
        float* pi = grad_states + (size_t)ga * 4;
        atomicAdd(pi + 0, d0[0]); atomicAdd(pi + 1, d0[1]); atomicAdd(pi + 2, d0[2]); atomicAdd(pi + 3, d0[3]);
        if (gj >= 0) {
          float* pj = grad_states + (size_t)gj * 4;
          atomicAdd(pj + 0, -d0[0]); atomicAdd(pj + 1, -d0[1]); atomicAdd(pj + 2, -d0[2]); atomicAdd(pj + 3, -d0[3]);
        }
      }
      named_bar_sync(1, WORKERS);                      // xs / xpart / dz1s are rewritten by the next tile
    }
    float* gp = slices + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;
    gp[CT_EW1 + wc * 5 + wf] += accw1a;                // wf = 0..3
    if (wf == 0) gp[CT_EW1 + wc * 5 + 4] += accw1b;
    else if (wf == 1) gp[CT_EB1 + wc] += accw1b;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

// ---- host side ----
namespace {
constexpr long long CE_CHUNK_TILES = 1 << 20;   // records are 4 KB per tile
constexpr int CE_GRID = 2 * NUM_SMS;           // two co-resident CTAs per SM
static inline size_t a1k(size_t x) { return (x + 1023) & ~(size_t)1023; }
inline long long chunk_tiles() {               // MACBF_BWD_CHUNK_TILES=<n>: see cbf_bwd_tc.cu
  const char* e = getenv("MACBF_BWD_CHUNK_TILES");
  const long long v = e ? atoll(e) : 0;
  return (v > 0 && v < CE_CHUNK_TILES) ? v : CE_CHUNK_TILES;
}
}
bool ctrl_edge_bwd_tc_applicable(int K) { return K >= 1 && K <= 128; }
int ctrl_edge_bwd_tc_slices() { return CE_GRID; }

// [argk: B*N*128 bytes][tile records]
size_t ctrl_edge_bwd_tc_scratch_bytes(int B, int N, int K) {
  const int TA = max(1, 128 / K);
  const long long nt = ((long long)B * N + TA - 1) / TA;
  const long long ct = nt < CE_CHUNK_TILES ? nt : CE_CHUNK_TILES;
  return a1k((size_t)B * N * 128) + (size_t)ct * cebtc::REC_BYTES + 2048;
}
uint8_t* ctrl_edge_bwd_tc_argk(void* scratch) { return (uint8_t*)(((uintptr_t)scratch + 1023) & ~(uintptr_t)1023); }

// dpool: d/dpool [B*N,128]; argk: winners from ctrl_edge_tc; slices: CE_GRID gradient slices (already zeroed)
int ctrl_edge_bwd_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, const float* dpool,
                     int B, int N, int M, int K, float* slices, float* grad_states, void* scratch, cudaStream_t st,
                     const uint8_t* argk_saved) {
  const long long agents = (long long)B * N;
  const int TA = max(1, 128 / K);
  const long long nt = (agents + TA - 1) / TA;
  const uint8_t* argk_own = ctrl_edge_bwd_tc_argk(scratch);
  unsigned char* rec = (unsigned char*)argk_own + a1k((size_t)agents * 128);
  const uint8_t* argk = argk_saved ? argk_saved : argk_own;      // winners kept by the training forward, or recomputed
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cebtc::smem_bytes));
  const long long CH = chunk_tiles();
  for (long long t0 = 0; t0 < nt; t0 += CH) {
    const long long t1 = (t0 + CH < nt) ? t0 + CH : nt;
    const long long ct = t1 - t0;
    const int grid = (int)(ct < CE_GRID ? ct : (long long)CE_GRID);
    ctrl_edge_bwd_tc_kernel<<<grid, cebtc::THREADS, cebtc::smem_bytes, st>>>(states, obstacles, idx, params, dpool, argk, N, M, K,
                                                                           agents, TA, t0, t1, rec, slices, grad_states);
    MACBF_CHECK_LAUNCH("ctrl_edge_bwd_tc_kernel");
    const int wgrid = (int)(ct < NUM_SMS ? ct : (long long)NUM_SMS);
    if (int rc = wgrad_tc_ctrl(rec, cebtc::REC_BYTES, ct, slices, wgrid, st, params, cebtc::XREC, dpool, argk, K, TA, agents, t0))
      return rc;
  }
  return 0;
}

}  // namespace macbf
