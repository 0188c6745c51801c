// K3b' (tensor-core path) - backward of the controller edge net + masked max pooling (autograd of
// controller.py:33-46, REPAIRED), given d/dpool from the node backward.
//
// ctrl_edge_bwd_tc_kernel, one persistent CTA per SM (16 worker warps + 1 MMA-issue warp), per tile of
// 128 edge rows (floor(128/K) agents):
//   feature warps    edge features (controller.py:33-40) -> tf32 hi/lo -> smem A0
//   MMA              layer 1  D1 = A0 W1^T (SS, K = 8)
//   workers          a1 = relu(D1 + b1): hi/lo -> TMEM (layer-2 operand), bf16 hi/lo image (wgrad operand), ReLU bits
//   MMA              layer 2  D2 = a1 W2^T (A from TMEM), two 64-channel chunks
//   workers          per chunk: y = relu(D2 + b2) -> smem stage; per (agent, channel) the first arg-max over the
//                    K slots of y * mask (torch.max backward); dZ2[row][c] = d/dpool[agent][c] on the winning row if
//                    the winner is > 0, else 0: hi/lo -> TMEM in place of D2, bf16 image
//   MMA              dA1 = dZ2 W2 (A from TMEM, B = W2^T resident next to W2), cross terms per chunk, hi*hi last
//   workers          dZ1 = dA1 [z1 > 0]; dW1, db1; d/dstates through [dx, dy, dvx, dvy] (fp32 atomics)
// dW2 = dZ2^T a1 and db2 are GEMMs over the EDGES: wgrad_tc_kernel<1> (cbf_bwd_tc.cu) reads the two images
// with bulk copies.  3xTF32 everywhere except the bf16x3 weight-gradient GEMM (see cbf_bwd_tc.cu).
#include "common.cuh"
#include "tc_common.cuh"

namespace macbf {

int wgrad_tc_ctrl(const unsigned char* rec, size_t rec_stride, long long ntiles, float* slices, int grid, cudaStream_t st);

namespace cebtc {
constexpr int WORKERS = 512;
constexpr int THREADS = WORKERS + 32;
constexpr int R = 128;
constexpr int MAX_TA = 32;                     // K >= 4
// TMEM: [0,64) D1 -> a1_hi  [64,128) a1_lo  [128,256) D2 -> dZ2_hi  [256,384) dZ2_lo  [384,448) dA1
constexpr uint32_t C_A1HI = 0, C_A1LO = 64, C_D2 = 128, C_Z2LO = 256, C_DA1 = 384;
constexpr uint32_t BLK = 128 * 128;
constexpr uint32_t Z2_HI = 0, Z2_LO = 2 * BLK, A1_HI = 4 * BLK, A1_LO = 5 * BLK, REC_BYTES = 6 * BLK;
constexpr int YS = 65, DZS = 129;
constexpr int oW2HI = 0;
constexpr int oW2LO = oW2HI + 128 * 64;
constexpr int oW2THI = oW2LO + 128 * 64;       // B[n = in (64)][k = out (128)] = W2[k][n]
constexpr int oW2TLO = oW2THI + 64 * 128;
constexpr int oW1HI = oW2TLO + 64 * 128;       // [64][32], k < 5 used
constexpr int oW1LO = oW1HI + 64 * 32;
constexpr int oA0HI = oW1LO + 64 * 32;         // [128][32]; after layer 1 has read it the region is reused:
constexpr int oA0LO = oA0HI + 128 * 32;
constexpr int oXs = oA0HI;                     //   xs    [128][8]   features + bias column
constexpr int oXpart = oXs + 128 * 8;          //   xpart [4][4][128]
constexpr int oGsel = oXpart + 4 * 4 * 128;    //   gsel  [MAX_TA][64] routed d/dpool of the current chunk
constexpr int oWin = oGsel + MAX_TA * 64;      //   win   [MAX_TA][64] winning slot (int)
static_assert(oWin + MAX_TA * 64 <= oA0LO + 128 * 32, "alias region overflow");
constexpr int oStage = oA0LO + 128 * 32;       // y stage [128][65]  /  dZ1^T stage [64][129]
constexpr int oW1P = oStage + 128 * YS;        // W1 plain [64][8]
constexpr int oB1 = oW1P + 64 * 8;
constexpr int oB2 = oB1 + 64;
constexpr int oMk = oB2 + 128;                 // radius mask per row
constexpr int total_floats = oMk + 128;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cebtc

__global__ void __launch_bounds__(cebtc::THREADS, 1)
ctrl_edge_bwd_tc_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                        const int32_t* __restrict__ idx, const float* __restrict__ params,
                        const float* __restrict__ dpool, int N, int M, int K, long long agents_total, int TA,
                        float obs_radius, long long tile_begin, long long tile_end, unsigned char* __restrict__ rec,
                        float* __restrict__ slices, float* __restrict__ grad_states) {
  using namespace cebtc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W2hi = sm + oW2HI; float* W2lo = sm + oW2LO; float* W2Thi = sm + oW2THI; float* W2Tlo = sm + oW2TLO;
  float* W1hi = sm + oW1HI; float* W1lo = sm + oW1LO; float* A0hi = sm + oA0HI; float* A0lo = sm + oA0LO;
  float* xs = sm + oXs; float* xpart = sm + oXpart; float* gsel = sm + oGsel; int* win = reinterpret_cast<int*>(sm + oWin);
  float* ystage = sm + oStage; float* dz1s = sm + oStage;
  float* W1 = sm + oW1P; float* B1 = sm + oB1; float* B2 = sm + oB2; float* mks = sm + oMk;
  __shared__ uint64_t a0_full, a1_full, z2_full[2], d1_full, d2_full[2], da1_full;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&a0_full, 4); mbar_init(&a1_full, 16); mbar_init(&z2_full[0], 16); mbar_init(&z2_full[1], 16);
    mbar_init(&d1_full, 1); mbar_init(&d2_full[0], 1); mbar_init(&d2_full[1], 1); mbar_init(&da1_full, 1);
    fence_mbar_init();
  }
  fill_weight_sw128<128, 64>(params + CT_EW2, 64, W2hi, W2lo);
  fill_weight_sw128_strided<64, 128>(params + CT_EW2, 1, 64, W2Thi, W2Tlo);
  for (int t = threadIdx.x; t < 64 * 32; t += THREADS) {
    const int n = t >> 5, pos = (t >> 2) & 7, e = t & 3;
    const int k = ((pos ^ (n & 7)) << 2) + e;
    float hi = 0.f, lo = 0.f;
    if (k < 5) tf32_split(__ldg(params + CT_EW1 + n * 5 + k), hi, lo);
    W1hi[t] = hi; W1lo[t] = lo;
  }
  for (int t = threadIdx.x; t < 128 * 32; t += THREADS) { A0hi[t] = 0.f; A0lo[t] = 0.f; }
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) W1[t] = ((t & 7) < 5) ? __ldg(params + CT_EW1 + (t >> 3) * 5 + (t & 7)) : 0.f;
  for (int t = threadIdx.x; t < 64; t += THREADS) B1[t] = __ldg(params + CT_EB1 + t);
  for (int t = threadIdx.x; t < 128; t += THREADS) B2[t] = __ldg(params + CT_EB2 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long first = tile_begin + blockIdx.x;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1hi)), d_w1lo = smem_desc_sw128(smem_u32(W1lo));
    const uint64_t d_a0hi = smem_desc_sw128(smem_u32(A0hi)), d_a0lo = smem_desc_sw128(smem_u32(A0lo));
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2hi)), d_w2lo = smem_desc_sw128(smem_u32(W2lo));
    const uint64_t d_wthi = smem_desc_sw128(smem_u32(W2Thi)), d_wtlo = smem_desc_sw128(smem_u32(W2Tlo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC = idesc_tf32(128, 64);
    uint32_t par = 0;
    for (long long tile = first; tile < tile_end; tile += gridDim.x, par ^= 1) {
      mbar_wait(&a0_full, par);
      fence_after_sync();
      if (elect_one()) mma_tf32_ss(tb + C_A1HI, d_a0lo, d_w1hi, IDESC, 0u);
      if (elect_one()) mma_tf32_ss(tb + C_A1HI, d_a0hi, d_w1lo, IDESC, 1u);
      if (elect_one()) mma_tf32_ss(tb + C_A1HI, d_a0hi, d_w1hi, IDESC, 1u);
      if (elect_one()) mma_commit(&d1_full);
      mbar_wait(&a1_full, par);
      fence_after_sync();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t a_col = tb + (term == 0 ? C_A1LO : C_A1HI);
          const uint64_t b_desc = desc_advance((term == 1) ? d_w2lo : d_w2hi, c * (64 * 128));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (128 * 128) + (ks & 3) * 32);
            if (elect_one()) mma_tf32_ts(tb + C_D2 + c * 64, a_col + ks * 8, bd, IDESC, (term | ks) ? 1u : 0u);
          }
        }
        if (elect_one()) mma_commit(&d2_full[c]);
      }
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&z2_full[c], par);
        fence_after_sync();
#pragma unroll
        for (int term = 0; term < 2; ++term) {
          const uint32_t a_col = tb + (term == 0 ? C_Z2LO : C_D2) + c * 64;
          const uint64_t b_desc = desc_advance((term == 1) ? d_wtlo : d_wthi, (2 * c) * (64 * 128));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (64 * 128) + (ks & 3) * 32);
            if (elect_one()) mma_tf32_ts(tb + C_DA1, a_col + ks * 8, bd, IDESC, (c | term | ks) ? 1u : 0u);
          }
        }
      }
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) {
        const uint64_t bd = desc_advance(d_wthi, (ks >> 2) * (64 * 128) + (ks & 3) * 32);
        if (elect_one()) mma_tf32_ts(tb + C_DA1, tb + C_D2 + ks * 8, bd, IDESC, 1u);
      }
      if (elect_one()) mma_commit(&da1_full);
    }
  } else {
    // ================================= worker warps =================================
    const int q = warp & 3, cgrp = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int c0 = cgrp * 16;
    const int a_loc = r / K, k_loc = r - a_loc * K;
    const int wc = threadIdx.x & 63, wf = threadIdx.x >> 6;     // dW1 / db1 output: channel wc, column wf (5 = bias)
    float accw1 = 0.f;
    // endpoints of the next tile are fetched while this one is processed (feature warps only)
    int pf_j = 0;
    float4 pf_si = make_float4(0.f, 0.f, 0.f, 0.f);
    bool pf_valid = false;
    auto prefetch = [&](long long t) {
      pf_valid = false;
      if (cgrp == 0 && t < tile_end) {
        const long long ga = t * TA + a_loc;
        if (a_loc < TA && ga < agents_total) {
          pf_j = __ldg(idx + (size_t)(t * TA) * K + r);
          pf_si = __ldg(reinterpret_cast<const float4*>(states) + ga);
          pf_valid = true;
        }
      }
    };
    prefetch(first);
    uint32_t par = 0;
    for (long long tile = first; tile < tile_end; tile += gridDim.x, par ^= 1) {
      unsigned char* tr = rec + (size_t)(tile - tile_begin) * REC_BYTES;
      const long long a0 = tile * TA;
      const int na = (int)min((long long)TA, agents_total - a0);
      const int rows = na * K;
      // ---- features -> A0 (feature warps) ----
      float f[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      long long gi = -1, gj = -1;
      if (cgrp == 0) {
        float mk = 0.f;
        if (pf_valid) {
          const long long ga = a0 + a_loc;
          const int b = (int)(ga / N), i = (int)(ga - (long long)b * N);
          const float4 sj = load_candidate(states, obstacles, b, pf_j, N, M);
          const EdgeGeom g = edge_geom(pf_si, sj, i, pf_j);
          f[0] = g.dx; f[1] = g.dy; f[2] = g.dvx; f[3] = g.dvy; f[4] = g.eye;
          mk = (ctrl_dist(g.dx, g.dy) < obs_radius) ? 1.f : 0.f;     // controller.py:38-40
          gi = ga;
          gj = (pf_j < N) ? (long long)b * N + pf_j : -1;
        }
        mks[r] = mk;
        float hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) tf32_split(e < 5 ? f[e] : 0.f, hi[e], lo[e]);
        const int p0 = (0 ^ (r & 7)) << 2, p1 = (1 ^ (r & 7)) << 2;
        *reinterpret_cast<float4*>(A0hi + r * 32 + p0) = make_float4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<float4*>(A0hi + r * 32 + p1) = make_float4(hi[4], hi[5], hi[6], hi[7]);
        *reinterpret_cast<float4*>(A0lo + r * 32 + p0) = make_float4(lo[0], lo[1], lo[2], lo[3]);
        *reinterpret_cast<float4*>(A0lo + r * 32 + p1) = make_float4(lo[4], lo[5], lo[6], lo[7]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a0_full);
        prefetch(tile + gridDim.x);
      }
      // ---- layer-1 epilogue: a1 -> TMEM hi/lo, image, ReLU bits ----
      uint32_t mb1 = 0;
      {
        mbar_wait(&d1_full, par);
        fence_after_sync();
        float v[16];
        tmem_ld16(tlane + C_A1HI + c0, v);
        wait_ld();
        uint4 bh[2], bl[2];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float x[8], hi[8], lo[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            x[e] = fmaxf(v[half * 8 + e] + B1[c0 + half * 8 + e], 0.f);
            mb1 |= (x[e] > 0.f ? 1u : 0u) << (half * 8 + e);
            tf32_split_trunc(x[e], hi[e], lo[e]);
          }
          tmem_st8(tlane + C_A1HI + c0 + half * 8, hi);
          tmem_st8(tlane + C_A1LO + c0 + half * 8, lo);
          bf16_split8(x, bh[half], bl[half]);
        }
        img_store16(tr + A1_HI, r, cgrp, bh[0], bh[1]);
        img_store16(tr + A1_LO, r, cgrp, bl[0], bl[1]);
        wait_st();
        warp_arrive_after_tmem(&a1_full);
      }
      // A0 has been read by layer 1 (d1_full observed): its storage now holds xs / xpart / gsel / win
      if (cgrp == 0) {
        *reinterpret_cast<float4*>(xs + r * 8) = make_float4(f[0], f[1], f[2], f[3]);
        *reinterpret_cast<float4*>(xs + r * 8 + 4) = make_float4(f[4], 1.f, 0.f, 0.f);
      }
      // ---- layer-2 epilogue + max-pool backward, per 64-channel chunk ----
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&d2_full[c], par);
        fence_after_sync();
        {
          float v[16];
          tmem_ld16(tlane + C_D2 + c * 64 + c0, v);
          wait_ld();
#pragma unroll
          for (int e = 0; e < 16; ++e) ystage[r * YS + c0 + e] = fmaxf(v[e] + B2[c * 64 + c0 + e], 0.f);
        }
        named_bar_sync(1, WORKERS);
        // first arg-max over the K slots of y * mask (controller.py:43-46); only a strictly positive winner
        // carries gradient (through "* mask" and relu')
        for (int item = threadIdx.x; item < na * 64; item += WORKERS) {
          const int cc = item & 63, a = item >> 6;
          const float* col = ystage + (a * K) * YS + cc;
          const float* mrow = mks + a * K;
          float best = col[0] * mrow[0];
          int bi = 0;
          for (int k = 1; k < K; ++k) {
            const float y = col[k * YS] * mrow[k];
            if (y > best) { best = y; bi = k; }
          }
          gsel[a * 64 + cc] = (best > 0.f) ? __ldg(dpool + (a0 + a) * 128 + c * 64 + cc) : 0.f;
          win[a * 64 + cc] = bi;
        }
        named_bar_sync(1, WORKERS);
        {
          uint4 bh[2], bl[2];
          const bool live = r < rows;
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            float x[8], hi[8], lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int cc = c0 + half * 8 + e;
              x[e] = (live && win[a_loc * 64 + cc] == k_loc) ? gsel[a_loc * 64 + cc] : 0.f;
              tf32_split(x[e], hi[e], lo[e]);
            }
            tmem_st8(tlane + C_D2 + c * 64 + c0 + half * 8, hi);
            tmem_st8(tlane + C_Z2LO + c * 64 + c0 + half * 8, lo);
            bf16_split8(x, bh[half], bl[half]);
          }
          img_store16(tr + Z2_HI + c * BLK, r, cgrp, bh[0], bh[1]);
          img_store16(tr + Z2_LO + c * BLK, r, cgrp, bl[0], bl[1]);
          wait_st();
          warp_arrive_after_tmem(&z2_full[c]);
        }
      }
      // ---- dZ1 = dA1 * [z1 > 0]; d/dfeatures partials; stage dZ1^T for dW1 (the y stage is free: every thread has
      //      passed the second barrier of chunk 1) ----
      mbar_wait(&da1_full, par);
      fence_after_sync();
      {
        float v[16];
        tmem_ld16(tlane + C_DA1 + c0, v);
        wait_ld();
        float px[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float dz = ((mb1 >> e) & 1u) ? v[e] : 0.f;
          dz1s[(c0 + e) * DZS + r] = dz;
          const float4 wa = *reinterpret_cast<const float4*>(W1 + (c0 + e) * 8);
          px[0] = fmaf(dz, wa.x, px[0]); px[1] = fmaf(dz, wa.y, px[1]); px[2] = fmaf(dz, wa.z, px[2]); px[3] = fmaf(dz, wa.w, px[3]);
        }
#pragma unroll
        for (int ff = 0; ff < 4; ++ff) xpart[(cgrp * 4 + ff) * R + r] = px[ff];
      }
      fence_before_sync();
      named_bar_sync(1, WORKERS);
      if (wf < 6) {
        const float* dzr = dz1s + wc * DZS;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
        for (int e = 0; e < R; e += 4) {
          s0 = fmaf(dzr[e + 0], xs[(e + 0) * 8 + wf], s0);
          s1 = fmaf(dzr[e + 1], xs[(e + 1) * 8 + wf], s1);
          s2 = fmaf(dzr[e + 2], xs[(e + 2) * 8 + wf], s2);
          s3 = fmaf(dzr[e + 3], xs[(e + 3) * 8 + wf], s3);
        }
        accw1 += (s0 + s1) + (s2 + s3);
      }
      if (grad_states != nullptr && cgrp == 0 && r < rows && gi != gj) {   // the self edge contributes +g - g = 0
        float d0[4];
#pragma unroll
        for (int ff = 0; ff < 4; ++ff)
          d0[ff] = (xpart[(0 * 4 + ff) * R + r] + xpart[(1 * 4 + ff) * R + r]) + (xpart[(2 * 4 + ff) * R + r] + xpart[(3 * 4 + ff) * R + r]);
        float* pi = grad_states + (size_t)gi * 4;
        atomicAdd(pi + 0, d0[0]); atomicAdd(pi + 1, d0[1]); atomicAdd(pi + 2, d0[2]); atomicAdd(pi + 3, d0[3]);
        if (gj >= 0) {
          float* pj = grad_states + (size_t)gj * 4;
          atomicAdd(pj + 0, -d0[0]); atomicAdd(pj + 1, -d0[1]); atomicAdd(pj + 2, -d0[2]); atomicAdd(pj + 3, -d0[3]);
        }
      }
      named_bar_sync(1, WORKERS);                      // xs / xpart / stage / A0 are rewritten by the next tile
    }
    float* gp = slices + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;
    if (wf < 5) gp[CT_EW1 + wc * 5 + wf] += accw1;
    else if (wf == 5) gp[CT_EB1 + wc] += accw1;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

// ---- host side ----
namespace {
constexpr long long CE_CHUNK_TILES = 16384;    // 16384 x 96 KB = 1.5 GB of records per chunk
}
bool ctrl_edge_bwd_tc_applicable(int K) { return K >= 4 && K <= 128; }

size_t ctrl_edge_bwd_tc_scratch_bytes(int B, int N, int K) {
  const int TA = max(1, 128 / K);
  const long long nt = ((long long)B * N + TA - 1) / TA;
  const long long ct = nt < CE_CHUNK_TILES ? nt : CE_CHUNK_TILES;
  return (size_t)ct * cebtc::REC_BYTES + 1024;
}

// dpool: d/dpool [B*N,128]; slices: NUM_SMS gradient slices (already zeroed); rec_scratch: records
int ctrl_edge_bwd_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, const float* dpool,
                     int B, int N, int M, int K, double obs_radius, float* slices, float* grad_states, void* rec_scratch,
                     cudaStream_t st) {
  const long long agents = (long long)B * N;
  const int TA = max(1, 128 / K);
  const long long nt = (agents + TA - 1) / TA;
  unsigned char* rec = (unsigned char*)(((uintptr_t)rec_scratch + 1023) & ~(uintptr_t)1023);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_edge_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cebtc::smem_bytes));
  for (long long t0 = 0; t0 < nt; t0 += CE_CHUNK_TILES) {
    const long long t1 = (t0 + CE_CHUNK_TILES < nt) ? t0 + CE_CHUNK_TILES : nt;
    const long long ct = t1 - t0;
    const int grid = (int)(ct < NUM_SMS ? ct : (long long)NUM_SMS);
    ctrl_edge_bwd_tc_kernel<<<grid, cebtc::THREADS, cebtc::smem_bytes, st>>>(states, obstacles, idx, params, dpool, N, M, K, agents,
                                                                           TA, (float)obs_radius, t0, t1, rec, slices,
                                                                           grad_states);
    MACBF_CHECK_LAUNCH("ctrl_edge_bwd_tc_kernel");
    if (int rc = wgrad_tc_ctrl(rec, cebtc::REC_BYTES, ct, slices, grid, st)) return rc;
  }
  return 0;
}

}  // namespace macbf
