// K2b (tensor-core path) - backward of the fused CBF head (autograd of cbf.py:21-45, REPAIRED).
//
// Three kernels per chunk of tiles (a tile = 128 edge rows; records in cbf_bwd_layout.cuh):
//
//   1. cbf_forward_tc_kernel<true> (cbf_tc.cu)  recompute the forward on the tensor core; dump the
//      layer-2 activations as bf16 hi/lo operand images, the ReLU bit masks, the edge features and
//      gm = dL/dh * radius mask; accumulate dW4, db4.
//   2. cbf_dgrad_tc_kernel   delta chain  dZ3 = gm w4 [z3>0] -> dA2 = dZ3 W3 -> dZ2 = dA2 [z2>0] -> dA1 = dZ2 W2
//      -> dZ1 = dA1 [z1>0]: the two GEMMs are 3xTF32 tcgen05.mma with A (deltas) in tensor memory and
//      B = the TRANSPOSED weights resident in shared memory (same machinery as the forward's layers 2, 3).
//      dZ2 is dumped as bf16 hi/lo images; dW1, db1 and d/dstates (6 features, fp32 atomics) are finished
//      here on the CUDA cores.
//   3. wgrad_cbf_ring_kernel (cbf_wgrad_ring.cu; the generic wgrad_tc_kernel below now serves the controller nets)
//      dW3^T = A2^T dZ3, dW2 = dZ2^T A1, db3, db2: GEMMs whose contraction index is the
//      EDGE, so both operands come from shared memory as MN-major bf16 images: the expensive ones (A2, dZ2) are
//      staged by bulk copies (cp.async.bulk + mbarrier, two 96-KB stage buffers), the cheap ones (a1 from the
//      edge features, dZ3 from gm and the ReLU bits) are rebuilt on chip by generator warps; 3 products (hi*hi,
//      lo*hi, hi*lo) per GEMM with fp32 accumulation in tensor memory across all tiles of the CTA; bias
//      gradients ride along as GEMMs against a constant block of ones.  1 KB of HBM traffic per edge.
//
// Every kernel adds into its CTA's private slice of the gradient partial buffer (flat state_dict layout);
// reduce_partials_kernel adds the slices in order, so parameter gradients are run-to-run deterministic.
// Precision: dgrad chain 3xTF32 (~5e-7); wgrad operands carry 16 mantissa bits (~7e-6 scale-relative on dW2,
// dW3, measured against an fp64 reference on the fixtures) - inside the 5e-5 gradient tolerance of the tests.
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "cbf_bwd_layout.cuh"
#include "ctrl_node_bwd_layout.cuh"

namespace macbf {

int reduce_partials(const float* partial, int nslices, int P, float* out, cudaStream_t st);
int wgrad_cbf_ring(const unsigned char* rec, long long ntiles, float* slices, const float* params, int grid, cudaStream_t st);
int cbf_forward_tc_dump(const float* states, const float* obstacles, const int32_t* idx, const float* params, int N, int M,
                        int K, long long agents, double dist_min, double obs_radius, long long tile_begin,
                        long long tile_end, const float* grad_h, unsigned char* rec, float* slices, cudaStream_t st,
                        float* h, float* a3buf);

// ================================ kernel 2: delta chain ================================
namespace cbfdg {
constexpr int WORKERS = 512;
constexpr int THREADS = WORKERS + 32;
constexpr int THREADS_RECORD = THREADS + 96;   // + three dW4 / db4 warps: 20 warps keep the 96-register budget of 17
constexpr int R = 128;
// TMEM columns: [0,64) dZ3_hi  [64,128) dZ3_lo  [128,256) dA2 -> dZ2_hi (in place)  [256,384) dZ2_lo  [384,448) dA1
constexpr uint32_t C_Z3HI = 0, C_Z3LO = 64, C_DA2 = 128, C_Z2LO = 256, C_DA1 = 384;
constexpr int DZS = 132;                       // row stride of the transposed dZ1 stage: 16-B aligned rows, = 4 (mod 32) words:
                                               // scalar stores down a column and LDS.128 along a row are conflict-free
constexpr int oW3THI = 0;                      // B[n = in (128)][k = out (64)] = W3[k][n]
constexpr int oW3TLO = oW3THI + 128 * 64;
constexpr int oW2THI = oW3TLO + 128 * 64;      // B[n = in (64)][k = out (128)] = W2[k][n]
constexpr int oW2TLO = oW2THI + 64 * 128;
constexpr int oW1 = oW2TLO + 64 * 128;         // [64][8] plain
constexpr int oW4 = oW1 + 64 * 8;
// double-buffered hand-off of a tile's dZ1 to the (deferred) dW1 / d-states pass:
constexpr int oDz1 = oW4 + 64;                 // [2][64][129]  dZ1^T
constexpr int oXs = oDz1 + 2 * 64 * DZS;       // [2][8][128]   per feature: f0..f5, 1 (bias column), d of every row
constexpr int oXpart = oXs + 2 * 128 * 8;      // [2][4][5][128] d/d(f0,f1,f2,f3,f5) partials of the 4 column groups
constexpr int total_floats = oXpart + 2 * 4 * 5 * 128;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cbfdg

template <bool RECORD>
__global__ void __launch_bounds__(RECORD ? cbfdg::THREADS_RECORD : cbfdg::THREADS, 1)
cbf_dgrad_tc_kernel(const int32_t* __restrict__ idx, const float* __restrict__ params, int N, int K,
                    long long agents_total, int TA, long long tile_begin, long long tile_end,
                    unsigned char* __restrict__ rec, float* __restrict__ slices, float* __restrict__ grad_states,
                    const float* __restrict__ grad_h, const float* __restrict__ a3buf) {
  // RECORD: the records come from the recording forward (cbf_forward_tc_kernel<true> without dL/dh): gm = dL/dh *
  // mask is formed here (and written to the record for the wgrad kernel), and dW4 / db4 = sum gm * a3, sum gm are
  // accumulated here from the recorded layer-3 activations.
  using namespace cbfdg;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W3Thi = sm + oW3THI; float* W3Tlo = sm + oW3TLO;
  float* W2Thi = sm + oW2THI; float* W2Tlo = sm + oW2TLO;
  float* W1 = sm + oW1; float* W4 = sm + oW4;
  float* dz1s = sm + oDz1; float* xs = sm + oXs; float* xpart = sm + oXpart;
  // workers -> MMA: z3_full, z2_full[2] (16 arrivals each); MMA -> workers: da2_full[2], da1_full
  // workers -> workers: dz_full[b] (16 arrivals): buffer b of dz1s / xs / xpart is complete
  __shared__ uint64_t z3_full, z2_full[2], da2_full[2], da1_full, dz_full[2];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&z3_full, 16); mbar_init(&z2_full[0], 16); mbar_init(&z2_full[1], 16);
    mbar_init(&da2_full[0], 1); mbar_init(&da2_full[1], 1); mbar_init(&da1_full, 1);
    mbar_init(&dz_full[0], 16); mbar_init(&dz_full[1], 16);
    fence_mbar_init();
  }
  {   // both transposed weights: all loads in flight, then the splits and stores
    constexpr int NT = RECORD ? THREADS_RECORD : THREADS;
    WeightFill<128, 64, NT> f3;
    WeightFill<64, 128, NT> f2;
    f3.template load<true>(params + CBF_W3, 1, 128);
    f2.template load<true>(params + CBF_W2, 1, 64);
    f3.template store<true>(W3Thi, W3Tlo);
    f2.template store<true>(W2Thi, W2Tlo);
  }
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) W1[t] = ((t & 7) < 6) ? __ldg(params + CBF_W1 + (t >> 3) * 6 + (t & 7)) : 0.f;
  for (int t = threadIdx.x; t < 64; t += THREADS) W4[t] = __ldg(params + CBF_W4 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const long long first = tile_begin + blockIdx.x;

  if (warp == WORKERS / 32) {
    // =============================== MMA-issue warp ===============================
    const uint64_t d_w3hi = smem_desc_sw128(smem_u32(W3Thi)), d_w3lo = smem_desc_sw128(smem_u32(W3Tlo));
    const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2Thi)), d_w2lo = smem_desc_sw128(smem_u32(W2Tlo));
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t IDESC = idesc_tf32(128, 64);
    // dA2[128 x 128] = dZ3[128 x 64] * W3 in two 64-column chunks (B rows 64c.. of each 32-wide K block)
    auto issue_da2 = [&](uint32_t parity) {
      mbar_wait(&z3_full, parity);
      fence_after_sync();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
#pragma unroll
        for (int term = 0; term < 3; ++term) {         // lo*hi, hi*lo, hi*hi (small terms first)
          const uint32_t a_col = tb + (term == 0 ? C_Z3LO : C_Z3HI);
          const uint64_t b_desc = desc_advance((term == 1) ? d_w3lo : d_w3hi, c * (64 * 128));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (128 * 128) + (ks & 3) * 32);
            if (elect_one()) mma_tf32_ts(tb + C_DA2 + c * 64, a_col + ks * 8, bd, IDESC, (term | ks) ? 1u : 0u);
          }
        }
        if (elect_one()) mma_commit(&da2_full[c]);
      }
    };
    issue_da2(0);
    uint32_t par = 0;
    for (long long tile = first; tile < tile_end; tile += gridDim.x, par ^= 1) {
      // dA1[128 x 64] = dZ2[128 x 128] * W2: cross terms per 64-wide K chunk as it arrives, hi*hi last
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&z2_full[c], par);
        fence_after_sync();
#pragma unroll
        for (int term = 0; term < 2; ++term) {
          const uint32_t a_col = tb + (term == 0 ? C_Z2LO : C_DA2) + c * 64;
          const uint64_t b_desc = desc_advance((term == 1) ? d_w2lo : d_w2hi, (2 * c) * (64 * 128));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bd = desc_advance(b_desc, (ks >> 2) * (64 * 128) + (ks & 3) * 32);
            if (elect_one()) mma_tf32_ts(tb + C_DA1, a_col + ks * 8, bd, IDESC, (c | term | ks) ? 1u : 0u);
          }
        }
      }
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) {
        const uint64_t bd = desc_advance(d_w2hi, (ks >> 2) * (64 * 128) + (ks & 3) * 32);
        if (elect_one()) mma_tf32_ts(tb + C_DA1, tb + C_DA2 + ks * 8, bd, IDESC, 1u);
      }
      if (elect_one()) mma_commit(&da1_full);
      // dA2 of the NEXT tile runs while the workers finish this one (dZ1, dW1, d/dstates); the tensor core executes
      // in order, so its writes to the dA2 columns follow this tile's reads of dZ2_hi from the same columns.
      if (tile + gridDim.x < tile_end) issue_da2(par ^ 1);
    }
  } else if (RECORD && warp > WORKERS / 32) {
    // ======================= dW4 / db4 warps (records of the recording forward) =======================
    // dW4[c] = sum_e gm_e a3_e[c] (cbf.py:19) straight from the recorded layer-3 activations: three warps stream this
    // CTA's tiles on their own (no barrier shared with the delta chain).  a3buf is [tile][combo = channel / 4][row] float4;
    // warp w owns the combos [cb0, cb0 + ncb), a lane the rows lane + 32 k: every load instruction reads 512 B in a row.
    const int w = warp - WORKERS / 32 - 1;
    const int cb0 = (w == 0) ? 0 : (w == 1 ? 6 : 11), ncb = (w == 0) ? 6 : 5;
    float acc[6][4];
    float accb = 0.f;
#pragma unroll
    for (int c = 0; c < 6; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
    for (long long t = first; t < tile_end; t += gridDim.x) {
      const unsigned char* tr = rec + (size_t)(t - tile_begin) * cbfrec::BYTES;
      float gm[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int r = lane + 32 * k;
        gm[k] = 0.f;
        if (r < TA * K && t * TA + r / K < agents_total)
          gm[k] = __ldg(grad_h + (size_t)(t * TA) * K + r) * __ldg(reinterpret_cast<const float*>(tr + cbfrec::GMASK) + r);
      }
      const float4* src = reinterpret_cast<const float4*>(a3buf) + ((size_t)(t - tile_begin) * 16 + cb0) * R + lane;
      // three batches of two combos: the 8 loads of a batch are all issued before the first of them is used (one
      // memory round trip per batch; a plain loop over the combos made it one round trip per combo, 7 us per tile)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int c = 2 * b + (i >> 2), k = i & 3;
          v[i] = (c < ncb) ? __ldcs(src + c * R + 32 * k) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int c = 2 * b + (i >> 2), k = i & 3;
          acc[c][0] = fmaf(gm[k], v[i].x, acc[c][0]); acc[c][1] = fmaf(gm[k], v[i].y, acc[c][1]);
          acc[c][2] = fmaf(gm[k], v[i].z, acc[c][2]); acc[c][3] = fmaf(gm[k], v[i].w, acc[c][3]);
        }
      }
      if (w == 0) accb += (gm[0] + gm[1]) + (gm[2] + gm[3]);
    }
    float* gp = slices + (size_t)blockIdx.x * MACBF_CBF_PARAMS;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float sw = warp_sum(acc[c][e]);
        if (lane == 0 && c < ncb) gp[CBF_W4 + (cb0 + c) * 4 + e] += sw;
      }
    }
    if (w == 0) {
      const float sb = warp_sum(accb);
      if (lane == 0) gp[CBF_B4] += sb;
    }
  } else {
    // ================================= worker warps =================================
    const int q = warp & 3, cgrp = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    const int c0 = cgrp * 16;
    const int wc = threadIdx.x & 63, wf = threadIdx.x >> 6;     // dW1 / db1 output of this thread: channel wc, column wf (6 = bias)
    float accw1 = 0.f;
    // dZ3 = gm * w4 * [z3 > 0] of tile t (this thread: row r, channels c0 .. c0+15) -> TMEM hi/lo + image
    // Record words are fetched one tile ahead of their use (they come from HBM: ~1 us of latency otherwise
    // sits at the head of every tile).
    struct RecA { uint32_t mb1, mb2a, mb2b; float4 x0, x1; };
    struct RecZ { uint32_t mb3; float gm; };
    auto load_a = [&](long long t) {
      RecA a;
      a.mb1 = a.mb2a = a.mb2b = 0u; a.x0 = make_float4(0.f, 0.f, 0.f, 0.f); a.x1 = a.x0;
      if (t < tile_end) {
        const unsigned char* tr = rec + (size_t)(t - tile_begin) * cbfrec::BYTES;
        const uint32_t* masks = reinterpret_cast<const uint32_t*>(tr + cbfrec::MASKS);
        a.mb1 = masks[(0 * 4 + cgrp) * R + r]; a.mb2a = masks[(1 * 4 + cgrp) * R + r]; a.mb2b = masks[(2 * 4 + cgrp) * R + r];
        if (cgrp == 0) {
          const float4* xr = reinterpret_cast<const float4*>(tr + cbfrec::XREC) + 2 * r;
          a.x0 = xr[0]; a.x1 = xr[1];
        }
      }
      return a;
    };
    auto load_z = [&](long long t) {
      RecZ z;
      z.mb3 = 0u; z.gm = 0.f;
      if (t < tile_end) {
        unsigned char* tr = rec + (size_t)(t - tile_begin) * cbfrec::BYTES;
        z.mb3 = reinterpret_cast<const uint32_t*>(tr + cbfrec::MASKS)[(3 * 4 + cgrp) * R + r];
        if constexpr (!RECORD) {
          z.gm = reinterpret_cast<const float*>(tr + cbfrec::GM)[r];
        } else {
          const long long a = t * TA + r / K;           // rows beyond the tile's agents are padding (mask 0)
          if (r < TA * K && a < agents_total)
            z.gm = __ldg(grad_h + (size_t)(t * TA) * K + r) * reinterpret_cast<const float*>(tr + cbfrec::GMASK)[r];
          if (cgrp == 0) reinterpret_cast<float*>(tr + cbfrec::GM)[r] = z.gm;      // for the wgrad kernel's dZ3
        }
      }
      return z;
    };
    auto build_z3 = [&](long long, const RecZ& z) {
      const uint32_t mb3 = z.mb3;
      const float gm = z.gm;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float x[8], hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          x[e] = ((mb3 >> (half * 8 + e)) & 1u) ? gm * W4[c0 + half * 8 + e] : 0.f;
        }
        tf32_split_trunc8(x, hi, lo);
        tmem_st8(tlane + C_Z3HI + c0 + half * 8, hi);
        tmem_st8(tlane + C_Z3LO + c0 + half * 8, lo);
      }
      wait_st();     // (no dZ3 image: the wgrad kernel rebuilds it from gm and the ReLU bits)
      warp_arrive_after_tmem(&z3_full);
    };
    // Deferred tail of tile ordinal `ord` (buffer ord & 1): dW1 / db1 partial sums and d/dstates.  It runs one tile
    // later, in the shadow of the tensor core's dA1 GEMM, after an ACQUIRE of dz_full (complete long before) - no
    // CTA-wide barrier.  The buffer is rewritten at ordinal ord + 2, whose dA1 needs every warp's z2 arrival of that
    // tile, which each warp issues after this call: write-after-read is ordered by the MMA barriers.
    auto finish = [&](long long t, int ord) {
      const int b = ord & 1;
      mbar_wait(&dz_full[b], (ord >> 1) & 1);
      const float* dzb = dz1s + b * (64 * DZS);
      const float* xsb = xs + b * (128 * 8);
      const float* xpb = xpart + b * (4 * 5 * 128);
      if (wf < 7) {     // dW1[c][f], db1[c] += sum over the tile's rows (rows without an edge have dZ1 = 0)
        const float4* dzr = reinterpret_cast<const float4*>(dzb + wc * DZS);
        const float4* xr = reinterpret_cast<const float4*>(xsb + wf * R);
        float2 s01 = make_float2(0.f, 0.f), s23 = s01;    // packed f32x2 FMAs (same four partial sums)
#pragma unroll 8
        for (int e4 = 0; e4 < R / 4; ++e4) {
          const float4 d = dzr[e4], x = xr[e4];
          s01 = __ffma2_rn(make_float2(d.x, d.y), make_float2(x.x, x.y), s01);
          s23 = __ffma2_rn(make_float2(d.z, d.w), make_float2(x.z, x.w), s23);
        }
        accw1 += (s01.x + s01.y) + (s23.x + s23.y);
      }
      // features -> states (cbf.py:23-30): +grad to agent i, -grad to neighbour j
      if (grad_states != nullptr && cgrp == 0) {
        const long long a0 = t * TA;
        const int rows = (int)min((long long)TA, agents_total - a0) * K;
        if (r < rows) {
          float d0[5];
#pragma unroll
          for (int f = 0; f < 5; ++f)
            d0[f] = (xpb[(0 * 5 + f) * R + r] + xpb[(1 * 5 + f) * R + r]) + (xpb[(2 * 5 + f) * R + r] + xpb[(3 * 5 + f) * R + r]);
          const float inv_d = 1.f / xsb[7 * R + r];
          const float gx = d0[0] + d0[4] * xsb[0 * R + r] * inv_d;
          const float gy = d0[1] + d0[4] * xsb[1 * R + r] * inv_d;
          const long long ga = a0 + r / K;
          const int b_env = (int)(ga / N);
          const int j = __ldg(idx + a0 * K + r);
          const long long gj = (j < N) ? (long long)b_env * N + j : -1;
          if (gj != ga) {   // the self edge contributes +g - g = 0
            float* pi = grad_states + (size_t)ga * 4;
            atomicAdd(pi + 0, gx); atomicAdd(pi + 1, gy); atomicAdd(pi + 2, d0[2]); atomicAdd(pi + 3, d0[3]);
            if (gj >= 0) {
              float* pj = grad_states + (size_t)gj * 4;
              atomicAdd(pj + 0, -gx); atomicAdd(pj + 1, -gy); atomicAdd(pj + 2, -d0[2]); atomicAdd(pj + 3, -d0[3]);
            }
          }
        }
      }
    };
    build_z3(first, load_z(first));
    RecA ra = load_a(first);
    RecZ rz = load_z(first + gridDim.x);
    uint32_t par = 0;
    int it = 0;
    long long tile = first;
    for (; tile < tile_end; tile += gridDim.x, par ^= 1, ++it) {
      unsigned char* tr = rec + (size_t)(tile - tile_begin) * cbfrec::BYTES;
      const uint32_t mb1 = ra.mb1;
      const uint32_t mb2[2] = {ra.mb2a, ra.mb2b};
      const float4 x0 = ra.x0, x1 = ra.x1;
      const int buf = it & 1;
      ra = load_a(tile + gridDim.x);                    // consumed next iteration
      const RecZ rz_next = load_z(tile + 2 * gridDim.x);
      uint4 bh[2], bl[2];
      // ---- dZ2 = dA2 * [z2 > 0] per 64-column chunk: hi in place, lo aside, image dumped ----
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        mbar_wait(&da2_full[c], par);
        fence_after_sync();
        float v[16];
        tmem_ld16(tlane + C_DA2 + c * 64 + c0, v);
        wait_ld();
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float x[8], hi[8], lo[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            x[e] = ((mb2[c] >> (half * 8 + e)) & 1u) ? v[half * 8 + e] : 0.f;
          }
          tf32_split_trunc8(x, hi, lo);
          tmem_st8(tlane + C_DA2 + c * 64 + c0 + half * 8, hi);
          tmem_st8(tlane + C_Z2LO + c * 64 + c0 + half * 8, lo);
          bf16_split8(x, bh[half], bl[half]);
        }
        img_store16(tr + cbfrec::Z2_HI + c * cbfrec::BLK, r, cgrp, bh[0], bh[1]);
        img_store16(tr + cbfrec::Z2_LO + c * cbfrec::BLK, r, cgrp, bl[0], bl[1]);
        wait_st();
        warp_arrive_after_tmem(&z2_full[c]);
      }
      // ---- dZ3 of the next tile (the dZ3 columns are free: both dA2 chunks of this tile have been observed) ----
      if (tile + gridDim.x < tile_end) build_z3(tile + gridDim.x, rz);
      rz = rz_next;
      // ---- deferred tail of the previous tile, while the tensor core finishes this tile's dA1 ----
      if (it > 0) finish(tile - gridDim.x, it - 1);
      // ---- dZ1 = dA1 * [z1 > 0]; d/dfeatures partials; stage dZ1^T for dW1 ----
      mbar_wait(&da1_full, par);
      fence_after_sync();
      {
        float v[16];
        tmem_ld16(tlane + C_DA1 + c0, v);
        wait_ld();
        float2 p01 = make_float2(0.f, 0.f), p23 = p01;  // d/d(f0, f1), d/d(f2, f3) as packed f32x2 FMAs; p4 = d/df5
        float p4 = 0.f;                                 // (the identity indicator f4 has no gradient)
        float* dzb = dz1s + buf * (64 * DZS);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float dz = ((mb1 >> e) & 1u) ? v[e] : 0.f;
          dzb[(c0 + e) * DZS + r] = dz;
          const float4 wa = *reinterpret_cast<const float4*>(W1 + (c0 + e) * 8);
          const float2 dz2 = make_float2(dz, dz);
          p01 = __ffma2_rn(dz2, make_float2(wa.x, wa.y), p01); p23 = __ffma2_rn(dz2, make_float2(wa.z, wa.w), p23);
          p4 = fmaf(dz, W1[(c0 + e) * 8 + 5], p4);
        }
        float* xpb = xpart + buf * (4 * 5 * 128);
#pragma unroll
        for (int f = 0; f < 5; ++f) xpb[(cgrp * 5 + f) * R + r] = f == 0 ? p01.x : f == 1 ? p01.y : f == 2 ? p23.x : f == 3 ? p23.y : p4;
        if (cgrp == 0) {
          float* xsb = xs + buf * (128 * 8);
          xsb[0 * R + r] = x0.x; xsb[1 * R + r] = x0.y; xsb[2 * R + r] = x0.z; xsb[3 * R + r] = x0.w;
          xsb[4 * R + r] = x1.x; xsb[5 * R + r] = x1.y; xsb[6 * R + r] = 1.f; xsb[7 * R + r] = x1.z;
        }
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&dz_full[buf]);
    }
    if (it > 0) finish(tile - gridDim.x, it - 1);
    float* gp = slices + (size_t)blockIdx.x * MACBF_CBF_PARAMS;
    if (wf < 6) gp[CBF_W1 + wc * 6 + wf] += accw1;
    else if (wf == 6) gp[CBF_B1 + wc] += accw1;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

// ================================ kernel 3: weight gradients ================================
namespace cbfwg {
constexpr int THREADS = 192;                   // warp 0: bulk-copy producer, warp 1: MMA issue, warps 2..5: final epilogue
constexpr int MAX_THREADS = 64 + 512;          // with on-chip operand generation: up to 16 generator warps (warps 2..17)
constexpr uint32_t STAGE = cbfrec::STAGE_BYTES;
constexpr uint32_t ONES = 2 * STAGE;           // bf16 ones: 2 blocks x 16 rows x 128 B = 4 KB (A operand M = 128, K = 16)
constexpr size_t smem_bytes = 2 * (size_t)STAGE + 4096 + 1024;
// TMEM columns: [0,64) dW3^T[in][out]  [64,128) dW2[out][in]  [128,192) db3 (every lane)  [192,208) db2 (column 0)
//               [256,320) dW1^T[in][out] (NST = 3)
constexpr uint32_t C_W3 = 0, C_W2 = 64, C_B3 = 128, C_B2 = 192, C_W1 = 256;
}  // namespace cbfwg

// Controller edge net only (NST = 1): with `dpool` set, the A operand - dZ2[row][c] = d/dpool[agent][c] on the row whose
// slot won the max-pool, else 0 - is rebuilt from d/dpool and the winners too; nothing but the edge features is then
// read from the record and the kernel stops being HBM-bound.
struct WgradRoute {
  const float* dpool;        // [agents][128] or NULL
  const uint8_t* argk;       // [agents][128] winning slot, 255 = none
  int K, TA;                 // rows per agent, agents per tile
  long long agents, tile0;   // total agents; global index of the record chunk's first tile
};

// (The CBF head's weight gradients moved to wgrad_cbf_ring_kernel, cbf_wgrad_ring.cu, in round 2; NST = 2 and gen_w4 below
// describe that former instantiation and are no longer launched.)
// NST = 2: CBF (stage 0 = G3: dW3^T, db3; stage 1 = G2: dW2, db2).  NST = 1: controller edge net (one G2-type stage
// per tile: A = dZ2 (128 channels), B = A1 (64 channels) -> dW2[out][in], db2).  NST = 3: controller node net (G3, G2 and
// a third stage A = pool (128 channels), B = dZ1 (64 channels) -> dW1[:, :128]^T, rows of W1 being `ldW1` floats long).
// `rec_stride` bytes between tile records, P floats per gradient slice, off* = offsets of the parameter tensors in a slice.
//
// gen_w1 != NULL (NST = 1, 2): the B operand of the G2-type stage - the layer-1 activations a1 = relu(W1 x + b1), 64
// channels - is not read from the record but REBUILT here from the tile's edge features (float[128][8] at byte
// `gen_xoff` of the record) by the four otherwise idle epilogue warps, written as bf16 hi/lo images straight into the
// stage buffer: a 6 -> 64 layer costs nothing next to an HBM-bound copy, and 256 B per edge of traffic (and the
// producer kernel's work of dumping them) disappear.  gen_w1: W1 [64][gen_nf] row-major, gen_b1: b1 [64].
// gen_w4 != NULL (NST = 2, CBF): likewise the B operand of the G3-type stage, dZ3 = gm * w4 * [z3 > 0], is rebuilt from
// the record's gm and layer-3 ReLU bits (the same product the dgrad kernel feeds its own GEMM).
template <int NST>
__global__ void __launch_bounds__(cbfwg::MAX_THREADS, 1)
wgrad_tc_kernel(const unsigned char* __restrict__ rec, size_t rec_stride, long long ntiles, float* __restrict__ slices, int P,
                int offW3, int offB3, int offW2, int offB2, int offW1, int ldW1, const float* __restrict__ gen_w1,
                const float* __restrict__ gen_b1, int gen_nf, uint32_t gen_xoff, const float* __restrict__ gen_w4,
                WgradRoute route, uint32_t rec_stage) {
  using namespace cbfwg;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[2], empty[2], done;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float gW1[64 * 8];
  __shared__ float gB1[64];
  const bool genA = (NST == 1) && route.dpool != nullptr && gen_w1 != nullptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool gen = (NST != 3) && gen_w1 != nullptr;
  constexpr uint32_t TMEM_COLS = (NST == 3) ? 512 : 256;
  if (warp == 0) tmem_alloc(&tmem_base_s, TMEM_COLS);
  if (threadIdx.x == 32) {
    // a stage is full when its bulk copies have landed and, with `gen`, the four generator warps have passed it
    const int gwarps = (int)(blockDim.x >> 5) - 2;
    mbar_init(&full[0], gen ? 1 + gwarps : 1); mbar_init(&full[1], gen ? 1 + gwarps : 1);
    mbar_init(&empty[0], 1); mbar_init(&empty[1], 1); mbar_init(&done, 1);
    fence_mbar_init();
  }
  if (gen) {
    for (int t = threadIdx.x; t < 64 * 8; t += blockDim.x) gW1[t] = ((t & 7) < gen_nf) ? __ldg(gen_w1 + (t >> 3) * gen_nf + (t & 7)) : 0.f;
    for (int t = threadIdx.x; t < 64; t += blockDim.x) gB1[t] = __ldg(gen_b1 + t);
  }
  for (int t = threadIdx.x; t < 1024; t += blockDim.x) reinterpret_cast<uint32_t*>(sm + ONES)[t] = 0x3f803f80u;   // bf16 1.0 pairs
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  // this CTA's tiles: blockIdx.x, blockIdx.x + grid, ...; two stages (G3, G2) per tile through a 2-deep ring
  if (warp == 0) {
    uint32_t n = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int s = 0; s < NST; ++s, ++n) {
        const uint32_t buf = n & 1;
        if (n >= 2) mbar_wait(&empty[buf], ((n >> 1) - 1) & 1);
        if (elect_one()) {
          const unsigned char* src = rec + (size_t)tile * rec_stride + (size_t)s * rec_stage;   // bytes of a stage in the record
          // with `gen` the last third of a G2-type stage (the a1 images) is produced on chip
          const int parts = genA ? 0 : (gen && ((NST == 1) || s == 1)) ? 2 : 3;
          mbar_expect_tx(&full[buf], parts * (STAGE / 3));
          for (int p = 0; p < parts; ++p) bulk_g2s(sm + buf * STAGE + p * (STAGE / 3), src + p * (STAGE / 3), STAGE / 3, &full[buf]);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t ID64 = idesc_bf16_mn(128, 64), ID16 = idesc_bf16_mn(128, 16);   // M = 128 needs N % 16 == 0
    const uint64_t d_ones_a = smem_desc_mn_sw128(smem_u32(sm + ONES), 2048);     // M = 128: two 64-wide blocks 2 KB apart
    const uint64_t d_ones_b = smem_desc_mn_sw128(smem_u32(sm + ONES), 2048);
    uint32_t n = 0;
    bool first_tile = true;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, first_tile = false) {
      for (int s0 = 0; s0 < NST; ++s0, ++n) {
        const int s = (NST == 1) ? 1 : s0;                // 0: G3-type, 1: G2-type, 2: pool x dZ1 (G3-shaped, no bias)
        const uint32_t buf = n & 1;
        mbar_wait(&full[buf], (n >> 1) & 1);
        fence_after_sync();
        const uint32_t base = smem_u32(sm + buf * STAGE);
        // A: 128 channels (two 64-wide blocks, 16 KB apart), B: 64 channels; hi at +0 / +4 BLK, lo at +2 / +5 BLK
        const uint64_t a_hi = smem_desc_mn_sw128(base, cbfrec::BLK), a_lo = smem_desc_mn_sw128(base + 2 * cbfrec::BLK, cbfrec::BLK);
        const uint64_t b_hi = smem_desc_mn_sw128(base + 4 * cbfrec::BLK, cbfrec::BLK), b_lo = smem_desc_mn_sw128(base + 5 * cbfrec::BLK, cbfrec::BLK);
        const uint32_t dcol = tb + (s == 0 ? C_W3 : (s == 1 ? C_W2 : C_W1));
        const uint32_t acc0 = first_tile ? 0u : 1u;
#pragma unroll
        for (int term = 0; term < 3; ++term) {          // lo*hi, hi*lo, hi*hi
          const uint64_t ad = (term == 0) ? a_lo : a_hi, bd = (term == 1) ? b_lo : b_hi;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)                // 16 edges per MMA: two 8-row groups = 2048 B
            if (elect_one()) mma_f16_ss(dcol, desc_advance(ad, ks * 2048), desc_advance(bd, ks * 2048), ID64, (term | ks) ? 1u : acc0);
        }
        if (s == 0) {
          // db3[n] = sum_e dZ3[e][n]: ones (A, M = 128) x dZ3 (B)
#pragma unroll
          for (int term = 0; term < 2; ++term)
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              if (elect_one()) mma_f16_ss(tb + C_B3, d_ones_a, desc_advance(term == 0 ? b_lo : b_hi, ks * 2048), ID64, (term | ks) ? 1u : acc0);
        } else if (s == 1) {
          // db2[m] = sum_e dZ2[e][m]: dZ2 (A) x ones (B, N = 16)
#pragma unroll
          for (int term = 0; term < 2; ++term)
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              if (elect_one()) mma_f16_ss(tb + C_B2, desc_advance(term == 0 ? a_lo : a_hi, ks * 2048), d_ones_b, ID16, (term | ks) ? 1u : acc0);
        }
        if (elect_one()) mma_commit(&empty[buf]);       // the stage may be overwritten once these MMAs have read it
      }
    }
    if (elect_one()) mma_commit(&done);
  } else {
    if (gen) {
      // ---- generators: operand images built on chip; thread = (row r of the tile, share `part` of the 16-byte chunks) ----
      const int gidx = (int)threadIdx.x - 64, r = gidx & 127, part = gidx >> 7, nparts = ((int)blockDim.x - 64) >> 7;
      uint32_t n = 0;
      float nx_v[8];
      int nx_w[8];
      auto load_route = [&](long long t) {             // d/dpool and winner of (agent (gidx >> 7) + 4 j, channel gidx & 127) of tile t
        const long long a0 = (route.tile0 + t) * route.TA;
        const int ch = gidx & 127;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int a = (gidx >> 7) + 4 * j;
          nx_v[j] = 0.f; nx_w[j] = 255;
          if (t < ntiles && a < route.TA && a0 + a < route.agents) {
            nx_w[j] = __ldg(route.argk + (a0 + a) * 128 + ch);
            nx_v[j] = __ldg(route.dpool + (a0 + a) * 128 + ch);
          }
        }
      };
      // Every global load of a stage is requested right AFTER the previous stage has been published (and, for the first
      // stage, here): fence.proxy.async waits for the thread's outstanding global loads, so a load requested between the
      // start of a stage and its fence puts a DRAM round trip on the hand-off of every stage.
      float4 xa_n = make_float4(0.f, 0.f, 0.f, 0.f), xb_n = xa_n;
      auto load_x = [&](long long t, int s0n) {          // edge features of this thread's row for stage (t, s0n), if it is a G2-type one
        if (t < ntiles && ((NST == 1) || s0n == 1)) {
          const float4* xr = reinterpret_cast<const float4*>(rec + (size_t)t * rec_stride + gen_xoff) + 2 * r;
          xa_n = __ldg(xr); xb_n = __ldg(xr + 1);
        }
      };
      if (genA) load_route(blockIdx.x);
      load_x(blockIdx.x, 0);
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int s0 = 0; s0 < NST; ++s0, ++n) {
          const uint32_t buf = n & 1;
          const bool g2 = (NST == 1) || s0 == 1;
          const float4 xa = xa_n, xb = xb_n;
          // dZ2 is sparse - per agent and channel ONE row (the max-pool winner) carries d/dpool, the rest is zero: the
          // image is zero-filled with wide stores and the winners are scattered into it (512 generator threads:
          // thread = (agent g >> 7 + 4 j of the tile, channel g & 127)); loads go out before the stage is waited for
          float sc_v[8];
          int sc_w[8];                                   // winner slot (255: none / no such agent)
          const int sc_n = genA ? (route.TA + 3) >> 2 : 0;
          if (genA) {   // this tile's values were fetched one tile ago (nx_*)
#pragma unroll
            for (int j = 0; j < 8; ++j) { sc_v[j] = nx_v[j]; sc_w[j] = nx_w[j]; }
          }
          if (n >= 2) mbar_wait(&empty[buf], ((n >> 1) - 1) & 1);
          if (genA) {
            unsigned char* a_hi = sm + buf * STAGE;
            unsigned char* a_lo = sm + buf * STAGE + 2 * cbfrec::BLK;
            uint4* z = reinterpret_cast<uint4*>(a_hi);                  // hi and lo halves are contiguous: 64 KB
#pragma unroll
            for (int i = 0; i < 8; ++i) z[gidx + 512 * i] = make_uint4(0u, 0u, 0u, 0u);
            named_bar_sync(2, 512);                                     // every zero is in place before a winner lands
            const int ch = gidx & 127;
            for (int j = 0; j < sc_n; ++j) {
              int row = -1;
              float val = 0.f;
              if (j < 8) {
#pragma unroll
                for (int jj = 0; jj < 8; ++jj)
                  if (jj == j) { row = (sc_w[jj] < route.K) ? ((gidx >> 7) + 4 * jj) * route.K + sc_w[jj] : -1; val = sc_v[jj]; }
              } else {                                                  // K < 4: more than 32 agents per tile
                const long long a0 = (route.tile0 + tile) * route.TA;
                const int a = (gidx >> 7) + 4 * j;
                if (a < route.TA && a0 + a < route.agents) {
                  const int w = __ldg(route.argk + (a0 + a) * 128 + ch);
                  val = __ldg(route.dpool + (a0 + a) * 128 + ch);
                  row = (w < route.K) ? a * route.K + w : -1;
                }
              }
              if (row < 0) continue;
              const uint32_t h = bf16x2_rn(val, 0.f) & 0xffffu;
              const float lo_f = val - __uint_as_float(h << 16);
              const uint32_t l = bf16x2_rn(lo_f, 0.f) & 0xffffu;
              const uint32_t off = (uint32_t)(ch >> 6) * cbfrec::BLK + (uint32_t)row * 128u +
                                   (uint32_t)((((ch & 63) >> 3) ^ (row & 7)) << 4) + (uint32_t)((ch & 7) << 1);
              *reinterpret_cast<unsigned short*>(a_hi + off) = (unsigned short)h;
              *reinterpret_cast<unsigned short*>(a_lo + off) = (unsigned short)l;
            }
          }
          if (g2) {
            unsigned char* img_hi = sm + buf * STAGE + 4 * cbfrec::BLK;
            unsigned char* img_lo = sm + buf * STAGE + 5 * cbfrec::BLK;
            for (int c8 = part; c8 < 8; c8 += nparts) {
              float x[8];
#pragma unroll
              for (int e = 0; e < 8; e += 2) {           // two channels per packed f32x2 FMA: the forward's chain, same roundings
                const int ch = c8 * 8 + e;
                const float4 wa0 = *reinterpret_cast<const float4*>(gW1 + ch * 8), wa1 = *reinterpret_cast<const float4*>(gW1 + ch * 8 + 8);
                const float4 wb0 = *reinterpret_cast<const float4*>(gW1 + ch * 8 + 4), wb1 = *reinterpret_cast<const float4*>(gW1 + ch * 8 + 12);
                float2 z = make_float2(gB1[ch], gB1[ch + 1]);
                z = __ffma2_rn(make_float2(wa0.x, wa1.x), make_float2(xa.x, xa.x), z);
                z = __ffma2_rn(make_float2(wa0.y, wa1.y), make_float2(xa.y, xa.y), z);
                z = __ffma2_rn(make_float2(wa0.z, wa1.z), make_float2(xa.z, xa.z), z);
                z = __ffma2_rn(make_float2(wa0.w, wa1.w), make_float2(xa.w, xa.w), z);
                z = __ffma2_rn(make_float2(wb0.x, wb1.x), make_float2(xb.x, xb.x), z);
                z = __ffma2_rn(make_float2(wb0.y, wb1.y), make_float2(xb.y, xb.y), z);
                x[e] = fmaxf(z.x, 0.f); x[e + 1] = fmaxf(z.y, 0.f);
              }
              uint4 bh, bl;
              bf16_split8(x, bh, bl);
              const uint32_t off = (uint32_t)r * 128u + (uint32_t)((c8 ^ (r & 7)) << 4);
              *reinterpret_cast<uint4*>(img_hi + off) = bh;
              *reinterpret_cast<uint4*>(img_lo + off) = bl;
            }
            fence_proxy_async_smem();
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&full[buf]);
          // the next stage's inputs
          const bool last = s0 + 1 == NST;
          if (genA && last) load_route(tile + gridDim.x);
          load_x(last ? tile + gridDim.x : tile, last ? 0 : s0 + 1);
        }
      }
    }
    if (warp < 6) {
    // final epilogue (warps 2..5): accumulators -> this CTA's gradient slice
    mbar_wait(&done, 0);
    fence_after_sync();
    const int q = warp & 3;                             // TMEM lane quadrant this warp may read
    const int m = q * 32 + lane;
    const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
    float* gp = slices + (size_t)blockIdx.x * P;
    // slice[base + e * stride] += v[e]: all 16 old values are fetched before the first store (a plain `+=` loop makes
    // every load wait for the previous store - they may alias - and this serial tail then costs tens of microseconds)
    auto accum16 = [&](float* base, int stride, const float* v) {
      float old[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) old[e] = __ldcg(base + e * stride);
#pragma unroll
      for (int e = 0; e < 16; ++e) base[e * stride] = old[e] + v[e];
    };
    for (int c0 = 0; c0 < 64; c0 += 16) {
      float v[16];
      if (NST >= 2) {
        tmem_ld16(tlane + C_W3 + c0, v);                // dW3^T[m = in][n = out]
        wait_ld();
        accum16(gp + offW3 + c0 * 128 + m, 128, v);
      }
      tmem_ld16(tlane + C_W2 + c0, v);                  // dW2[m = out][n = in]
      wait_ld();
      accum16(gp + offW2 + m * 64 + c0, 1, v);
    }
    {
      float v[8];
      tmem_ld8(tlane + C_B2, v);
      wait_ld();
      gp[offB2 + m] += v[0];
    }
    if (NST == 3) {
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16];
        tmem_ld16(tlane + C_W1 + c0, v);                // dW1^T[m = pool channel][n = out]
        wait_ld();
        accum16(gp + offW1 + c0 * ldW1 + m, ldW1, v);
      }
    }
    if (NST >= 2 && q == 0) {                                       // every lane holds the same db3 row; lane m < 32 writes channels 2m, 2m+1
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16];
        tmem_ld16(tlane + C_B3 + c0, v);
        wait_ld();
        if (lane == 0) accum16(gp + offB3 + c0, 1, v);
      }
    }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, TMEM_COLS);
}

// controller node net (ctrl_node_bwd_tc.cu): three 96-KB stages per tile record
int wgrad_tc_node(const unsigned char* rec, long long ntiles, float* slices, int grid, cudaStream_t st) {
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cbfwg::smem_bytes));
  wgrad_tc_kernel<3><<<grid, cbfwg::THREADS, cbfwg::smem_bytes, st>>>(rec, cnrec::BYTES, ntiles, slices, MACBF_CTRL_PARAMS, CT_NW3,
                                                                     CT_NB3, CT_NW2, CT_NB2, CT_NW1, 132, nullptr, nullptr, 0, 0,
                                                                     nullptr, WgradRoute{}, cnrec::STAGE);
  MACBF_CHECK_LAUNCH("wgrad_tc_kernel<node>");
  return 0;
}

// controller edge net (ctrl_edge_bwd_tc.cu): one 96-KB stage per tile record
int wgrad_tc_ctrl(const unsigned char* rec, size_t rec_stride, long long ntiles, float* slices, int grid, cudaStream_t st,
                  const float* params, uint32_t xoff, const float* dpool, const uint8_t* argk, int K, int TA, long long agents,
                  long long tile0) {
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cbfwg::smem_bytes));
  wgrad_tc_kernel<1><<<grid, cbfwg::MAX_THREADS, cbfwg::smem_bytes, st>>>(rec, rec_stride, ntiles, slices, MACBF_CTRL_PARAMS, 0, 0,
                                                                     CT_EW2, CT_EB2, 0, 0, params + CT_EW1, params + CT_EB1, 5, xoff,
                                                                     nullptr, WgradRoute{dpool, argk, K, TA, agents, tile0}, 0);
  MACBF_CHECK_LAUNCH("wgrad_tc_kernel<ctrl>");
  return 0;
}

// ================================ host side ================================
namespace {
constexpr long long CHUNK_TILES = 8192;        // records of one chunk: 8192 x 141 KB = 1.1 GB
// MACBF_BWD_CHUNK_TILES=<n> shrinks the chunk (tests exercise the multi-chunk path on small problems)
inline long long chunk_tiles() {
  const char* e = getenv("MACBF_BWD_CHUNK_TILES");
  const long long v = e ? atoll(e) : 0;
  return (v > 0 && v < CHUNK_TILES) ? v : CHUNK_TILES;
}
inline long long cbf_bwd_tiles(long long agents, int K) {
  const int TA = max(1, 128 / K);
  return (agents + TA - 1) / TA;
}
}  // namespace

size_t cbf_backward_tc_scratch_bytes(int B, int N, int K) {
  const long long nt = cbf_bwd_tiles((long long)B * N, K);
  const long long ct = nt < CHUNK_TILES ? nt : CHUNK_TILES;
  return (size_t)NUM_SMS * MACBF_CBF_PARAMS * sizeof(float) + 1024 + (size_t)ct * cbfrec::BYTES;
}

int cbf_backward_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params,
                    const float* grad_h, int B, int N, int M, int K, double dist_min, double obs_radius,
                    float* grad_params, float* grad_states, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "cbf_backward: B*N must be < 2^31");
  MACBF_CHECK_ARG(K <= 128, "cbf_backward: tensor-core path needs K <= 128");
  const size_t need = cbf_backward_tc_scratch_bytes(B, N, K);
  MACBF_CHECK_ARG(scratch_bytes >= need, "cbf_backward: scratch too small (%zu < %zu)", scratch_bytes, need);
  const size_t slice_bytes = (size_t)NUM_SMS * MACBF_CBF_PARAMS * sizeof(float);
  float* slices = (float*)scratch;
  unsigned char* rec = (unsigned char*)(((uintptr_t)scratch + slice_bytes + 1023) & ~(uintptr_t)1023);
  const int TA = max(1, 128 / K);
  const long long nt = cbf_bwd_tiles(agents, K);
  MACBF_CHECK_CUDA(cudaMemsetAsync(slices, 0, slice_bytes, st));
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_dgrad_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cbfdg::smem_bytes));
  const long long CH = chunk_tiles();
  for (long long t0 = 0; t0 < nt; t0 += CH) {
    const long long t1 = (t0 + CH < nt) ? t0 + CH : nt;
    const long long ct = t1 - t0;
    const int grid = (int)(ct < NUM_SMS ? ct : (long long)NUM_SMS);
    if (int rc = cbf_forward_tc_dump(states, obstacles, idx, params, N, M, K, agents, dist_min, obs_radius, t0, t1, grad_h, rec,
                                     slices, st, nullptr, nullptr))
      return rc;
    cbf_dgrad_tc_kernel<false><<<grid, cbfdg::THREADS, cbfdg::smem_bytes, st>>>(idx, params, N, K, agents, TA, t0, t1, rec, slices,
                                                                               grad_states, nullptr, nullptr);
    MACBF_CHECK_LAUNCH("cbf_dgrad_tc_kernel");
    if (int rc = wgrad_cbf_ring(rec, ct, slices, params, grid, st)) return rc;
  }
  return reduce_partials(slices, NUM_SMS, MACBF_CBF_PARAMS, grad_params, st);
}

// ---- training step: the forward pass writes the records, the backward pass starts at the delta chain ----
// record buffer: [nt tile records of cbfrec::BYTES][nt x cbfrec::A3_BYTES layer-3 activations], nt = all tiles of the
// problem (173 KB per 128 / K agents: the caller decides whether it has the memory; macbf_cbf_backward needs none of it)
size_t cbf_record_bytes(int B, int N, int K) {
  const long long nt = cbf_bwd_tiles((long long)B * N, K);
  return (size_t)nt * (cbfrec::BYTES + cbfrec::A3_BYTES) + 1024;
}
size_t cbf_backward_record_scratch_bytes() { return (size_t)NUM_SMS * MACBF_CBF_PARAMS * sizeof(float); }

static inline unsigned char* rec_base(void* record) { return (unsigned char*)(((uintptr_t)record + 1023) & ~(uintptr_t)1023); }

int cbf_forward_record(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                       int M, int K, double dist_min, double obs_radius, float* h, void* record, size_t record_bytes,
                       cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents * K < (1ll << 31), "cbf_forward_record: B*N*K must be < 2^31");
  MACBF_CHECK_ARG(K <= 128, "cbf_forward_record: K=%d out of range (max 128)", K);
  MACBF_CHECK_ARG(record_bytes >= cbf_record_bytes(B, N, K), "cbf_forward_record: record buffer too small (%zu < %zu)",
                  record_bytes, cbf_record_bytes(B, N, K));
  const long long nt = cbf_bwd_tiles(agents, K);
  unsigned char* rec = rec_base(record);
  float* a3 = (float*)(rec + (size_t)nt * cbfrec::BYTES);
  return cbf_forward_tc_dump(states, obstacles, idx, params, N, M, K, agents, dist_min, obs_radius, 0, nt, nullptr, rec, nullptr,
                             st, h, a3);
}

int cbf_backward_record(const int32_t* idx, const float* params, const float* grad_h, int B, int N, int K, void* record,
                        size_t record_bytes, float* grad_params, float* grad_states, void* scratch, size_t scratch_bytes,
                        cudaStream_t st) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents * K < (1ll << 31) && K <= 128, "cbf_backward_record: B*N*K must be < 2^31 and K <= 128");
  MACBF_CHECK_ARG(record_bytes >= cbf_record_bytes(B, N, K), "cbf_backward_record: record buffer too small (%zu < %zu)",
                  record_bytes, cbf_record_bytes(B, N, K));
  MACBF_CHECK_ARG(scratch_bytes >= cbf_backward_record_scratch_bytes(), "cbf_backward_record: scratch too small (%zu < %zu)",
                  scratch_bytes, cbf_backward_record_scratch_bytes());
  const long long nt = cbf_bwd_tiles(agents, K);
  const int TA = max(1, 128 / K);
  unsigned char* rec = rec_base(record);
  const float* a3 = (const float*)(rec + (size_t)nt * cbfrec::BYTES);
  float* slices = (float*)scratch;
  MACBF_CHECK_CUDA(cudaMemsetAsync(slices, 0, cbf_backward_record_scratch_bytes(), st));
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(cbf_dgrad_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cbfdg::smem_bytes));
  const int grid = (int)(nt < NUM_SMS ? nt : (long long)NUM_SMS);
  cbf_dgrad_tc_kernel<true><<<grid, cbfdg::THREADS_RECORD, cbfdg::smem_bytes, st>>>(idx, params, N, K, agents, TA, 0, nt, rec, slices,
                                                                            grad_states, grad_h, a3);
  MACBF_CHECK_LAUNCH("cbf_dgrad_tc_kernel");
  if (int rc = wgrad_cbf_ring(rec, nt, slices, params, grid, st)) return rc;
  return reduce_partials(slices, NUM_SMS, MACBF_CBF_PARAMS, grad_params, st);
}

}  // namespace macbf
