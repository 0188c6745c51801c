// K3b (tensor-core path) - backward of the controller node net (autograd of controller.py:47-61 + the Euler step
// of train.py:70), given dL/du and dL/ds'.  Three kernels (records in ctrl_node_bwd_layout.cuh, 128 agents per tile):
//   1. ctrl_node_tc_kernel<true> (ctrl_node_tc.cu)  recompute the forward; head backward (gains, action) -> d/d(pre-
//      sigmoid), direct d/d(e, v); dump pool / a1 / a2 as bf16 hi/lo operand images + ReLU masks; dW4, db4.
//   2. ctrl_node_dgrad_tc_kernel (here)  dZ3 = (W4^T d_o) [z3>0] -> dA2 = dZ3 W3 -> dZ2 -> dA1 = dZ2 W2 -> dZ1 ->
//      dA0 = dZ1 W1: three 3xTF32 GEMMs with the deltas in tensor memory and the TRANSPOSED weights in shared memory
//      (W1^T padded to N = 144).  dA0[:, :128] replaces pool by d/dpool for the edge backward; dA0[:, 128:132] plus
//      the direct terms go to d/dstates, d/dgoals.  dZ3, dZ2, dZ1 are dumped as images; the 4 + 1 columns of dW1 / db1
//      that are not a GEMM against pool are reduced with warp shuffles (a CTA sees only a few tiles).
//   3. wgrad_tc_kernel<3> (cbf_bwd_tc.cu)  dW3, dW2, dW1[:, :128], db3, db2 as bf16x3 GEMMs over the agents.
// A CTA handles agents/128/148 tiles, so the phases of a tile run back to back with block barriers, as in the forward.
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "ctrl_node_bwd_layout.cuh"

namespace macbf {

int ctrl_node_tc_dump(const float* states, const float* goals, const float* pool, const float* params, long long agents,
                      double dt, const float* grad_u, const float* grad_s_next, unsigned char* rec, float* slices,
                      cudaStream_t st);
int wgrad_tc_node(const unsigned char* rec, long long ntiles, float* slices, int grid, cudaStream_t st);

namespace cnbtc {
constexpr int THREADS = 512;
constexpr int R = 128;
constexpr int N1 = 144;                       // 132 inputs padded to a legal MMA N (multiple of 16)
// TMEM: [0,64) dZ3_hi [64,128) dZ3_lo [128,256) dA2 -> dZ2_hi [256,384) dZ2_lo [384,448) dA1 -> dZ1_hi [448,512) dZ1_lo;
//       dA0 reuses [0,144) once dA1 has been formed
constexpr uint32_t C_Z3HI = 0, C_Z3LO = 64, C_DA2 = 128, C_Z2LO = 256, C_DA1 = 384, C_Z1LO = 448, C_DA0 = 0;
constexpr int oW3THI = 0;                     // B[n = in (128)][k = out (64)]
constexpr int oW3TLO = oW3THI + 128 * 64;
constexpr int oW2THI = oW3TLO + 128 * 64;     // B[n = in (64)][k = out (128)]
constexpr int oW2TLO = oW2THI + 64 * 128;
constexpr int oW1THI = oW2TLO + 64 * 128;     // B[n = in (144, rows >= 132 zero)][k = out (64)]
constexpr int oW1TLO = oW1THI + N1 * 64;
constexpr int oW4 = oW1TLO + N1 * 64;         // [4][64]
constexpr int oWsum = oW4 + 256;              // [16 warps][80]: per-warp sums of dZ1[c] * (e_x, e_y, v_x, v_y, 1)
constexpr int total_floats = oWsum + 16 * 80;
constexpr size_t smem_bytes = (size_t)total_floats * 4 + 1024;
}  // namespace cnbtc

__global__ void __launch_bounds__(cnbtc::THREADS, 1)
ctrl_node_dgrad_tc_kernel(const float* __restrict__ params, const float* __restrict__ grad_s_next, long long agents_total,
                          float dt, unsigned char* __restrict__ rec, float* __restrict__ pool_dz,
                          float* __restrict__ slices, float* __restrict__ grad_states, float* __restrict__ grad_goals) {
  using namespace cnbtc;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  float* W3Thi = sm + oW3THI; float* W3Tlo = sm + oW3TLO;
  float* W2Thi = sm + oW2THI; float* W2Tlo = sm + oW2TLO;
  float* W1Thi = sm + oW1THI; float* W1Tlo = sm + oW1TLO;
  float* W4 = sm + oW4; float* wsum = sm + oWsum;
  __shared__ uint64_t bar[3];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cgrp = warp >> 2;
  const int r = q * 32 + lane;

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 32) {
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_init(&bar[2], 1);
    fence_mbar_init();
  }
  {   // both transposed weights: all loads in flight, then the splits and stores
    WeightFill<128, 64, THREADS> f3;
    WeightFill<64, 128, THREADS> f2;
    f3.template load<true>(params + CT_NW3, 1, 128);
    f2.template load<true>(params + CT_NW2, 1, 64);
    f3.template store<true>(W3Thi, W3Tlo);
    f2.template store<true>(W2Thi, W2Tlo);
  }
  // W1^T: element (n = input, k = output) = W1[k][n], inputs 132..143 are zero rows
  for (int t = threadIdx.x; t < N1 * 16; t += THREADS) {
    const int n = t % N1, k4 = t / N1;
    const int kb = k4 >> 3, c = k4 & 7;
    float h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) tf32_split(n < 132 ? __ldg(params + CT_NW1 + (size_t)(4 * k4 + e) * 132 + n) : 0.f, h[e], l[e]);
    const int off = kb * (N1 * 32) + n * 32 + ((c ^ (n & 7)) << 2);
    *reinterpret_cast<float4*>(W1Thi + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(W1Tlo + off) = make_float4(l[0], l[1], l[2], l[3]);
  }
  for (int t = threadIdx.x; t < 256; t += THREADS) W4[t] = __ldg(params + CT_NW4 + t);
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
  const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
  const uint64_t d_w3hi = smem_desc_sw128(smem_u32(W3Thi)), d_w3lo = smem_desc_sw128(smem_u32(W3Tlo));
  const uint64_t d_w2hi = smem_desc_sw128(smem_u32(W2Thi)), d_w2lo = smem_desc_sw128(smem_u32(W2Tlo));
  const uint64_t d_w1hi = smem_desc_sw128(smem_u32(W1Thi)), d_w1lo = smem_desc_sw128(smem_u32(W1Tlo));
  constexpr uint32_t IDESC64 = idesc_tf32(128, 64), IDESC128 = idesc_tf32(128, 128), IDESC144 = idesc_tf32(128, N1);
  float accw1 = 0.f;       // thread t < 320 owns dW1[c = t & 63][128 + f] (f = t >> 6 < 4) or db1[c] (f = 4)

  const long long ntiles = (agents_total + R - 1) / R;
  uint32_t par = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, par ^= 1) {
    const long long a0 = tile * R;
    const bool valid = a0 + r < agents_total;
    unsigned char* tr = rec + (size_t)tile * cnrec::BYTES;
    const uint32_t* masks = reinterpret_cast<const uint32_t*>(tr + cnrec::MASKS);
    const uint32_t mb1 = masks[(0 * 4 + cgrp) * R + r], mb2 = masks[(1 * 4 + cgrp) * R + r], mb3 = masks[(2 * 4 + cgrp) * R + r];
    const float4* aux = reinterpret_cast<const float4*>(tr + cnrec::AUX) + 3 * r;
    const float4 d_o = aux[0];
    // ---- dZ3 = (W4^T d_o) * [z3 > 0]: this thread's 16 channels ----
    {
      const int c0 = cgrp * 16;
      uint4 bh[2], bl[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float x[8], hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int c = c0 + half * 8 + e;
          float sv = W4[c] * d_o.x;
          sv = fmaf(W4[64 + c], d_o.y, sv); sv = fmaf(W4[128 + c], d_o.z, sv); sv = fmaf(W4[192 + c], d_o.w, sv);
          x[e] = ((mb3 >> (half * 8 + e)) & 1u) ? sv : 0.f;
        }
        tf32_split_trunc8(x, hi, lo);
        tmem_st8(tlane + C_Z3HI + c0 + half * 8, hi);
        tmem_st8(tlane + C_Z3LO + c0 + half * 8, lo);
        bf16_split8(x, bh[half], bl[half]);
      }
      img_store16(tr + cnrec::Z3_HI, r, cgrp, bh[0], bh[1]);
      img_store16(tr + cnrec::Z3_LO, r, cgrp, bl[0], bl[1]);
      wait_st();
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- dA2 = dZ3 W3 ----
      fence_after_sync();
      mma_3xtf32_ts_commit<8>(tb + C_DA2, tb + C_Z3HI, tb + C_Z3LO, d_w3hi, d_w3lo, (128 * 128), IDESC128, &bar[0]);
    }
    mbar_wait(&bar[0], par);
    fence_after_sync();
#pragma unroll
    for (int half16 = 0; half16 < 2; ++half16) {   // dZ2 = dA2 * [z2 > 0] on this thread's 32 channels
      const int c0 = cgrp * 32 + half16 * 16;
      float v[16];
      tmem_ld16(tlane + C_DA2 + c0, v);
      wait_ld();
      uint4 bh[2], bl[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float x[8], hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          x[e] = ((mb2 >> (half16 * 16 + half * 8 + e)) & 1u) ? v[half * 8 + e] : 0.f;
        }
        tf32_split_trunc8(x, hi, lo);
        tmem_st8(tlane + C_DA2 + c0 + half * 8, hi);
        tmem_st8(tlane + C_Z2LO + c0 + half * 8, lo);
        bf16_split8(x, bh[half], bl[half]);
      }
      img_store16(tr + cnrec::Z2_HI + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + half16, bh[0], bh[1]);
      img_store16(tr + cnrec::Z2_LO + (cgrp >> 1) * cnrec::BLK, r, (cgrp & 1) * 2 + half16, bl[0], bl[1]);
    }
    wait_st();
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- dA1 = dZ2 W2 ----
      fence_after_sync();
      mma_3xtf32_ts_commit<16>(tb + C_DA1, tb + C_DA2, tb + C_Z2LO, d_w2hi, d_w2lo, (64 * 128), IDESC64, &bar[1]);
    }
    mbar_wait(&bar[1], par);
    fence_after_sync();
    {  // dZ1 = dA1 * [z1 > 0] on 16 channels; per-warp sums for the (e, v) columns of dW1 and for db1
      const int c0 = cgrp * 16;
      const float4 ev = aux[2];
      float v[16];
      tmem_ld16(tlane + C_DA1 + c0, v);
      wait_ld();
      uint4 bh[2], bl[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float x[8], hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          x[e] = ((mb1 >> (half * 8 + e)) & 1u) ? v[half * 8 + e] : 0.f;
          const float s0 = warp_sum(x[e] * ev.x), s1 = warp_sum(x[e] * ev.y), s2 = warp_sum(x[e] * ev.z),
                      s3 = warp_sum(x[e] * ev.w), s4 = warp_sum(x[e]);
          if (lane == 0) {
            float* ws = wsum + warp * 80 + (half * 8 + e) * 5;
            ws[0] = s0; ws[1] = s1; ws[2] = s2; ws[3] = s3; ws[4] = s4;
          }
        }
        tf32_split_trunc8(x, hi, lo);
        tmem_st8(tlane + C_DA1 + c0 + half * 8, hi);
        tmem_st8(tlane + C_Z1LO + c0 + half * 8, lo);
        bf16_split8(x, bh[half], bl[half]);
      }
      img_store16(tr + cnrec::Z1_HI, r, cgrp, bh[0], bh[1]);
      img_store16(tr + cnrec::Z1_LO, r, cgrp, bl[0], bl[1]);
      wait_st();
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) {   // ---- dA0 = dZ1 W1 (N = 144) ----
      fence_after_sync();
      mma_3xtf32_ts_commit<8>(tb + C_DA0, tb + C_DA1, tb + C_Z1LO, d_w1hi, d_w1lo, (N1 * 128), IDESC144, &bar[2]);
    }
    if (threadIdx.x < 320) {   // (e, v) columns of dW1 and db1: the 4 lane-quadrant warps of channel group cg, in order
      const int c = threadIdx.x & 63, f = threadIdx.x >> 6, cg = c >> 4, e = c & 15;
      accw1 += (wsum[(cg * 4 + 0) * 80 + e * 5 + f] + wsum[(cg * 4 + 1) * 80 + e * 5 + f]) +
               (wsum[(cg * 4 + 2) * 80 + e * 5 + f] + wsum[(cg * 4 + 3) * 80 + e * 5 + f]);
    }
    mbar_wait(&bar[2], par);
    fence_after_sync();
    {  // d/dpool (replaces pool for the edge backward): this thread's 32 channels
      float* prow = pool_dz + (a0 + r) * 128 + cgrp * 32;
#pragma unroll
      for (int half16 = 0; half16 < 2; ++half16) {
        float v[16];
        tmem_ld16(tlane + C_DA0 + cgrp * 32 + half16 * 16, v);
        wait_ld();
        if (valid) {
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4)
            reinterpret_cast<float4*>(prow + half16 * 16)[j4] = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
        }
      }
      if (cgrp == 3) {   // d/d(p - g), d/dv through layer 1 + the direct terms through the action and the Euler step
        float v[8];
        tmem_ld8(tlane + C_DA0 + 128, v);
        wait_ld();
        if (valid) {
          const float4 dir = aux[1];
          const float gex = dir.x + v[0], gey = dir.y + v[1], gvx = dir.z + v[2], gvy = dir.w + v[3];
          float4 gsn = make_float4(0.f, 0.f, 0.f, 0.f);
          if (grad_s_next != nullptr) gsn = __ldg(reinterpret_cast<const float4*>(grad_s_next) + a0 + r);
          if (grad_states != nullptr) {
            float4* p = reinterpret_cast<float4*>(grad_states) + a0 + r;
            float4 g = *p;
            // e = p - g ; s'_p = p + dt v ; s'_v = v + dt u (u handled in the head backward)
            g.x += gex + gsn.x; g.y += gey + gsn.y;
            g.z += gvx + gsn.z + dt * gsn.x; g.w += gvy + gsn.w + dt * gsn.y;
            *p = g;
          }
          if (grad_goals != nullptr) {
            float2* p = reinterpret_cast<float2*>(grad_goals) + a0 + r;
            float2 g = *p;
            g.x -= gex; g.y -= gey;
            *p = g;
          }
        }
      }
    }
    fence_before_sync();
    __syncthreads();   // TMEM and wsum are rewritten by the next tile
  }
  float* gp = slices + (size_t)blockIdx.x * MACBF_CTRL_PARAMS;
  if (threadIdx.x < 256) gp[CT_NW1 + (threadIdx.x & 63) * 132 + 128 + (threadIdx.x >> 6)] += accw1;
  else if (threadIdx.x < 320) gp[CT_NB1 + (threadIdx.x & 63)] += accw1;
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

// ---- host side ----
namespace {
constexpr long long CN_CHUNK_TILES = 4096;     // 4096 x 300 KB = 1.2 GB of records at a time
inline long long chunk_tiles() {               // MACBF_BWD_CHUNK_TILES=<n>: see cbf_bwd_tc.cu
  const char* e = getenv("MACBF_BWD_CHUNK_TILES");
  const long long v = e ? atoll(e) : 0;
  return (v > 0 && v < CN_CHUNK_TILES) ? v : CN_CHUNK_TILES;
}
}
size_t ctrl_node_bwd_tc_scratch_bytes(int B, int N) {
  const long long nt = ((long long)B * N + 127) / 128;
  return (size_t)(nt < CN_CHUNK_TILES ? nt : CN_CHUNK_TILES) * cnrec::BYTES + 1024;
}

// pool: pooled features in, d/dpool out; slices: gradient slices (zeroed by the caller); scratch: tile records
int ctrl_node_bwd_tc(const float* states, const float* goals, const float* params, const float* grad_u,
                     const float* grad_s_next, int B, int N, double dt, float* pool, float* slices, float* grad_states,
                     float* grad_goals, void* scratch, cudaStream_t st) {
  const long long agents = (long long)B * N;
  unsigned char* rec = (unsigned char*)(((uintptr_t)scratch + 1023) & ~(uintptr_t)1023);
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(ctrl_node_dgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cnbtc::smem_bytes));
  // every per-agent array is contiguous, so a chunk of tiles is just an offset into each of them
  const long long CH = chunk_tiles() * 128;
  for (long long a0 = 0; a0 < agents; a0 += CH) {
    const long long na = (agents - a0 < CH) ? agents - a0 : CH;
    const long long nt = (na + 127) / 128;
    const float* gu = grad_u ? grad_u + a0 * 2 : nullptr;
    const float* gsn = grad_s_next ? grad_s_next + a0 * 4 : nullptr;
    if (int rc = ctrl_node_tc_dump(states + a0 * 4, goals + a0 * 2, pool + a0 * 128, params, na, dt, gu, gsn, rec, slices, st))
      return rc;
    const int grid = (int)(nt < NUM_SMS ? nt : (long long)NUM_SMS);
    ctrl_node_dgrad_tc_kernel<<<grid, cnbtc::THREADS, cnbtc::smem_bytes, st>>>(
        params, gsn, na, (float)dt, rec, pool + a0 * 128, slices, grad_states ? grad_states + a0 * 4 : nullptr,
        grad_goals ? grad_goals + a0 * 2 : nullptr);
    MACBF_CHECK_LAUNCH("ctrl_node_dgrad_tc_kernel");
    if (int rc = wgrad_tc_node(rec, nt, slices, grid, st)) return rc;
  }
  return 0;
}

}  // namespace macbf
