// K2b-3 - weight gradients of the CBF head (dW3, db3, dW2, db2) from the tile records, as a ring of HALF-TILE stages.
//
// Same arithmetic as wgrad_tc_kernel (cbf_bwd_tc.cu: both operands MN-major bf16 hi/lo images, 3 products per GEMM, fp32
// accumulators in tensor memory across all tiles of the CTA, bias gradients as GEMMs against a block of ones, a1 and dZ3
// rebuilt on chip), but a stage is 64 edge rows instead of 128: four 48-KB stages fit where two 96-KB ones did, so twice
// as many bulk copies are in flight per SM.  With two stages the kernel was paced by the hand-off chain of one buffer
// (MMA done -> copy lands -> generators done -> MMA): 913 MB in 197 us = 4.6 TB/s.
//
//   warp 0      producer: per half-stage four 8-KB bulk copies (hi / lo x two 64-channel blocks) of A2 (stage G3) or dZ2 (G2),
//               and the generators' inputs (gm + ReLU words, or the edge features) into a 2-KB side buffer: a generator
//               warp must not have global loads of its own in flight when it publishes a stage - fence.proxy.async
//               waits for them, which put a DRAM round trip on every half-stage of this serial resource
//   warp 1      MMA issue: per half-stage 3 x 4 MMAs (M = 128 channels, N = 64, K = 16 edges) + 2 x 4 bias MMAs
//   warps 2..17 generators: the B operand of the half-stage - dZ3 = gm w4 [z3 > 0] (G3) or a1 = relu(W1 x + b1) (G2) -
//               as bf16 hi/lo images straight into the stage buffer (thread = one row x one 8-channel chunk, its rows of
//               W1 and w4 in registers); warps 2..5 also run the final epilogue
// Stage order per tile: G3 rows 0-63, G3 rows 64-127, G2 rows 0-63, G2 rows 64-127.
#include "common.cuh"
#include "tc_common.cuh"
#include "cbf_bwd_layout.cuh"

namespace macbf {

namespace cbfwr {
constexpr int GEN_WARPS = 16;
constexpr int THREADS = 64 + 32 * GEN_WARPS;
constexpr int NBUF = 4;
constexpr uint32_t HALF = 64 * 128;                  // one 64-channel block of one image half, 64 rows: 8 KB
constexpr uint32_t STAGE = 6 * HALF;                 // [A hi b0][A hi b1][A lo b0][A lo b1][B hi][B lo] = 48 KB
constexpr uint32_t ONES = NBUF * STAGE;              // bf16 ones: 4 KB
// generator inputs of a half-stage, bulk-copied next to its images: G3: gm[64] + 4 x layer-3 ReLU words[64]; G2: features[64][8]
constexpr uint32_t INPUTS = ONES + 4096, IN_BYTES = 2048;
constexpr size_t smem_bytes = (size_t)NBUF * STAGE + 4096 + NBUF * IN_BYTES + 1024;
// TMEM columns: [0,64) dW3^T[in][out]  [64,128) dW2[out][in]  [128,192) db3 (every lane)  [192,208) db2 (column 0)
constexpr uint32_t C_W3 = 0, C_W2 = 64, C_B3 = 128, C_B2 = 192;
}  // namespace cbfwr

__global__ void __launch_bounds__(cbfwr::THREADS, 1)
wgrad_cbf_ring_kernel(const unsigned char* __restrict__ rec, long long ntiles, float* __restrict__ slices,
                      const float* __restrict__ params) {
  using namespace cbfwr;
  using namespace tc;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[NBUF], empty[NBUF], in_full[NBUF], done;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float gW1[64 * 8];
  __shared__ float gB1[64];
  __shared__ float gW4[64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  if (threadIdx.x == 32) {
    for (int b = 0; b < NBUF; ++b) { mbar_init(&full[b], 1 + GEN_WARPS); mbar_init(&empty[b], 1); mbar_init(&in_full[b], 1); }   // copies + the generator warps
    mbar_init(&done, 1);
    fence_mbar_init();
  }
  for (int t = threadIdx.x; t < 64 * 8; t += THREADS) gW1[t] = ((t & 7) < 6) ? __ldg(params + CBF_W1 + (t >> 3) * 6 + (t & 7)) : 0.f;
  for (int t = threadIdx.x; t < 64; t += THREADS) { gB1[t] = __ldg(params + CBF_B1 + t); gW4[t] = __ldg(params + CBF_W4 + t); }
  for (int t = threadIdx.x; t < 1024; t += THREADS) reinterpret_cast<uint32_t*>(sm + ONES)[t] = 0x3f803f80u;   // bf16 1.0 pairs
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;

  if (warp == 0) {
    // ================================ producer ================================
    uint32_t n = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const unsigned char* tr = rec + (size_t)tile * cbfrec::BYTES;
#pragma unroll 1
      for (int s = 0; s < 4; ++s, ++n) {                  // s >> 1: 0 = G3 (A2 images), 1 = G2 (dZ2 images); s & 1: row half
        const uint32_t buf = n & (NBUF - 1);
        if (n >= NBUF) mbar_wait(&empty[buf], ((n / NBUF) - 1) & 1);
        if (elect_one()) {
          const unsigned char* src = tr + (size_t)(s >> 1) * cbfrec::REC_STAGE + (size_t)(s & 1) * HALF;
          unsigned char* in = sm + INPUTS + buf * IN_BYTES;
          const int h = s & 1;
          if ((s >> 1) == 0) {                            // G3: gm[64] at +0, ReLU words of channel group cg at +256 (1 + cg)
            mbar_expect_tx(&in_full[buf], 5 * 256);
            bulk_g2s(in, tr + cbfrec::GM + h * 256, 256, &in_full[buf]);
#pragma unroll
            for (int cg = 0; cg < 4; ++cg)
              bulk_g2s(in + 256 * (1 + cg), tr + cbfrec::MASKS + (size_t)(3 * 4 + cg) * 512 + h * 256, 256, &in_full[buf]);
          } else {                                        // G2: features float[64][8]
            mbar_expect_tx(&in_full[buf], 2048);
            bulk_g2s(in, tr + cbfrec::XREC + h * 2048, 2048, &in_full[buf]);
          }
          mbar_expect_tx(&full[buf], 4 * HALF);
#pragma unroll
          for (int p = 0; p < 4; ++p) bulk_g2s(sm + buf * STAGE + p * HALF, src + (size_t)p * cbfrec::BLK, HALF, &full[buf]);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issue ================================
    const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
    constexpr uint32_t ID64 = idesc_bf16_mn(128, 64), ID16 = idesc_bf16_mn(128, 16);
    const uint64_t d_ones = smem_desc_mn_sw128(smem_u32(sm + ONES), 2048);      // M = 128: two 64-wide blocks 2 KB apart
    uint32_t n = 0;
    bool first_tile = true;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, first_tile = false) {
#pragma unroll 1
      for (int s = 0; s < 4; ++s, ++n) {
        const uint32_t buf = n & (NBUF - 1);
        mbar_wait(&full[buf], (n / NBUF) & 1);
        fence_after_sync();
        const uint32_t base = smem_u32(sm + buf * STAGE);
        const uint64_t a_hi = smem_desc_mn_sw128(base, HALF), a_lo = smem_desc_mn_sw128(base + 2 * HALF, HALF);
        const uint64_t b_hi = smem_desc_mn_sw128(base + 4 * HALF, HALF), b_lo = smem_desc_mn_sw128(base + 5 * HALF, HALF);
        const bool g3 = (s >> 1) == 0;
        const uint32_t dcol = tb + (g3 ? C_W3 : C_W2);
        const uint32_t acc0 = (first_tile && (s & 1) == 0) ? 0u : 1u;   // the first half-stage of each type starts the accumulators
        if (elect_one()) {
#pragma unroll
          for (int term = 0; term < 3; ++term) {          // lo*hi, hi*lo, hi*hi
            const uint64_t ad = (term == 0) ? a_lo : a_hi, bd = (term == 1) ? b_lo : b_hi;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)                // 16 edges per MMA: two 8-row groups = 2048 B
              mma_f16_ss(dcol, desc_advance(ad, ks * 2048), desc_advance(bd, ks * 2048), ID64, (term | ks) ? 1u : acc0);
          }
          if (g3) {
            // db3[n] = sum_e dZ3[e][n]: ones (A, M = 128) x dZ3 (B)
#pragma unroll
            for (int term = 0; term < 2; ++term)
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                mma_f16_ss(tb + C_B3, d_ones, desc_advance(term == 0 ? b_lo : b_hi, ks * 2048), ID64, (term | ks) ? 1u : acc0);
          } else {
            // db2[m] = sum_e dZ2[e][m]: dZ2 (A) x ones (B, N = 16)
#pragma unroll
            for (int term = 0; term < 2; ++term)
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                mma_f16_ss(tb + C_B2, desc_advance(term == 0 ? a_lo : a_hi, ks * 2048), d_ones, ID16, (term | ks) ? 1u : acc0);
          }
          mma_commit(&empty[buf]);                        // the stage may be overwritten once these MMAs have read it
        }
        __syncwarp();
      }
    }
    if (elect_one()) mma_commit(&done);
    __syncwarp();
  } else {
    // ================================ generators ================================
    // thread = (row rl of the half-stage, 16-byte chunk c8 = 8 channels); its 8 rows of W1 (+ b1) and of w4 live in registers
    const int gidx = (int)threadIdx.x - 64, rl = gidx & 63, c8 = gidx >> 6;
    float w1r[8][7], w4r[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
#pragma unroll
      for (int f = 0; f < 6; ++f) w1r[e][f] = gW1[(c8 * 8 + e) * 8 + f];
      w1r[e][6] = gB1[c8 * 8 + e];
      w4r[e] = gW4[c8 * 8 + e];
    }
    uint32_t n = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll 1
      for (int s = 0; s < 4; ++s, ++n) {
        const uint32_t buf = n & (NBUF - 1);
        const bool g3 = (s >> 1) == 0;
        const unsigned char* in = sm + INPUTS + buf * IN_BYTES;
        if (n >= NBUF) mbar_wait(&empty[buf], ((n / NBUF) - 1) & 1);
        mbar_wait(&in_full[buf], (n / NBUF) & 1);         // landed long ago (the ring is four deep): an acquire
        float x[8];
        if (g3) {
          const float gm = reinterpret_cast<const float*>(in)[rl];
          const uint32_t bits = reinterpret_cast<const uint32_t*>(in + 256 * (1 + (c8 >> 1)))[rl] >> ((c8 & 1) * 8);
#pragma unroll
          for (int e = 0; e < 8; ++e) x[e] = ((bits >> e) & 1u) ? gm * w4r[e] : 0.f;
        } else {
          const float4 xa = reinterpret_cast<const float4*>(in)[2 * rl], xb = reinterpret_cast<const float4*>(in)[2 * rl + 1];
          const float xf[6] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y};
#pragma unroll
          for (int e = 0; e < 8; e += 2) {                // two channels per packed f32x2 FMA: the forward's chain, same roundings
            float2 z = make_float2(w1r[e][6], w1r[e + 1][6]);
#pragma unroll
            for (int f = 0; f < 6; ++f) z = __ffma2_rn(make_float2(w1r[e][f], w1r[e + 1][f]), make_float2(xf[f], xf[f]), z);
            x[e] = fmaxf(z.x, 0.f); x[e + 1] = fmaxf(z.y, 0.f);
          }
        }
        uint4 bh, bl;
        bf16_split8(x, bh, bl);
        const uint32_t off = (uint32_t)rl * 128u + (uint32_t)((c8 ^ (rl & 7)) << 4);
        *reinterpret_cast<uint4*>(sm + buf * STAGE + 4 * HALF + off) = bh;
        *reinterpret_cast<uint4*>(sm + buf * STAGE + 5 * HALF + off) = bl;
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[buf]);
      }
    }
    if (warp < 6) {
      // final epilogue (warps 2..5): accumulators -> this CTA's gradient slice
      mbar_wait(&done, 0);
      fence_after_sync();
      const int q = warp & 3;                             // TMEM lane quadrant this warp may read
      const int m = q * 32 + lane;
      const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
      float* gp = slices + (size_t)blockIdx.x * MACBF_CBF_PARAMS;
      auto accum16 = [&](float* base, int stride, const float* v) {   // all old values are fetched before the first store
        float old[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) old[e] = __ldcg(base + e * stride);
#pragma unroll
        for (int e = 0; e < 16; ++e) base[e * stride] = old[e] + v[e];
      };
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16];
        tmem_ld16(tlane + C_W3 + c0, v);                  // dW3^T[m = in][n = out]
        wait_ld();
        accum16(gp + CBF_W3 + c0 * 128 + m, 128, v);
        tmem_ld16(tlane + C_W2 + c0, v);                  // dW2[m = out][n = in]
        wait_ld();
        accum16(gp + CBF_W2 + m * 64 + c0, 1, v);
      }
      {
        float v[8];
        tmem_ld8(tlane + C_B2, v);
        wait_ld();
        gp[CBF_B2 + m] += v[0];
      }
      if (q == 0) {                                       // every lane holds the same db3 row
        for (int c0 = 0; c0 < 64; c0 += 16) {
          float v[16];
          tmem_ld16(tlane + C_B3 + c0, v);
          wait_ld();
          if (lane == 0) accum16(gp + CBF_B3 + c0, 1, v);
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

int wgrad_cbf_ring(const unsigned char* rec, long long ntiles, float* slices, const float* params, int grid, cudaStream_t st) {
  MACBF_CHECK_CUDA(cudaFuncSetAttribute(wgrad_cbf_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cbfwr::smem_bytes));
  wgrad_cbf_ring_kernel<<<grid, cbfwr::THREADS, cbfwr::smem_bytes, st>>>(rec, ntiles, slices, params);
  MACBF_CHECK_LAUNCH("wgrad_cbf_ring_kernel");
  return 0;
}

}  // namespace macbf
