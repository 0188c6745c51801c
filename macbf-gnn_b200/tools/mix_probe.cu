// Probe for a cheaper fp32-exact GEMM on tcgen05: D = A * W^T with
//   hi*hi as kind::tf32 (A hi in tensor memory, one 32-bit column per k),
//   lo*hi and hi*lo as kind::f16 MMAs at twice the TF32 rate: a_lo as fp16 and (a_hi * 2^-8) as fp16 in tensor
//   memory (two k per 32-bit column), W_hi as fp16 and (W_lo * 2^8) as fp16 in shared memory (SWIZZLE_128B K-major,
//   64 halves per 128-byte row), all accumulated into ONE fp32 accumulator in TMEM (cross terms first).
// Checks the operand packing (PACK = 0: lower k in the low half of a TMEM column) and the accuracy against fp64.
// Measured on B200: 2.7e-7 .. 4.9e-7 scale-relative, identical to 3 x TF32, with 2/3 of the tensor-core time.  Tried in
// cbf_tc.cu and ctrl_tc.cu (round 2): numerics pass, but both kernels then become bound by their epilogues, which have
// to produce three operand images instead of two (K2 102 -> 105 us, K3a 57 -> 55 us), so it is not in the product yet.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o mix_probe tools/mix_probe.cu && ./mix_probe
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_fp16.h>

#include "../csrc/tc_common.cuh"

using namespace macbf::tc;

__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {   // fp16 x fp16 -> fp32, both K-major
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float lo_k, float hi_k, int pack) {
  const __half2 h = pack == 0 ? __floats2half2_rn(lo_k, hi_k) : __floats2half2_rn(hi_k, lo_k);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// MODE 0: 3 x tf32 (reference path), 1: tf32 hi*hi + fp16 cross terms, 2: all three terms as fp16
template <int N, int K, int MODE>
__global__ void __launch_bounds__(128, 1) mix_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                     float* __restrict__ D, int pack) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* w_hi = reinterpret_cast<float*>(base);                       // tf32 hi  [N x K] 4 B
  float* w_lo = w_hi + N * K;                                         // tf32 lo
  __half* w_hi16 = reinterpret_cast<__half*>(w_lo + N * K);            // fp16 hi  [N x K] 2 B
  __half* w_lo16 = w_hi16 + N * K;                                    // fp16 (lo * 2^8)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  fill_weight_sw128<N, K>(W, K, w_hi, w_lo);
  for (int t = threadIdx.x; t < N * K; t += blockDim.x) {              // fp16 images: slab = 64 k, chunk = 8 halves
    const int n = t / K, k = t % K;
    const float w = W[n * K + k];
    const float hi = tf32_hi(w), lo = w - hi;
    const int off = (k / 64) * (N * 64) + n * 64 + ((((k % 64) / 8) ^ (n & 7)) * 8) + (k % 8);
    w_hi16[off] = __float2half_rn(hi);
    w_lo16[off] = __float2half_rn(lo * 256.f);
  }
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  // TMEM columns: A_hi tf32 [0,K)  A_lo tf32 [K,2K)  A_lo16 [2K, 2K+K/2)  A_hi16s [2K+K/2, 3K)  D [3K, 3K+N)
  const uint32_t cHi = 0, cLo = K, cLo16 = 2 * K, cHi16 = 2 * K + K / 2, cD = 3 * K;
  const int r = threadIdx.x;
  for (int c0 = 0; c0 < K; c0 += 16) {
    float hi[16], lo[16];
    uint32_t lo16[8], hi16[8];
#pragma unroll
    for (int q = 0; q < 16; ++q) tf32_split_trunc(A[r * K + c0 + q], hi[q], lo[q]);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      lo16[q] = pack_h2(lo[2 * q], lo[2 * q + 1], pack);
      hi16[q] = pack_h2(hi[2 * q] * (1.f / 256.f), hi[2 * q + 1] * (1.f / 256.f), pack);
    }
    tmem_st16(tbase + lane_base + cHi + c0, hi);
    tmem_st16(tbase + lane_base + cLo + c0, lo);
    tmem_st8(tbase + lane_base + cLo16 + c0 / 2, reinterpret_cast<const float*>(lo16));
    tmem_st8(tbase + lane_base + cHi16 + c0 / 2, reinterpret_cast<const float*>(hi16));
  }
  wait_st();
  fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_after_sync();
    const uint32_t i32 = idesc_tf32(128, N), i16 = idesc_f16(128, N);
    uint32_t acc = 0;
    if (MODE == 0) {
      for (int term = 0; term < 3; ++term) {
        const uint32_t a_col = (term == 0) ? cLo : cHi;
        const float* b = (term == 1) ? w_lo : w_hi;
        for (int ks = 0; ks < K / 8; ++ks) {
          mma_tf32_ts(tbase + cD, tbase + a_col + ks * 8, smem_desc_sw128(smem_u32(b) + (ks >> 2) * (N * 128) + (ks & 3) * 32), i32, acc);
          acc = 1;
        }
      }
    } else {
      for (int term = 0; term < 2; ++term) {              // a_lo * W_hi, (a_hi 2^-8) * (W_lo 2^8): fp16, K = 16 per MMA
        const uint32_t a_col = (term == 0) ? cLo16 : cHi16;
        const __half* b = (term == 0) ? w_hi16 : w_lo16;
        for (int ks = 0; ks < K / 16; ++ks) {
          mma_f16_ts(tbase + cD, tbase + a_col + ks * 8, smem_desc_sw128(smem_u32(b) + (ks >> 2) * (N * 128) + (ks & 3) * 32), i16, acc);
          acc = 1;
        }
      }
      for (int ks = 0; ks < K / 8; ++ks)                  // hi * hi in tf32, last
        mma_tf32_ts(tbase + cD, tbase + cHi + ks * 8, smem_desc_sw128(smem_u32(w_hi) + (ks >> 2) * (N * 128) + (ks & 3) * 32), i32, 1u);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  for (int c0 = 0; c0 < N; c0 += 16) {
    float v[16];
    tmem_ld16(tbase + lane_base + cD + c0, v);
    wait_ld();
#pragma unroll
    for (int q = 0; q < 16; ++q) D[r * N + c0 + q] = v[q];
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

template <int N, int K, int MODE>
static double run(const char* name, int pack, int dist) {
  std::vector<float> A(128 * K), W(N * K), D(128 * N);
  srand(7);
  for (auto& v : A) {
    const float u = (float)rand() / RAND_MAX;
    if (dist == 0) v = u * 2.f - 1.f;                                  // uniform [-1, 1]
    else v = (rand() & 1) ? 0.f : powf(10.f, u * 5.f - 4.f);          // ReLU-like: zeros and 1e-4 .. 10
  }
  for (auto& v : W) v = ((float)rand() / RAND_MAX * 2.f - 1.f) * 0.125f;
  float *dA, *dW, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dW, W.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, D.size() * 4);
  const size_t smem = (size_t)N * K * (4 + 4 + 2 + 2) + 1024;
  cudaFuncSetAttribute(mix_kernel<N, K, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  mix_kernel<N, K, MODE><<<1, 128, smem>>>(dA, dW, dD, pack);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); return 1e9; }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0, maxref = 0;
  for (int r = 0; r < 128; ++r)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[r * K + k] * (double)W[n * K + k];
      maxerr = fmax(maxerr, fabs(s - (double)D[r * N + n]));
      maxref = fmax(maxref, fabs(s));
    }
  printf("%-28s N=%3d K=%3d pack=%d dist=%d  max|err|=%.3e  max|ref|=%.3e  scale-rel=%.3e\n", name, N, K, pack, dist, maxerr,
         maxref, maxerr / maxref);
  cudaFree(dA); cudaFree(dW); cudaFree(dD);
  return maxerr / maxref;
}

int main() {
  for (int dist = 0; dist < 2; ++dist) {
    run<128, 64, 0>("3xtf32 layer2", 0, dist);
    run<128, 64, 1>("tf32 + 2 fp16 layer2", 0, dist);
    run<128, 64, 1>("tf32 + 2 fp16 layer2", 1, dist);
    run<64, 128, 0>("3xtf32 layer3", 0, dist);
    run<64, 128, 1>("tf32 + 2 fp16 layer3", 0, dist);
    run<64, 128, 1>("tf32 + 2 fp16 layer3", 1, dist);
  }
  return 0;
}
