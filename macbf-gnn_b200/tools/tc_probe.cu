// Stand-alone probe for the tcgen05 building blocks of csrc/tc_common.cuh: one 128 x N x K GEMM
// (A from tensor memory, B from SWIZZLE_128B shared memory, 3xTF32 split) against an fp64 CPU
// reference.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_probe tools/tc_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../csrc/tc_common.cuh"

using namespace macbf::tc;

template <int N, int K, int PASSES, bool TRUNC_A = false>
__global__ void __launch_bounds__(128, 1) probe_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                       float* __restrict__ D) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // SWIZZLE_128B atoms must start on a 1024-B boundary of the shared address space
  unsigned char* aligned = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* w_hi = reinterpret_cast<float*>(aligned);
  float* w_lo = w_hi + N * K;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t NCOLS = 512;
  if (warp == 0) tmem_alloc(&tmem_base_s, NCOLS);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  fill_weight_sw128<N, K>(W, K, w_hi, w_lo);
  // make the generic-proxy smem writes visible to the async (tensor core) proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const uint32_t cA_hi = 0, cA_lo = K, cD = 2 * K;
  // A row r -> TMEM lane r
  const int r = threadIdx.x;
  for (int c0 = 0; c0 < K; c0 += 16) {
    float hi[16], lo[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      float v = A[r * K + c0 + q];
      if (TRUNC_A) { hi[q] = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); lo[q] = v - hi[q]; } else { tf32_split(v, hi[q], lo[q]); }
    }
    tmem_st16(tbase + lane_base + cA_hi + c0, hi);
    tmem_st16(tbase + lane_base + cA_lo + c0, lo);
  }
  wait_st();
  fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(128, N);
    uint32_t acc = 0;
    for (int pass = 0; pass < PASSES; ++pass) {
      // small cross terms first: the accumulator is still tiny while they are added, so the tensor
      // core's accumulate rounding costs nothing there; the hi*hi pass goes last
      const int term = (PASSES == 1) ? 2 : pass;       // 0: lo*hi, 1: hi*lo, 2: hi*hi
      const uint32_t a_col = (term == 0) ? cA_lo : cA_hi;
      const float* b = (term == 1) ? w_lo : w_hi;
      for (int ks = 0; ks < K / 8; ++ks) {
        const int kb = ks >> 2, kin = ks & 3;
        const uint32_t baddr = smem_u32(b) + kb * (N * 128) + kin * 32;
        mma_tf32_ts(tbase + cD, tbase + a_col + ks * 8, smem_desc_sw128(baddr), idesc, acc);
        acc = 1;
      }
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
  if (warp == 0) tmem_dealloc(tbase, NCOLS);
  (void)lane;
}

template <int N, int K, int PASSES, bool TRUNC_A = false>
static double run(const char* name) {
  std::vector<float> A(128 * K), W(N * K), D(128 * N);
  srand(7);
  for (auto& v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
  for (auto& v : W) v = ((float)rand() / RAND_MAX * 2.f - 1.f) * 0.125f;
  float *dA, *dW, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dW, W.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, D.size() * 4);
  const size_t smem = 2 * N * K * 4 + 1024;
  cudaFuncSetAttribute(probe_kernel<N, K, PASSES, TRUNC_A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  probe_kernel<N, K, PASSES, TRUNC_A><<<1, 128, smem>>>(dA, dW, dD);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("%s: CUDA error %s\n", name, cudaGetErrorString(e));
    return 1e9;
  }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0, maxref = 0;
  for (int r = 0; r < 128; ++r)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[r * K + k] * (double)W[n * K + k];
      maxerr = fmax(maxerr, fabs(s - (double)D[r * N + n]));
      maxref = fmax(maxref, fabs(s));
    }
  printf("%s: N=%d K=%d passes=%d  max|err|=%.3e  max|ref|=%.3e  scale-rel=%.3e\n", name, N, K, PASSES, maxerr, maxref,
         maxerr / maxref);
  cudaFree(dA); cudaFree(dW); cudaFree(dD);
  return maxerr / maxref;
}

int main() {
  double e1 = run<128, 64, 1>("tf32x1 layer2");
  double e3 = run<128, 64, 3>("tf32x3 layer2");
  double f1 = run<64, 128, 1>("tf32x1 layer3");
  double f3 = run<64, 128, 3>("tf32x3 layer3");
  double g3 = run<128, 32, 3>("tf32x3 k32");
  run<128, 64, 3, true>("tf32x3 layer2 truncA");
  run<64, 128, 3, true>("tf32x3 layer3 truncA");
  bool ok = e1 < 5e-3 && f1 < 5e-3 && e3 < 2e-6 && f3 < 2e-6 && g3 < 2e-6;
  printf("%s\n", ok ? "PROBE OK" : "PROBE FAILED");
  return ok ? 0 : 1;
}
