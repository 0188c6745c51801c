// Measured denominators for the roofline fractions of the tensor-core kernels: back-to-back tcgen05.mma on every SM
// (one CTA per SM, one issuing thread, operands resident in shared memory / tensor memory, accumulator in TMEM), plus
// an FP32 FFMA loop.  Prints one JSON object; bench.py reads the committed copy (profiles/r2_peak_probe.json).
//
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o peak_probe tools/peak_probe.cu && ./peak_probe
//
// kind::tf32 : K = 8 per instruction (32 B of a 128-B swizzled row), SS mode (A, B from shared memory) and TS mode
//              (A from tensor memory, as the fused MLP kernels use it); kind::f16 with bf16 operands: K = 16.
// The MMA shapes are the ones the kernels issue: M = 128, N = 64 / 128 / 256.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../csrc/tc_common.cuh"

using namespace macbf::tc;

__host__ __device__ constexpr uint32_t idesc_bf16_kmajor(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// MODE 0: tf32 SS, 1: tf32 TS, 2: bf16 SS
template <int MODE, int N>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int groups, int group_len) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* a_s = base;                 // 128 rows x 128 B
  unsigned char* b_s = base + 128 * 128;     // N rows x 128 B
  __shared__ uint64_t bar[2];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  // operands: ones (fp32 1.0 is exact in tf32; bf16 1.0 = 0x3f80)
  const uint32_t one = (MODE == 2) ? 0x3f803f80u : 0x3f800000u;
  for (int i = threadIdx.x; i < (128 + N) * 32; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = one;
  fence_proxy_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  if (MODE == 1) {   // A operand in TMEM columns [256, 288): lane = row, column = k
    float v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = 1.0f;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    tmem_st16(tbase + lane_base + 256, v);
    tmem_st16(tbase + lane_base + 272, v);
    wait_st();
    fence_before_sync();
  }
  __syncthreads();
  if (warp == 0) {
    fence_after_sync();
    const uint32_t idesc = (MODE == 2) ? idesc_bf16_kmajor(128, N) : idesc_tf32(128, N);
    const uint64_t a_desc = smem_desc_sw128(smem_u32(a_s));
    const uint64_t b_desc = smem_desc_sw128(smem_u32(b_s));
    const uint32_t d = tbase, a_t = tbase + 256;
    uint32_t phase[2] = {0, 0};
    for (int g = 0; g < groups; ++g) {
      if (elect_one()) {
        for (int i = 0; i < group_len; ++i) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {    // the four 32-byte k steps of one 128-byte swizzled row
            if (MODE == 0) mma_tf32_ss(d, desc_advance(a_desc, 32 * kk), desc_advance(b_desc, 32 * kk), idesc, 1u);
            else if (MODE == 1) mma_tf32_ts(d, a_t + 8 * kk, desc_advance(b_desc, 32 * kk), idesc, 1u);
            else mma_f16_ss(d, desc_advance(a_desc, 32 * kk), desc_advance(b_desc, 32 * kk), idesc, 1u);
          }
        }
        mma_commit(&bar[g & 1]);
      }
      __syncwarp();
      if (g >= 1) {                           // stay one group ahead of the tensor core
        mbar_wait(&bar[(g - 1) & 1], phase[(g - 1) & 1]);
        phase[(g - 1) & 1] ^= 1;
      }
    }
    mbar_wait(&bar[(groups - 1) & 1], phase[(groups - 1) & 1]);
    fence_after_sync();
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

template <bool PACKED>
__global__ void __launch_bounds__(1024, 1) ffma_rate_kernel(float* out, int iters, float a, float b) {
  float2 acc[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) acc[q] = make_float2(threadIdx.x * 1e-3f + q, q * 0.5f);
  const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if (PACKED) acc[q] = __ffma2_rn(acc[q], a2, b2);
      else { acc[q].x = fmaf(acc[q].x, a, b); acc[q].y = fmaf(acc[q].y, a, b); }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) s += acc[q].x + acc[q].y;
  if (s == 123.456f) out[0] = s;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <typename F>
static float best_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int r = 0; r < reps + 2; ++r) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (r >= 2 && ms < best) best = ms;
  }
  return best;
}

template <int MODE, int N>
static double mma_tflops(int sms) {
  const int groups = 64, group_len = 64;          // 64 x 64 x 4 = 16,384 MMAs per CTA
  const size_t smem = (128 + N) * 128 + 1024;
  CK(cudaFuncSetAttribute(mma_rate_kernel<MODE, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  float ms = best_ms([&] { mma_rate_kernel<MODE, N><<<sms, 128, smem>>>(groups, group_len); }, 5);
  const double k_per = (MODE == 2) ? 16.0 : 8.0;
  const double flop = 2.0 * 128 * N * k_per * 4.0 * groups * group_len * sms;
  return flop / (ms * 1e-3) / 1e12;
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  int clk = 0;
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  float* out;
  CK(cudaMalloc(&out, 4));
  const int iters = 1 << 15;
  float ms_f = best_ms([&] { ffma_rate_kernel<false><<<sms * 2, 1024>>>(out, iters, 1.0001f, 0.5f); }, 5);
  float ms_p = best_ms([&] { ffma_rate_kernel<true><<<sms * 2, 1024>>>(out, iters, 1.0001f, 0.5f); }, 5);
  const double ffma_flop = 2.0 * 16 * (double)iters * 1024 * sms * 2;
  const double tf32_ss64 = mma_tflops<0, 64>(sms), tf32_ss128 = mma_tflops<0, 128>(sms), tf32_ss256 = mma_tflops<0, 256>(sms);
  const double tf32_ts64 = mma_tflops<1, 64>(sms), tf32_ts128 = mma_tflops<1, 128>(sms), tf32_ts256 = mma_tflops<1, 256>(sms);
  const double bf16_64 = mma_tflops<2, 64>(sms), bf16_128 = mma_tflops<2, 128>(sms), bf16_256 = mma_tflops<2, 256>(sms);
  double tf32_best = tf32_ss256 > tf32_ts256 ? tf32_ss256 : tf32_ts256;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_max\": %.0f,\n", prop.name, sms, clk / 1000.0);
  printf(" \"how\": \"one CTA per SM, one thread issuing 16384 tcgen05.mma (M=128, K=8 tf32 / 16 bf16 per instruction) "
         "back to back into one TMEM accumulator, operands resident; best of 5, CUDA events; no clock lock\",\n");
  printf(" \"tf32_ss_tflops\": {\"n64\": %.1f, \"n128\": %.1f, \"n256\": %.1f},\n", tf32_ss64, tf32_ss128, tf32_ss256);
  printf(" \"tf32_ts_tflops\": {\"n64\": %.1f, \"n128\": %.1f, \"n256\": %.1f},\n", tf32_ts64, tf32_ts128, tf32_ts256);
  printf(" \"bf16_ss_tflops\": {\"n64\": %.1f, \"n128\": %.1f, \"n256\": %.1f},\n", bf16_64, bf16_128, bf16_256);
  printf(" \"tf32_mma_tflops\": %.1f, \"bf16_mma_tflops\": %.1f,\n", tf32_best, bf16_256);
  printf(" \"ffma_tflops\": %.2f, \"ffma2_tflops\": %.2f}\n", ffma_flop / (ms_f * 1e-3) / 1e12, ffma_flop / (ms_p * 1e-3) / 1e12);
  return 0;
}
