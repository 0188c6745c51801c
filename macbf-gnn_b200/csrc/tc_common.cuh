// tcgen05 / TMEM primitives (inline PTX, sm_100a) used by the tensor-core edge-MLP kernels.
//
// Numerics: the MLPs must match fp32 within 1e-5, which single-pass TF32 cannot.
// Every GEMM therefore runs as a 3xTF32 split: x = hi + lo with hi = x truncated to the 19 bits
// a kind::tf32 MMA reads (sign, 8 exponent, 10 mantissa) and lo = x - hi (exact in fp32), and
//   A*B ~= A_hi*B_hi + A_lo*B_hi + A_hi*B_lo          (lo*lo ~ 2^-22 relative, dropped)
// accumulated in fp32 in tensor memory.
//
// Operand placement: A (activations, 128 rows = 128 TMEM lanes) lives in TENSOR MEMORY, written
// by the threads that produced it with tcgen05.st (one 32-bit column per k); B (weights, N x K,
// K-major = the torch (out,in) layout) lives in shared memory in the canonical SWIZZLE_128B
// K-major layout: per 32-wide K block a slab of N rows x 128 B, 16-byte chunk c of row n stored
// at chunk position c ^ (n & 7); 8-row groups are 1024 B apart (SBO).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace macbf {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- TMEM allocation (one warp; power-of-two columns >= 32) --------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols));
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- mbarrier -----------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// warp-collective: every lane has finished its TMEM traffic -> one arrival per warp
__device__ __forceinline__ void warp_arrive_after_tmem(uint64_t* bar) {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
// named barrier among `nthreads` threads (id 1..15; id 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// all previously issued tcgen05.mma of this thread arrive on `bar` when they complete
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// ---- descriptors ----------------------------------------------------------------------------
// K-major, SWIZZLE_128B shared-memory operand; `addr` = byte address of the first row at the
// wanted k offset (slab base 1024-B aligned; k advances by adding bytes inside the 128-B row).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);   // start address, 16-B units
  d |= (uint64_t)1 << 16;                  // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;        // stride byte offset: 8 rows x 128 B
  d |= (uint64_t)1 << 46;                  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                  // layout: SWIZZLE_128B
  return d;
}
// kind::tf32, fp32 accumulate, A and B K-major, M x N
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// One lane of a converged warp.  Issue tcgen05.mma / commit under `if (elect_one())` with all
// operands computed warp-uniformly OUTSIDE the branch: the compiler then keeps descriptors and
// TMEM addresses in uniform registers.  Under `if (lane == 0)` it cannot prove uniformity and
// wraps every UTCHMMA in a ~16-instruction R2UR "waterfall" loop (measured: ~100 cycles per MMA,
// which made the single issuing thread the bottleneck of the whole kernel).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
// descriptor of the same operand `bytes` further along (bytes % 16 == 0, no carry out of the
// 14-bit address field for shared-memory windows < 256 KB)
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }

// D[tmem] (+)= A[tmem] * B[smem]^T   (issued by ONE thread)
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T   (both operands K-major SWIZZLE_128B descriptors)
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major operand with 32-byte rows (K = 8 tf32): SWIZZLE_32B, the two 16-byte chunks of row r swapped when
// (r >> 2) & 1; 8-row groups 256 B apart.
__device__ __forceinline__ uint64_t smem_desc_sw32(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(256 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)6 << 61;
  return d;
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld1(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r[0]) : "r"(taddr));
}
// TS-mode MMA with the shared-memory descriptor given as two 32-bit halves: stepping along K only changes the low word
// (the 14-bit address field), so the issuing warp does one 32-bit add per MMA instead of 64-bit descriptor arithmetic.
__device__ __forceinline__ void mma_tf32_ts_split(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 bd;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 bd, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], bd, %4, p;\n\t"
      "}\n" ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One whole 3xTF32 product  D = A * B^T  (A: hi / lo column blocks in tensor memory, B: hi / lo SWIZZLE_128B K-major
// operands whose 32-wide K blocks are `slab_bytes` apart; KS k-steps of 8; small terms first) issued under ONE
// elect_one(), then committed to `bar`.  Call with the whole warp converged.
template <int KS>
__device__ __forceinline__ void mma_3xtf32_ts_commit(uint32_t d_tmem, uint32_t a_hi, uint32_t a_lo, uint64_t b_hi, uint64_t b_lo,
                                                     uint32_t slab_bytes, uint32_t idesc, uint64_t* bar) {
  const uint32_t bhi_lo = (uint32_t)b_hi, blo_lo = (uint32_t)b_lo, b_hi32 = (uint32_t)(b_hi >> 32);
  if (elect_one()) {
#pragma unroll
    for (int term = 0; term < 3; ++term) {           // lo*hi, hi*lo, hi*hi
      const uint32_t a_col = (term == 0) ? a_lo : a_hi;
      const uint32_t bl = (term == 1) ? blo_lo : bhi_lo;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
        mma_tf32_ts_split(d_tmem, a_col + ks * 8, bl + (((ks >> 2) * slab_bytes + (ks & 3) * 32) >> 4), b_hi32, idesc,
                          (term | ks) ? 1u : 0u);
    }
    mma_commit(bar);
  }
  __syncwarp();
}
// generic-proxy shared-memory writes -> visible to the tensor core (async proxy)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM <-> registers: 32 lanes x 32 bit, N consecutive columns per thread -----------------
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}

// ---- 3xTF32 split ---------------------------------------------------------------------------
// Round-to-nearest split (cvt.rna.tf32): |lo| <= 2^-11 |x| with random sign, so the dropped
// lo*lo terms are ~2^-22 and average out; a truncating split (x & 0xFFFFE000) leaves one-sided
// 2^-20 errors that add up coherently (measured 2e-6 vs 3e-7 scale-relative on a K=128 GEMM).
// The result has its 13 low mantissa bits clear, so the tensor core reads it exactly.
// cvt.rna.tf32.f32 expands to ~5 instructions (NaN/Inf handling); for finite inputs the same
// rounding (nearest, ties away from zero) is an integer add of half an ulp and a mask.
__device__ __forceinline__ float tf32_hi(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}
// Cheap split for ACTIVATIONS (2 instructions): hi = x truncated to tf32, lo = x - hi (exact; the
// tensor core reads its top 19 bits).  The weights keep the round-to-nearest split, so the
// dropped a_lo*b_lo products still carry random signs; measured on the probe: 3.1e-7 / 5.1e-7
// scale-relative (K = 64 / 128) against 3.1e-7 / 4.7e-7 with both operands rounded.
__device__ __forceinline__ void tf32_split_trunc(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}
// Packed fp32 pairs (Blackwell FFMA2/FADD2): one issue slot does two lanes of math; the epilogues
// of the tensor-core kernels are issue-bound, so bias adds, the layer-1 FMAs and the lo = x - hi
// subtraction (as hi * -1 + x, exact) go through these.
__device__ __forceinline__ float2 relu2(float2 x) { return make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)); }
__device__ __forceinline__ void split2_trunc(float2 x, float2& hi, float2& lo) {
  hi.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
  hi.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
  lo = __ffma2_rn(hi, make_float2(-1.f, -1.f), x);
}
// The truncating split of 8 values as packed operations: 1.5 instructions per value (tf32_split below takes 5); used
// for the deltas of the backward kernels (activations of the backward pass: same accuracy argument as above).
__device__ __forceinline__ void tf32_split_trunc8(const float* x, float* hi, float* lo) {
#pragma unroll
  for (int e = 0; e < 8; e += 2) {
    float2 h, l;
    split2_trunc(make_float2(x[e], x[e + 1]), h, l);
    hi[e] = h.x; hi[e + 1] = h.y; lo[e] = l.x; lo[e + 1] = l.y;
  }
}
// split x into (hi, lo), both exactly representable in tf32
__device__ __forceinline__ void tf32_split(float x, float& hi, float& lo) {
  hi = tf32_hi(x);
  lo = tf32_hi(x - hi);
}

// Fill one SWIZZLE_128B K-major operand pair (hi, lo) from a torch-layout weight matrix
// W[n * ld + k], n < N, k < K (K % 32 == 0): slab kb holds k in [32 kb, 32 kb + 32).
template <int N, int K>
__device__ __forceinline__ void fill_weight_sw128(const float* __restrict__ W, int ld, float* __restrict__ hi,
                                                  float* __restrict__ lo) {
  constexpr int CHUNKS = N * (K / 4);
  for (int t = threadIdx.x; t < CHUNKS; t += blockDim.x) {
    const int n = t / (K / 4), k4 = t % (K / 4);     // 16-byte chunk k4 of row n
    const int kb = k4 >> 3, c = k4 & 7;
    float w[4], h[4], l[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      w[q] = __ldg(W + (size_t)n * ld + 4 * k4 + q);
      tf32_split(w[q], h[q], l[q]);
    }
    const int off = kb * (N * 32) + n * 32 + ((c ^ (n & 7)) << 2);   // in floats
    *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + off) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

// The same fill in two phases, so that a kernel prologue can put the loads of ALL its weight matrices in flight before the
// first split / store (one memory round trip instead of one per loop iteration and matrix: the prologue of a persistent
// kernel is serial time on every SM).  element (n, k) = W[n * stride_n + k * stride_k]; THREADS = blockDim.x.
template <int N, int K, int THREADS>
struct WeightFill {
  static constexpr int CHUNKS = N * (K / 4);
  static constexpr int IT = (CHUNKS + THREADS - 1) / THREADS;
  float w[IT][4];
  // consecutive threads -> consecutive k4 of a row (row-major weights, stride_k == 1) or consecutive n (NFAST: transposed views)
  // k_valid < K: the matrix has only k_valid columns, the rest of the operand is zero padding
  template <bool NFAST>
  __device__ __forceinline__ void load(const float* __restrict__ W, int stride_n, int stride_k, int k_valid = K) {
#pragma unroll
    for (int i = 0; i < IT; ++i) {
      const int t = (int)threadIdx.x + i * THREADS;
      const int n = NFAST ? t % N : t / (K / 4), k4 = NFAST ? t / N : t % (K / 4);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        w[i][q] = (t < CHUNKS && 4 * k4 + q < k_valid) ? __ldg(W + (size_t)n * stride_n + (size_t)(4 * k4 + q) * stride_k) : 0.f;
    }
  }
  template <bool NFAST>
  __device__ __forceinline__ void store(float* __restrict__ hi, float* __restrict__ lo) const {
#pragma unroll
    for (int i = 0; i < IT; ++i) {
      const int t = (int)threadIdx.x + i * THREADS;
      if (t >= CHUNKS) continue;
      const int n = NFAST ? t % N : t / (K / 4), k4 = NFAST ? t / N : t % (K / 4);
      const int kb = k4 >> 3, c = k4 & 7;
      float h[4], l[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) tf32_split(w[i][q], h[q], l[q]);
      const int off = kb * (N * 32) + n * 32 + ((c ^ (n & 7)) << 2);
      *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<float4*>(lo + off) = make_float4(l[0], l[1], l[2], l[3]);
    }
  }
};

// Same, for a strided view: element (n, k) = W[n * stride_n + k * stride_k]  (a transposed weight for dgrad).
template <int N, int K>
__device__ __forceinline__ void fill_weight_sw128_strided(const float* __restrict__ W, int stride_n, int stride_k,
                                                          float* __restrict__ hi, float* __restrict__ lo) {
  constexpr int CHUNKS = N * (K / 4);
  for (int t = threadIdx.x; t < CHUNKS; t += blockDim.x) {
    const int n = t % N, k4 = t / N;                 // consecutive threads -> consecutive n (coalesced when stride_n == 1)
    const int kb = k4 >> 3, c = k4 & 7;
    float h[4], l[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) tf32_split(__ldg(W + (size_t)n * stride_n + (size_t)(4 * k4 + q) * stride_k), h[q], l[q]);
    const int off = kb * (N * 32) + n * 32 + ((c ^ (n & 7)) << 2);
    *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + off) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

// ---- bf16 hi/lo images for the weight-gradient GEMMs (K = edges) ------------------------------
// x ~= hi + lo with hi = bf16_rn(x), lo = bf16_rn(x - hi): 16 mantissa bits, 2^-17 relative.  Eight
// consecutive channels of one edge row are one 16-byte chunk of an MN-major SWIZZLE_128B operand
// image: block (64 channels) -> 128 rows (edges) x 128 B, chunk c of row e at position c ^ (e & 7).
__device__ __forceinline__ uint32_t bf16x2_rn(float lo_elem, float hi_elem) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi_elem), "f"(lo_elem));   // d = {hi16: first src, lo16: second src}
  return d;
}
__device__ __forceinline__ void bf16_split8(const float* x, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    h[p] = bf16x2_rn(x[2 * p], x[2 * p + 1]);
    const float r0 = x[2 * p] - __uint_as_float(h[p] << 16);
    const float r1 = x[2 * p + 1] - __uint_as_float(h[p] & 0xffff0000u);
    l[p] = bf16x2_rn(r0, r1);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}
// Sixteen consecutive channels (chunks 2*pair, 2*pair+1) of row `row`: one aligned 32-byte sector of the image,
// written with a single 256-bit store (the swizzle only swaps the two 16-byte halves on odd rows).
__device__ __forceinline__ void st_global_v8(void* p, uint4 a, uint4 b) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x),
               "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}
__device__ __forceinline__ void img_store16(unsigned char* img, int row, int pair, uint4 c0, uint4 c1) {
  const uint32_t off = (uint32_t)row * 128u + (uint32_t)((((2 * pair) ^ (row & 7)) & ~1) << 4);
  if (row & 1) st_global_v8(img + off, c1, c0);
  else st_global_v8(img + off, c0, c1);
}

// MN-major SWIZZLE_128B shared-memory operand: 64 elements (bf16) contiguous along M/N per 128-B row, rows = K,
// 8-row groups 1024 B apart (SBO), 64-element blocks `lbo_bytes` apart (LBO).
__device__ __forceinline__ uint64_t smem_desc_mn_sw128(uint32_t addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 with bf16 operands, fp32 accumulate, A and B MN-major
__host__ __device__ constexpr uint32_t idesc_bf16_mn(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ---- bulk global -> shared copies completing on an mbarrier (no tensor map needed) -------------
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

}  // namespace tc
}  // namespace macbf
