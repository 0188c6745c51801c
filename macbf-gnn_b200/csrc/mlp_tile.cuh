// FP32 (FFMA) register-tiled layers over a shared-memory tile of R rows (edges or agents).
//
// Activations are CHANNEL-MAJOR in shared memory: A[k * RS + r] (row r contiguous), so a
// thread's 4 consecutive rows are one LDS.128 and a warp's reads are contiguous.  Weights stay
// in the torch layout W[n * WS + k] (Conv1d (out,in,1) / Linear (out,in)), read 4 k's at a time.
// 256 threads: rg = t & 15 picks the rows, cg = t >> 4 the output channels (interleaved by 16
// so the two channel groups of a warp hit different banks).
#pragma once
#include "common.cuh"

namespace macbf {

constexpr int MLP_THREADS = 256;

__device__ __forceinline__ float f4_get(const float4& v, int i) {
  return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w));
}

// C[n][r] = act(bias[n] + sum_k W[n][k] * A[k][r]),   n < NOUT (NOUT in {64,128}), KIN % 4 == 0
template <int R, int KIN, int NOUT, int RS_IN, int RS_OUT, int WS, bool RELU>
__device__ __forceinline__ void layer_nk(const float* __restrict__ A, const float* __restrict__ W,
                                         const float* __restrict__ bias, float* __restrict__ C) {
  static_assert(R == 128 || R == 64, "tile rows");
  static_assert(KIN % 4 == 0 && NOUT % 16 == 0, "shape");
  constexpr int RQ = R / 64;      // float4 row groups per thread (rows rg*4 + 64*i)
  constexpr int CPT = NOUT / 16;  // output channels per thread
  const int rg = threadIdx.x & 15, cg = threadIdx.x >> 4;
  float acc[CPT][RQ * 4];
#pragma unroll
  for (int j = 0; j < CPT; ++j)
#pragma unroll
    for (int i = 0; i < RQ * 4; ++i) acc[j][i] = 0.f;

#pragma unroll 1
  for (int k0 = 0; k0 < KIN; k0 += 4) {
    float4 w[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) w[j] = *reinterpret_cast<const float4*>(W + (cg + 16 * j) * WS + k0);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float a[RQ * 4];
#pragma unroll
      for (int i = 0; i < RQ; ++i) {
        float4 v = *reinterpret_cast<const float4*>(A + (k0 + kk) * RS_IN + rg * 4 + 64 * i);
        a[4 * i + 0] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
      }
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const float wv = f4_get(w[j], kk);
#pragma unroll
        for (int i = 0; i < RQ * 4; ++i) acc[j][i] = fmaf(wv, a[i], acc[j][i]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    const int n = cg + 16 * j;
    const float bv = bias[n];
#pragma unroll
    for (int i = 0; i < RQ; ++i) {
      float4 v;
      v.x = acc[j][4 * i + 0] + bv; v.y = acc[j][4 * i + 1] + bv;
      v.z = acc[j][4 * i + 2] + bv; v.w = acc[j][4 * i + 3] + bv;
      if (RELU) {
        v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
      }
      *reinterpret_cast<float4*>(C + n * RS_OUT + rg * 4 + 64 * i) = v;
    }
  }
}

// Backward data GEMM: D_in[n][r] = gate(G[n][r]) * sum_k W[k][n] * D_out[k][r]
//   (W is the same torch-layout buffer, now read k-major: k = output channel of the layer,
//    n = input channel).  gate = (G > 0) implements relu'(.) with G the saved activation;
//   the result overwrites G in place when C == G.  NOUT = layer input width (64 or 128).
template <int R, int KIN, int NOUT, int RS_IN, int RS_OUT, int WS, bool GATE>
__device__ __forceinline__ void layer_kn(const float* __restrict__ Dout, const float* __restrict__ W,
                                         const float* G, float* C) {
  static_assert(R == 128 || R == 64, "tile rows");
  constexpr int RQ = R / 64;
  constexpr int CPT = NOUT / 16;  // 4 or 8, contiguous channels cg*CPT + j
  static_assert(CPT % 4 == 0, "vector weight loads");
  const int rg = threadIdx.x & 15, cg = threadIdx.x >> 4;
  float acc[CPT][RQ * 4];
#pragma unroll
  for (int j = 0; j < CPT; ++j)
#pragma unroll
    for (int i = 0; i < RQ * 4; ++i) acc[j][i] = 0.f;
#pragma unroll 2
  for (int k = 0; k < KIN; ++k) {
    float w[CPT];
#pragma unroll
    for (int j4 = 0; j4 < CPT / 4; ++j4) {
      float4 v = *reinterpret_cast<const float4*>(W + k * WS + cg * CPT + 4 * j4);
      w[4 * j4 + 0] = v.x; w[4 * j4 + 1] = v.y; w[4 * j4 + 2] = v.z; w[4 * j4 + 3] = v.w;
    }
    float a[RQ * 4];
#pragma unroll
    for (int i = 0; i < RQ; ++i) {
      float4 v = *reinterpret_cast<const float4*>(Dout + k * RS_IN + rg * 4 + 64 * i);
      a[4 * i + 0] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
    }
#pragma unroll
    for (int j = 0; j < CPT; ++j)
#pragma unroll
      for (int i = 0; i < RQ * 4; ++i) acc[j][i] = fmaf(w[j], a[i], acc[j][i]);
  }
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    const int n = cg * CPT + j;
#pragma unroll
    for (int i = 0; i < RQ; ++i) {
      float* p = C + n * RS_OUT + rg * 4 + 64 * i;
      float4 v = make_float4(acc[j][4 * i + 0], acc[j][4 * i + 1], acc[j][4 * i + 2], acc[j][4 * i + 3]);
      if (GATE) {
        float4 g = *reinterpret_cast<const float4*>(G + n * RS_OUT + rg * 4 + 64 * i);
        v.x = g.x > 0.f ? v.x : 0.f; v.y = g.y > 0.f ? v.y : 0.f;
        v.z = g.z > 0.f ? v.z : 0.f; v.w = g.w > 0.f ? v.w : 0.f;
      }
      *reinterpret_cast<float4*>(p) = v;
    }
  }
}

// Weight-gradient GEMM: dW[n][k] += sum_r D[n][r] * A[k][r]  (n < NOUT, k < KVALID <= KIN), both
// operands channel-major in shared memory.  dW is the CTA's private slice of the global partial
// buffer in the flat state_dict layout (row stride WS_OUT = true fan-in): every element is
// owned by exactly one thread, so the read-modify-write needs no atomics and the summation
// order is fixed.  Also db[n] += sum_r D[n][r].  Padded rows must hold D == 0 (they do).
template <int R, int KIN, int NOUT, int RS_D, int RS_A, int WS_OUT, int KVALID>
__device__ __forceinline__ void grad_w(const float* __restrict__ D, const float* __restrict__ A,
                                       float* __restrict__ dW, float* __restrict__ db) {
  constexpr int KB = 16;             // threads along k
  constexpr int NB = 16;             // threads along n
  constexpr int KT = KIN / KB;
  constexpr int NT = NOUT / NB;
  static_assert(KIN % KB == 0 && NOUT % NB == 0, "shape");
  const int kb = threadIdx.x & 15, nb = threadIdx.x >> 4;
  float acc[NT][KT];
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int i = 0; i < KT; ++i) acc[j][i] = 0.f;
  float bsum[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) bsum[j] = 0.f;
#pragma unroll 1
  for (int r = 0; r < R; r += 4) {
    float4 d[NT], a[KT];
#pragma unroll
    for (int j = 0; j < NT; ++j) d[j] = *reinterpret_cast<const float4*>(D + (nb + NB * j) * RS_D + r);
#pragma unroll
    for (int i = 0; i < KT; ++i) a[i] = *reinterpret_cast<const float4*>(A + (kb + KB * i) * RS_A + r);
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      bsum[j] += (d[j].x + d[j].y) + (d[j].z + d[j].w);
#pragma unroll
      for (int i = 0; i < KT; ++i) {
        acc[j][i] = fmaf(d[j].x, a[i].x, acc[j][i]);
        acc[j][i] = fmaf(d[j].y, a[i].y, acc[j][i]);
        acc[j][i] = fmaf(d[j].z, a[i].z, acc[j][i]);
        acc[j][i] = fmaf(d[j].w, a[i].w, acc[j][i]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int n = nb + NB * j;
#pragma unroll
    for (int i = 0; i < KT; ++i) {
      const int k = kb + KB * i;
      if (k < KVALID) dW[n * WS_OUT + k] += acc[j][i];
    }
    if (kb == 0) db[n] += bsum[j];
  }
}

// copy a torch-layout weight matrix [NOUT][KIN] from global into padded shared [NOUT][WS],
// zero-filling k in [KIN, KPAD)
template <int NOUT, int KIN, int KPAD, int WS>
__device__ __forceinline__ void load_weight(const float* __restrict__ g, float* __restrict__ s) {
  for (int t = threadIdx.x; t < NOUT * KPAD; t += blockDim.x) {
    int n = t / KPAD, k = t % KPAD;
    s[n * WS + k] = (k < KIN) ? __ldg(g + n * KIN + k) : 0.f;
  }
}

}  // namespace macbf
