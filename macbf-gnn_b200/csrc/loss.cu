// K5 - time-to-collision mask (core.py:187-209) and the barrier / derivative / action loss
// reductions (core.py:89-184), plus the analytic seeds of the backward pass.
//
// Reduction is two-stage and deterministic: each CTA reduces a fixed chunk of one env's
// (agent, slot) rows to 10 partial sums in caller-provided scratch, a second single-CTA-per-env
// kernel adds the partials in index order and forms the 12 scalars.  HBM-bound: per edge it
// reads idx (4 B), h and h' (8 B) and writes the mask (1 B).
#include "common.cuh"

namespace macbf {

constexpr int LOSS_THREADS = 256;
constexpr int LOSS_ROWS_PER_CTA = 1024;
constexpr int LOSS_NACC = 10;
static_assert(LOSS_ROWS_PER_CTA % LOSS_THREADS == 0, "whole rows per thread");

static inline int loss_chunks(int N, int K) {
  long long rows = (long long)N * K;
  return (int)((rows + LOSS_ROWS_PER_CTA - 1) / LOSS_ROWS_PER_CTA);
}

__global__ void ttc_mask_kernel(const float* __restrict__ states, const float* __restrict__ obstacles,
                                const int32_t* __restrict__ idx, int N, int M, int K, long long edges,
                                float r2, float ttc, uint8_t* __restrict__ mask) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= edges) return;
  const long long ga = e / K;
  const int b = (int)(ga / N), i = (int)(ga % N);
  const int j = __ldg(idx + e);
  const float4 si = __ldg(reinterpret_cast<const float4*>(states) + ga);
  const float4 sj = load_candidate(states, obstacles, b, j, N, M);
  mask[e] = ttc_dangerous(edge_geom(si, sj, i, j), r2, ttc) ? 1 : 0;
}

// acc: 0 sum_dang relu(h+1e-3) | 1 sum_safe relu(-h) | 2 cnt_dang[h<=0] | 3 cnt_safe[h>0]
//      4 sum_dang relu(-d+1e-3) | 5 sum_safe relu(-d) | 6 cnt_dang[d>=0] | 7 cnt_safe[d>0]
//      8 n_dang | 9 sum_i | |a|^2 - |a_ref|^2 |
__global__ void __launch_bounds__(LOSS_THREADS)
loss_partial_kernel(const float* __restrict__ states, const float* __restrict__ goals,
                    const float* __restrict__ obstacles, const int32_t* __restrict__ idx,
                    const float* __restrict__ h, const float* __restrict__ h_next, const float* __restrict__ u,
                    int N, int M, int K, float r2, float ttc, float dt_alpha, uint8_t* __restrict__ mask,
                    float* __restrict__ partial, int chunk0, int nchunks, int agent_begin, int agent_end) {
  // the launched chunks are [chunk0, chunk0 + gridDim.x) of the env's nchunks; the action term covers the agents
  // [agent_begin, agent_end), striped over the launched chunks (the whole env unless a row range is set)
  const int b = blockIdx.y, chunk = chunk0 + blockIdx.x;
  const long long rows_env = (long long)N * K;
  const long long r_begin = (long long)chunk * LOSS_ROWS_PER_CTA;
  const long long r_end = min(rows_env, r_begin + LOSS_ROWS_PER_CTA);
  float acc[LOSS_NACC];
#pragma unroll
  for (int q = 0; q < LOSS_NACC; ++q) acc[q] = 0.f;

  // a thread's LOSS_ROWS_PER_CTA / LOSS_THREADS rows: all neighbour indices first, then all endpoint states, then the
  // arithmetic - three rounds of dependent loads per CTA instead of three per row
  constexpr int RPT = LOSS_ROWS_PER_CTA / LOSS_THREADS;
  int jj[RPT];
  float4 si_[RPT], sj_[RPT];
  float hv_[RPT], hn_[RPT];
#pragma unroll
  for (int q = 0; q < RPT; ++q) {
    const long long r = r_begin + threadIdx.x + q * LOSS_THREADS;
    jj[q] = -1; hv_[q] = 0.f; hn_[q] = 0.f;
    if (r < r_end) {
      const long long e = (long long)b * rows_env + r;
      jj[q] = __ldg(idx + e);
      hv_[q] = __ldg(h + e);
      if (h_next != nullptr) hn_[q] = __ldg(h_next + e);
      si_[q] = __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + (int)(r / K));
    }
  }
#pragma unroll
  for (int q = 0; q < RPT; ++q)
    if (jj[q] >= 0) sj_[q] = load_candidate(states, obstacles, b, jj[q], N, M);
#pragma unroll
  for (int q = 0; q < RPT; ++q) {
    const long long r = r_begin + threadIdx.x + q * LOSS_THREADS;
    if (r >= r_end) continue;
    const int i = (int)(r / K);
    const long long e = (long long)b * rows_env + r;
    const int j = jj[q];
    const bool dang = ttc_dangerous(edge_geom(si_[q], sj_[q], i, j), r2, ttc);
    mask[e] = dang ? 1 : 0;
    const float hv = hv_[q];
    if (dang) {
      acc[8] += 1.f;
      acc[0] += fmaxf(xadd(hv, 1e-3f), 0.f);   // core.py:114-115
      acc[2] += (hv <= 0.f) ? 1.f : 0.f;       // core.py:119-120
    } else {
      acc[1] += fmaxf(-hv, 0.f);               // core.py:116-117
      acc[3] += (hv > 0.f) ? 1.f : 0.f;        // core.py:121-122 (repair L6)
    }
    if (h_next != nullptr) {
      // core.py:139: h_next - h + TIME_STEP * ALPHA_CBF * h
      const float d = xadd(xsub(hn_[q], hv), xmul(dt_alpha, hv));
      if (dang) {
        acc[4] += fmaxf(xadd(-d, 1e-3f), 0.f);
        acc[6] += (d >= 0.f) ? 1.f : 0.f;
      } else {
        acc[5] += fmaxf(-d, 0.f);
        acc[7] += (d > 0.f) ? 1.f : 0.f;
      }
    }
  }
  if (u != nullptr) {
    // core.py:174-184: a_ref = -(p - g) - sqrt(3) v, agents striped over the env's chunks
    const float s3 = 1.7320508075688772f;
    for (int i = agent_begin + blockIdx.x * LOSS_THREADS + threadIdx.x; i < agent_end; i += gridDim.x * LOSS_THREADS) {
      const float4 s = __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + i);
      const float2 g = __ldg(reinterpret_cast<const float2*>(goals) + (size_t)b * N + i);
      const float2 a = __ldg(reinterpret_cast<const float2*>(u) + (size_t)b * N + i);
      const float rx = -(s.x - g.x) - s3 * s.z, ry = -(s.y - g.y) - s3 * s.w;
      acc[9] += fabsf((a.x * a.x + a.y * a.y) - (rx * rx + ry * ry));
    }
  }
  __shared__ float red[LOSS_THREADS / 32][LOSS_NACC];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < LOSS_NACC; ++q) {
    float v = warp_sum(acc[q]);
    if (lane == 0) red[warp][q] = v;
  }
  __syncthreads();
  if (threadIdx.x < LOSS_NACC) {
    float v = 0.f;
    for (int w = 0; w < LOSS_THREADS / 32; ++w) v += red[w][threadIdx.x];
    partial[((size_t)b * nchunks + chunk) * LOSS_NACC + threadIdx.x] = v;
  }
}

__global__ void loss_final_kernel(const float* __restrict__ partial, int nchunks, int N, int K,
                                  int has_next, int has_u, float* __restrict__ scalars) {
  const int b = blockIdx.x;
  __shared__ double acc[LOSS_NACC];
  {  // warp q sums accumulator q: lanes stride over the chunks, then a fixed shuffle tree (deterministic)
    const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double v = 0.0;
    for (int c = lane; c < nchunks; c += 32) v += (double)partial[((size_t)b * nchunks + c) * LOSS_NACC + q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) acc[q] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float nd = (float)acc[8];
    const float ns = (float)((double)N * K - acc[8]);
    const float dd = 1e-5f + nd, ds = 1e-5f + ns;   // core.py:114-117 denominators
    float* o = scalars + (size_t)b * MACBF_NUM_SCALARS;
    o[0] = (float)acc[0] / dd;
    o[1] = (float)acc[1] / ds;
    o[2] = nd > 0.f ? (float)acc[2] / dd : -1.f;     // core.py:124-127 (repair L7)
    o[3] = ns > 0.f ? (float)acc[3] / ds : -1.f;
    o[4] = has_next ? (float)acc[4] / dd : 0.f;
    o[5] = has_next ? (float)acc[5] / ds : 0.f;
    o[6] = has_next ? (nd > 0.f ? (float)acc[6] / dd : -1.f) : 0.f;
    o[7] = has_next ? (ns > 0.f ? (float)acc[7] / ds : -1.f) : 0.f;
    o[8] = has_u ? (float)(acc[9] / (double)N) : 0.f;
    o[9] = 10.f * (2.f * o[0] + o[1] + 2.f * o[4] + o[5] + 0.01f * o[8]);  // train.py:93,98
    o[10] = nd;
    o[11] = ns;
  }
}

// dL/dh, dL/dh_next, dL/du from d(objective)/d(scalars).  Hinge subgradients follow torch:
// clamp(min=0) passes the gradient where its argument >= 0, abs' = sign.
__global__ void loss_backward_kernel(const float* __restrict__ states, const float* __restrict__ goals,
                                     const float* __restrict__ h, const float* __restrict__ h_next,
                                     const float* __restrict__ u, const uint8_t* __restrict__ mask,
                                     const float* __restrict__ scalars, const float* __restrict__ gs, int N,
                                     int K, float dt_alpha, float* __restrict__ gh, float* __restrict__ ghn,
                                     float* __restrict__ gu, float* __restrict__ gst) {
  const int b = blockIdx.y;
  const long long rows_env = (long long)N * K;
  const float* sc = scalars + (size_t)b * MACBF_NUM_SCALARS;
  const float* g = gs + (size_t)b * MACBF_NUM_SCALARS;
  const float inv_d = 1.f / (1e-5f + sc[10]), inv_s = 1.f / (1e-5f + sc[11]);
  const float g0 = g[0] * inv_d, g1 = g[1] * inv_s, g4 = g[4] * inv_d, g5 = g[5] * inv_s, g8 = g[8] / (float)N;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows_env; r += stride) {
    const long long e = (long long)b * rows_env + r;
    const bool dang = mask[e] != 0;
    const float hv = h[e];
    float dh = 0.f;
    if (dang) dh += (xadd(hv, 1e-3f) >= 0.f) ? g0 : 0.f;
    else      dh += (-hv >= 0.f) ? -g1 : 0.f;
    if (h_next != nullptr) {
      const float d = xadd(xsub(h_next[e], hv), xmul(dt_alpha, hv));
      float gd;
      if (dang) gd = (xadd(-d, 1e-3f) >= 0.f) ? -g4 : 0.f;
      else      gd = (-d >= 0.f) ? -g5 : 0.f;
      if (ghn != nullptr) ghn[e] = gd;
      dh += gd * (dt_alpha - 1.f);
    } else if (ghn != nullptr) {
      ghn[e] = 0.f;
    }
    if (gh != nullptr) gh[e] = dh;
  }
  if (gu != nullptr || gst != nullptr) {
    const float s3 = 1.7320508075688772f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
      float2 o = make_float2(0.f, 0.f);
      float4 os = make_float4(0.f, 0.f, 0.f, 0.f);
      if (u != nullptr) {
        const float4 s = __ldg(reinterpret_cast<const float4*>(states) + (size_t)b * N + i);
        const float2 gl = __ldg(reinterpret_cast<const float2*>(goals) + (size_t)b * N + i);
        const float2 a = __ldg(reinterpret_cast<const float2*>(u) + (size_t)b * N + i);
        const float rx = -(s.x - gl.x) - s3 * s.z, ry = -(s.y - gl.y) - s3 * s.w;
        const float diff = (a.x * a.x + a.y * a.y) - (rx * rx + ry * ry);
        const float sg = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
        o = make_float2(g8 * sg * 2.f * a.x, g8 * sg * 2.f * a.y);
        // d|a_ref|^2/dp = -2 a_ref, d/dv = -2 sqrt(3) a_ref; the loss carries -|a_ref|^2
        os = make_float4(g8 * sg * 2.f * rx, g8 * sg * 2.f * ry, g8 * sg * 2.f * s3 * rx, g8 * sg * 2.f * s3 * ry);
      }
      if (gu != nullptr) reinterpret_cast<float2*>(gu)[(size_t)b * N + i] = o;
      if (gst != nullptr) reinterpret_cast<float4*>(gst)[(size_t)b * N + i] = os;
    }
  }
}

int ttc_mask(const float* states, const float* obstacles, const int32_t* idx, int B, int N, int M, int K,
             double dist_min, double ttc, uint8_t* mask, cudaStream_t st) {
  const long long edges = (long long)B * N * K;
  const float r2 = (float)(dist_min * dist_min);  // python: DIST_MIN_THRES ** 2 in double, then cast
  ttc_mask_kernel<<<(unsigned)((edges + 255) / 256), 256, 0, st>>>(states, obstacles, idx, N, M, K, edges, r2,
                                                                  (float)ttc, mask);
  MACBF_CHECK_LAUNCH("ttc_mask_kernel");
  return 0;
}

size_t loss_scratch_bytes(int B, int N, int K) {
  return (size_t)B * loss_chunks(N, K) * LOSS_NACC * sizeof(float);
}

// Argument checks shared by macbf_loss_forward and macbf_loss_forward_noalloc: loss_partial_kernel reads states as
// float4 and goals / u as float2, so a misaligned buffer must come back as an error code, not as a sticky fault.
int loss_validate(const char* fn, const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                  const float* h, const float* u, int B, int N, int M, int K, const uint8_t* mask,
                  const float* scalars) {
  MACBF_CHECK_ARG(B >= 1 && B <= 65535, "%s: B=%d out of range [1,65535]", fn, B);
  MACBF_CHECK_ARG(N >= 1 && M >= 0, "%s: bad extents N=%d M=%d", fn, N, M);
  MACBF_CHECK_ARG(K >= 1 && K <= MACBF_MAX_K && K <= N + M, "%s: K=%d out of range (max %d, candidates %d)", fn, K,
                  MACBF_MAX_K, N + M);
  MACBF_CHECK_ARG((long long)B * N * K < (1ll << 40), "%s: problem too large", fn);
  MACBF_CHECK_ARG(states != nullptr && idx != nullptr && h != nullptr && mask != nullptr && scalars != nullptr,
                  "%s: states, idx, h, ttc_mask and scalars must not be NULL", fn);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG((((uintptr_t)states) & 15) == 0, "%s: states must be 16-byte aligned", fn);
  MACBF_CHECK_ARG((((uintptr_t)goals) & 7) == 0 && (((uintptr_t)u) & 7) == 0 && (((uintptr_t)obstacles) & 7) == 0,
                  "%s: goals, u and obstacles must be 8-byte aligned", fn);
  MACBF_CHECK_ARG(((((uintptr_t)idx) | ((uintptr_t)h) | ((uintptr_t)scalars)) & 3) == 0,
                  "%s: idx, h and scalars must be 4-byte aligned", fn);
  return 0;
}

int loss_forward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                 const float* h, const float* h_next, const float* u, int B, int N, int M, int K,
                 double dist_min, double ttc, double dt, double alpha, uint8_t* mask, float* scalars,
                 void* scratch, size_t scratch_bytes, cudaStream_t st) {
  const int nchunks = loss_chunks(N, K);
  MACBF_CHECK_ARG(scratch_bytes >= loss_scratch_bytes(B, N, K), "loss_forward: scratch too small (%zu < %zu)",
                  scratch_bytes, loss_scratch_bytes(B, N, K));
  const float r2 = (float)(dist_min * dist_min);
  const float dt_alpha = (float)(dt * alpha);     // python: TIME_STEP * ALPHA_CBF in double, then cast
  int chunk0 = 0, nlaunch = nchunks, a_begin = 0, a_end = N;
  if (rows_active()) {
    // a row range of one env: only its chunks are computed (the caller zeroed the scratch; the partial sums of all
    // ranges are added - one all-reduce - before loss_final).  Range bounds are multiples of 1024 agents = whole chunks.
    MACBF_CHECK_ARG(B == 1, "loss_forward: a row range needs B = 1");
    a_begin = tl_row_range.begin;
    a_end = tl_row_range.end < N ? tl_row_range.end : N;
    if (a_end <= a_begin) return 0;
    MACBF_CHECK_ARG(((long long)a_begin * K) % LOSS_ROWS_PER_CTA == 0 && (a_end == N || ((long long)a_end * K) % LOSS_ROWS_PER_CTA == 0),
                    "loss_forward: row range [%d, %d) does not fall on reduction chunks (use multiples of 1024 agents)", a_begin, a_end);
    chunk0 = (int)(((long long)a_begin * K) / LOSS_ROWS_PER_CTA);
    nlaunch = (int)((((long long)a_end * K) + LOSS_ROWS_PER_CTA - 1) / LOSS_ROWS_PER_CTA) - chunk0;
  }
  dim3 grid(nlaunch, B);
  loss_partial_kernel<<<grid, LOSS_THREADS, 0, st>>>(states, goals, obstacles, idx, h, h_next,
                                                    (goals != nullptr) ? u : nullptr, N, M, K, r2, (float)ttc,
                                                    dt_alpha, mask, (float*)scratch, chunk0, nchunks, a_begin, a_end);
  MACBF_CHECK_LAUNCH("loss_partial_kernel");
  if (rows_active()) return 0;                    // loss_finalize() after the partial sums of all ranges were added
  loss_final_kernel<<<B, 32 * LOSS_NACC, 0, st>>>((const float*)scratch, nchunks, N, K, h_next != nullptr,
                                      (u != nullptr && goals != nullptr), scalars);
  MACBF_CHECK_LAUNCH("loss_final_kernel");
  return 0;
}

// second stage alone: partial sums (scratch) -> the 12 scalars per env
int loss_finalize(const void* scratch, int B, int N, int K, int has_next, int has_u, float* scalars, cudaStream_t st) {
  loss_final_kernel<<<B, 32 * LOSS_NACC, 0, st>>>((const float*)scratch, loss_chunks(N, K), N, K, has_next, has_u, scalars);
  MACBF_CHECK_LAUNCH("loss_final_kernel");
  return 0;
}

int loss_backward(const float* states, const float* goals, const float* h, const float* h_next, const float* u,
                  const uint8_t* mask, const float* scalars, const float* grad_scalars, int B, int N, int K,
                  double dt, double alpha, float* gh, float* ghn, float* gu, float* gst, cudaStream_t st) {
  const long long rows_env = (long long)N * K;
  int gx = (int)(((rows_env + 255) / 256) < 4ll * NUM_SMS ? ((rows_env + 255) / 256) : 4ll * NUM_SMS);
  if (gx < 1) gx = 1;
  dim3 grid(gx, B);
  loss_backward_kernel<<<grid, 256, 0, st>>>(states, goals, h, h_next, (goals != nullptr) ? u : nullptr, mask,
                                            scalars, grad_scalars, N, K, (float)(dt * alpha), gh, ghn, gu, gst);
  MACBF_CHECK_LAUNCH("loss_backward_kernel");
  return 0;
}

}  // namespace macbf

// ---------------------------------------------------------------------------------------------
// macbf_loss_forward_noalloc: the loss reduction for callers that cannot allocate device memory.
//
// Who needs this: code that runs where a device allocation is forbidden or impossible - inside
// a region the host framework has frozen (an allocator that is locked while a step is being
// recorded or replayed), in a real-time control loop that may not call into the driver's
// allocator because its latency is unbounded, or in an embedding host (a C or Fortran simulator
// calling this library directly) that simply owns no device allocator to size the scratch
// buffer macbf_loss_forward asks for.  Such callers still pass their inputs and outputs; only
// the reduction's partial sums, which no caller ever reads, live in the fixed array below.
//
// The array is part of the library image: every GPU that loads libmacbf_b200.so carries these
// 256 KiB until the process exits, whether or not this entry point is used.
// ---------------------------------------------------------------------------------------------
namespace macbf {

static __device__ __align__(256) float g_loss_static_scratch[MACBF_LOSS_NOALLOC_MAX_FLOATS];

// In host code `g_loss_static_scratch` is only a handle for the runtime, not a device address;
// the address has to be asked for, per device (each device has its own copy of the array).
constexpr int kMaxDevices = 64;
static float* g_loss_static_ptr[kMaxDevices] = {nullptr};

static int loss_static_scratch_for_current_device(float** out) {
  int dev = -1;
  MACBF_CHECK_CUDA(cudaGetDevice(&dev));
  MACBF_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "macbf_loss_forward_noalloc: device ordinal %d not supported", dev);
  if (g_loss_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_loss_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("macbf_loss_forward_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return 2;
    }
    g_loss_static_ptr[dev] = p;   // same value from every thread, so a racing first call is benign
  }
  *out = g_loss_static_ptr[dev];
  return 0;
}

}  // namespace macbf

extern "C" int macbf_loss_forward_noalloc(const float* states, const float* goals, const float* obstacles,
                                          const int32_t* idx, const float* h, const float* h_next, const float* u,
                                          int B, int N, int M, int K, double dist_min_thres, double ttc, double dt,
                                          double alpha, uint8_t* ttc_mask_out, float* scalars, void* stream) {
  using namespace macbf;
  const char* fn = "macbf_loss_forward_noalloc";
  if (loss_validate(fn, states, goals, obstacles, idx, h, u, B, N, M, K, ttc_mask_out, scalars)) return 1;
  // The fixed array exists once per device and its address is looked up for the CURRENT device, so the buffers
  // (and the stream) have to belong to the current device: refuse anything else instead of mixing two GPUs.
  {
    int cur = -1;
    MACBF_CHECK_CUDA(cudaGetDevice(&cur));
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, states) == cudaSuccess && at.type == cudaMemoryTypeDevice)
      MACBF_CHECK_ARG(at.device == cur, "%s: states lives on device %d but device %d is current; make the buffers' "
                      "device current (cudaSetDevice) before the call", fn, at.device, cur);
    else
      (void)cudaGetLastError();
  }
  const size_t need = loss_scratch_bytes(B, N, K);
  MACBF_CHECK_ARG(need <= sizeof(float) * (size_t)MACBF_LOSS_NOALLOC_MAX_FLOATS,
                  "%s: problem too large for the built-in scratch: B=%d N=%d K=%d needs %zu bytes of partial sums, "
                  "the fixed array holds %zu (use macbf_loss_forward with caller scratch instead)",
                  fn, B, N, K, need, sizeof(float) * (size_t)MACBF_LOSS_NOALLOC_MAX_FLOATS);
  float* scratch = nullptr;
  int rc = loss_static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // Every partial-sum slot that loss_final_kernel reads is written by loss_partial_kernel in this
  // same call, so nothing in the array has to start at zero and the previous call's contents
  // are simply overwritten.
  return loss_forward(states, goals, obstacles, idx, h, h_next, u, B, N, M, K, dist_min_thres, ttc, dt, alpha,
                      ttc_mask_out, scalars, scratch, need, (cudaStream_t)stream);
}
