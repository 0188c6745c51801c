// Scenario generation on the device (core.py:45-71 generate_data): starts uniform in a square of side
// sqrt(max(1, N/8)) and goals within +-0.5 of them, every pair farther apart than dist_min_thres - the
// reference's sequential dart throwing (an O(N^2) Python loop, 1.6 s at N = 4096), including its quirk that
// not-yet-placed rows sit at the origin and repel candidates.
//
// The algorithm is the SAME sequential process, run on a counter-based random stream so that it can be resolved
// in parallel: candidate t is accepted iff no ACCEPTED candidate s < t lies within the threshold.  Every candidate
// looks at the lower-numbered candidates around it (uniform grid, linked cell lists): a conflicting accepted one
// rejects it, a conflicting undecided one makes it wait, otherwise it is accepted; the fixed point is the
// sequential result (conflicts are rare, so a handful of rounds).  Goals: agent i takes the first draw of its own
// substream that clears the goals of all agents j < i; it can decide once every lower-numbered agent whose +-0.5
// box can reach its own is settled.  oracle/macbf_oracle.py:generate_data_stream restates the sequential process
// on the same stream; the test demands identical bits.  (The stream is splitmix64-based, not numpy's MT19937: the
// host generator core.generate_data stays the np.random-compatible one.)
//
// One CTA per env; fp32 arithmetic with uncontracted ops so the numpy restatement reproduces every decision.
#include "common.cuh"

namespace macbf {

namespace scen {
constexpr int THREADS = 1024;
constexpr int GMAX = 1024;       // position-conflict grid: cells per side
constexpr int G2MAX = 512;       // reach grid of the goal stage

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31;
  return z;
}
// two uniforms in [0,1) with 24 random bits each (exactly representable in fp32)
__device__ __forceinline__ float2 rng2(unsigned long long seed, int env, int stream, unsigned long long index) {
  unsigned long long h = mix64(seed ^ mix64((unsigned long long)env + 1ull));
  h = mix64(h ^ mix64((unsigned long long)stream + 0x1234567ull));
  h = mix64(h + index * 0x9E3779B97F4A7C15ull);
  return make_float2((float)(h >> 40) * 5.9604644775390625e-8f, (float)((h >> 8) & 0xFFFFFFull) * 5.9604644775390625e-8f);
}
__device__ __forceinline__ float dist2d(float ax, float ay, float bx, float by) {
  const float dx = xsub(ax, bx), dy = xsub(ay, by);
  return __fsqrt_rn(xadd(xmul(dx, dx), xmul(dy, dy)));
}
__device__ __forceinline__ int cell_of(float v, float inv_cs, int G) {
  int c = (int)(v * inv_cs);
  return c < 0 ? 0 : (c >= G ? G - 1 : c);
}
}  // namespace scen

struct ScenarioLayout {   // per-env workspace, in 4-byte words
  int T, G, G2;
  size_t o_cand, o_head, o_next, o_status, o_head2, o_next2, o_settled, o_goal, words;
};
static ScenarioLayout scenario_layout(int N, int T) {
  ScenarioLayout L;
  const double side = sqrt(fmax(1.0, N / 8.0));
  L.T = T;
  L.G = (int)fmin((double)scen::GMAX, fmax(1.0, floor(side / 0.25)));      // cell >= 0.25 >= any sensible threshold (checked)
  L.G2 = (int)fmin((double)scen::G2MAX, fmax(1.0, floor(side / 0.5)));
  size_t o = 0;
  auto take = [&o](size_t words) { const size_t at = o; o = (o + words + 3) & ~(size_t)3; return at; };   // 16-byte sections
  L.o_cand = take(2 * (size_t)T);
  L.o_head = take((size_t)L.G * L.G);
  L.o_next = take(T);
  L.o_status = take(T);
  L.o_head2 = take((size_t)L.G2 * L.G2);
  L.o_next2 = take(N);
  L.o_settled = take(N);
  L.o_goal = take(2 * (size_t)N);
  L.words = (o + 63) & ~(size_t)63;
  return L;
}

__global__ void __launch_bounds__(scen::THREADS)
scenario_kernel(unsigned long long seed, int N, float side, float thres, ScenarioLayout L, int* __restrict__ ws,
                float* __restrict__ states, float* __restrict__ goals, int32_t* __restrict__ accepted_out) {
  using namespace scen;
  const int env = blockIdx.x;
  int* w = ws + (size_t)env * L.words;
  float2* cand = reinterpret_cast<float2*>(w + L.o_cand);
  int* head = w + L.o_head; int* next = w + L.o_next; int* status = w + L.o_status;
  int* head2 = w + L.o_head2; int* next2 = w + L.o_next2; int* settled = w + L.o_settled;
  float2* goal = reinterpret_cast<float2*>(w + L.o_goal);
  float4* st = reinterpret_cast<float4*>(states) + (size_t)env * N;
  float2* gl = reinterpret_cast<float2*>(goals) + (size_t)env * N;
  const int T = L.T, G = L.G, G2 = L.G2;
  const float inv_cs = (float)G / side, inv_cs2 = (float)G2 / side;
  __shared__ int s_scan[32];
  __shared__ int s_base;

  // ---------------- stage 1: positions ----------------
  for (int c = threadIdx.x; c < G * G; c += THREADS) head[c] = -1;
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += THREADS) {
    const float2 u = rng2(seed, env, 0, (unsigned long long)t);
    const float2 c = make_float2(xmul(u.x, side), xmul(u.y, side));
    cand[t] = c;
    const bool far_from_origin = dist2d(c.x, c.y, 0.f, 0.f) > thres;      // unplaced rows are (0,0): core.py:47,54
    status[t] = far_from_origin ? 0 : 2;
    if (far_from_origin) next[t] = atomicExch(&head[cell_of(c.y, inv_cs, G) * G + cell_of(c.x, inv_cs, G)], t);
  }
  __syncthreads();
  for (;;) {
    int pending = 0;
    for (int t = threadIdx.x; t < T; t += THREADS) {
      if (status[t] != 0) continue;
      const float2 c = cand[t];
      const int cx = cell_of(c.x, inv_cs, G), cy = cell_of(c.y, inv_cs, G);
      int verdict = 1;                                   // accepted unless a lower-numbered neighbour objects
      for (int yy = max(cy - 1, 0); yy <= min(cy + 1, G - 1) && verdict != 2; ++yy)
        for (int xx = max(cx - 1, 0); xx <= min(cx + 1, G - 1) && verdict != 2; ++xx)
          for (int s = head[yy * G + xx]; s >= 0; s = next[s]) {
            if (s >= t) continue;
            const float2 o = cand[s];
            if (dist2d(o.x, o.y, c.x, c.y) > thres) continue;
            const int ss = reinterpret_cast<volatile int*>(status)[s];
            if (ss == 1) { verdict = 2; break; }
            if (ss == 0) verdict = 0;
          }
      if (verdict == 0) pending = 1;
      else reinterpret_cast<volatile int*>(status)[t] = verdict;
    }
    if (!__syncthreads_or(pending)) break;
  }
  // rank of every accepted candidate in stream order; the first N become the agents
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int t0 = 0; t0 < T; t0 += THREADS) {
    const int t = t0 + threadIdx.x;
    const int acc = (t < T && status[t] == 1) ? 1 : 0;
    int incl = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += v;
    }
    if ((threadIdx.x & 31) == 31) s_scan[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      int v = s_scan[threadIdx.x];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xffffffffu, v, o);
        if (threadIdx.x >= o) v += x;
      }
      s_scan[threadIdx.x] = v;
    }
    __syncthreads();
    const int rank = s_base + incl - acc + ((threadIdx.x >> 5) ? s_scan[(threadIdx.x >> 5) - 1] : 0);
    if (acc && rank < N) {
      const float2 c = cand[t];
      st[rank] = make_float4(c.x, c.y, 0.f, 0.f);
    }
    __syncthreads();
    if (threadIdx.x == 0) s_base += s_scan[31];
    __syncthreads();
  }
  const int n_acc = s_base;
  if (threadIdx.x == 0) accepted_out[env] = n_acc;
  if (n_acc < N) return;                                  // not enough candidates: the host retries with more
  __threadfence_block();
  __syncthreads();

  // ---------------- stage 2: goals ----------------
  for (int c = threadIdx.x; c < G2 * G2; c += THREADS) head2[c] = -1;
  for (int i = threadIdx.x; i < N; i += THREADS) settled[i] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < N; i += THREADS) {
    const float4 p = st[i];
    next2[i] = atomicExch(&head2[cell_of(p.y, inv_cs2, G2) * G2 + cell_of(p.x, inv_cs2, G2)], i);
  }
  __syncthreads();
  const float reach = 1.0f + thres + 1e-3f;               // |p_i - p_j|_inf beyond this: the goals cannot meet
  const int rc = (int)ceilf(reach * inv_cs2);
  for (int round = 1;; ++round) {
    int pending = 0;
    for (int i = threadIdx.x; i < N; i += THREADS) {
      if (settled[i] != 0) continue;
      const float4 p = st[i];
      const int cx = cell_of(p.x, inv_cs2, G2), cy = cell_of(p.y, inv_cs2, G2);
      const int y0 = max(cy - rc, 0), y1 = min(cy + rc, G2 - 1), x0 = max(cx - rc, 0), x1 = min(cx + rc, G2 - 1);
      bool ready = true;
      for (int yy = y0; yy <= y1 && ready; ++yy)
        for (int xx = x0; xx <= x1 && ready; ++xx)
          for (int j = head2[yy * G2 + xx]; j >= 0; j = next2[j]) {
            if (j >= i) continue;
            const float4 q = st[j];
            if (fabsf(q.x - p.x) > reach || fabsf(q.y - p.y) > reach) continue;
            const int sj = settled[j];
            if (sj == 0 || sj >= round) { ready = false; break; }     // settled in an EARLIER round only
          }
      if (!ready) { pending = 1; continue; }
      float2 g = make_float2(0.f, 0.f);
      for (unsigned long long a = 0;; ++a) {
        const float2 u = rng2(seed, env, 1, ((unsigned long long)i << 20) | a);
        g = make_float2(xadd(xsub(u.x, 0.5f), p.x), xadd(xsub(u.y, 0.5f), p.y));      // core.py:59
        bool ok = dist2d(g.x, g.y, 0.f, 0.f) > thres;                                 // unplaced goal rows are (0,0)
        for (int yy = y0; yy <= y1 && ok; ++yy)
          for (int xx = x0; xx <= x1 && ok; ++xx)
            for (int j = head2[yy * G2 + xx]; j >= 0; j = next2[j]) {
              if (j >= i) continue;
              const float4 q = st[j];
              if (fabsf(q.x - p.x) > reach || fabsf(q.y - p.y) > reach) continue;
              const float2 gj = goal[j];
              if (dist2d(gj.x, gj.y, g.x, g.y) <= thres) { ok = false; break; }
            }
        if (ok || a >= (1ull << 20) - 1) break;
      }
      goal[i] = g;
      gl[i] = g;
      settled[i] = round;
    }
    if (!__syncthreads_or(pending)) break;
  }
}

size_t scenario_workspace_bytes(int B, int N, int T) { return (size_t)B * scenario_layout(N, T).words * sizeof(int); }

int scenario_generate(unsigned long long seed, int B, int N, int T, double dist_min_thres, float* states, float* goals,
                      int32_t* accepted, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  MACBF_CHECK_ARG(dist_min_thres > 0.0 && dist_min_thres <= 0.25, "generate_scenario: dist_min_thres=%g must be in (0, 0.25]",
                  dist_min_thres);
  MACBF_CHECK_ARG(T >= N, "generate_scenario: %d candidates for %d agents", T, N);
  const ScenarioLayout L = scenario_layout(N, T);
  MACBF_CHECK_ARG(workspace_bytes >= (size_t)B * L.words * sizeof(int), "generate_scenario: workspace too small (%zu < %zu)",
                  workspace_bytes, (size_t)B * L.words * sizeof(int));
  const float side = (float)sqrt(fmax(1.0, N / 8.0));
  scenario_kernel<<<B, scen::THREADS, 0, st>>>(seed, N, side, (float)dist_min_thres, L, (int*)workspace, states, goals, accepted);
  MACBF_CHECK_LAUNCH("scenario_kernel");
  return 0;
}

// ---- obstacle point sets (core.py:7-42) ----------------------------------------------------------------------
// `num` points per obstacle on a circle (kind 0: size = (radius, unused)) or on a rectangle's perimeter, clockwise
// from the top-left corner (kind 1: size = (a, b)); float64 arithmetic in numpy's operation order (np.linspace with
// endpoint=False is i * ((stop - start) / n) + start, the centre is added last), one thread per point.
__device__ __forceinline__ double lin(double start, double stop, int n, int i) {
  return __dadd_rn(__dmul_rn((double)i, __ddiv_rn(__dsub_rn(stop, start), (double)n)), start);
}
__global__ void obstacle_points_kernel(const int32_t* __restrict__ kinds, const double* __restrict__ centers,
                                       const double* __restrict__ sizes, int n_obstacles, int num, double* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_obstacles * num) return;
  const int o = t / num, i = t - o * num;
  const double cx = centers[2 * o], cy = centers[2 * o + 1];
  double x, y;
  if (kinds[o] == 0) {
    const double r = sizes[2 * o];
    const double th = lin(0.0, __dmul_rn(3.141592653589793, 2.0), num, i);
    x = __dmul_rn(cos(th), r); y = __dmul_rn(sin(th), r);
  } else {
    const double a = sizes[2 * o], b = sizes[2 * o + 1];
    const int n1 = (int)__ddiv_rn(__dmul_rn((double)(num / 2), a), __dadd_rn(a, b));     // int(num // 2 * a / (a + b))
    const int n2 = num / 2 - n1, n3 = n1;
    const double ha = __ddiv_rn(a, 2.0), hb = __ddiv_rn(b, 2.0);
    if (i < n1) { x = lin(-ha, ha, n1, i); y = hb; }                                      // top
    else if (i < n1 + n2) { x = ha; y = lin(hb, -hb, n2, i - n1); }                       // right
    else if (i < n1 + n2 + n3) { x = lin(ha, -ha, n3, i - n1 - n2); y = -hb; }            // bottom
    else { x = -ha; y = lin(-hb, hb, num - n1 - n2 - n3, i - n1 - n2 - n3); }             // left
  }
  out[2 * t] = __dadd_rn(x, cx);
  out[2 * t + 1] = __dadd_rn(y, cy);
}

int obstacle_points(const int32_t* kinds, const double* centers, const double* sizes, int n_obstacles, int num, double* out,
                    cudaStream_t st) {
  const int total = n_obstacles * num;
  obstacle_points_kernel<<<(total + 127) / 128, 128, 0, st>>>(kinds, centers, sizes, n_obstacles, num, out);
  MACBF_CHECK_LAUNCH("obstacle_points_kernel");
  return 0;
}

}  // namespace macbf
