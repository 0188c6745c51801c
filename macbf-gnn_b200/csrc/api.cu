// extern "C" surface of libmacbf_b200.so (declared in include/macbf_b200.h).
#include <stdarg.h>

#include "common.cuh"

namespace macbf {

static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;
thread_local const int* tl_skip_flag = nullptr;
thread_local QueryPart tl_query_part = {0, 1};
thread_local RowRange tl_row_range = {0, -1};
thread_local GraphPair tl_graph_pair = {0, 0.f, nullptr};
thread_local uint8_t* tl_save_winners = nullptr;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches += n; }

// ---- per-launch timing -------------------------------------------------------------------------
constexpr int kTimingMaxMarks = 4096;
struct TimingState {
  bool on = false;
  cudaStream_t stream = nullptr;
  int used = 0;
  cudaEvent_t ev[kTimingMaxMarks];
  int created = 0;
  const char* name[kTimingMaxMarks];
};
static thread_local TimingState* g_timing = nullptr;

void timing_mark(const char* name) {
  TimingState* t = g_timing;
  if (t == nullptr || !t->on || t->used >= kTimingMaxMarks) return;
  if (t->used >= t->created) {
    if (cudaEventCreate(&t->ev[t->created]) != cudaSuccess) { (void)cudaGetLastError(); return; }
    ++t->created;
  }
  if (cudaEventRecord(t->ev[t->used], t->stream) != cudaSuccess) { (void)cudaGetLastError(); return; }
  t->name[t->used++] = name;
}

// implemented in the kernel translation units
int neighbor_topk(const float*, const float*, int, int, int, int, int32_t*, cudaStream_t);
int dense_topk(const float*, int, int, int, int, int32_t*, cudaStream_t);
int obstacle_points(const int32_t*, const double*, const double*, int, int, double*, cudaStream_t);
size_t rollout_workspace_bytes(int, int, int, int);
size_t rollout_backward_scratch_bytes(int, int, int, int);
int rollout_backward(const float*, const float*, const float*, const float*, const float*, const float*, int, int, int, int, int,
                     double, double, float*, float*, float*, void*, size_t, cudaStream_t);
int rollout_forward(const float*, const float*, const float*, const float*, const float*, int, int, int, int, int, int, double, double,
                    double, float*, float*, int32_t*, void*, size_t, cudaStream_t);
size_t neighbor_grid_workspace_bytes(int, int, int);
bool neighbor_grid_applicable(int, int, int);
int neighbor_topk_grid(const float*, const float*, int, int, int, int, int32_t*, void*, size_t, cudaStream_t);
bool neighbor_grid_pair_applicable(int, int, int);
size_t neighbor_grid_pair_workspace_bytes(int, int, int);
int neighbor_topk_grid_pair(const float*, const float*, int, int, int, int, double, int32_t*, int32_t*, void*, size_t, cudaStream_t);

// grid search when it applies (K <= 128, enough candidates) and workspace was given, else brute force
static int neighbor_any(const float* states, const float* obstacles, int B, int N, int M, int top_k, int32_t* idx,
                        void* ws, size_t ws_bytes, cudaStream_t st) {
  if (ws != nullptr && neighbor_grid_applicable(N, M, top_k))
    return neighbor_topk_grid(states, obstacles, B, N, M, top_k, idx, ws, ws_bytes, st);
  return neighbor_topk(states, obstacles, B, N, M, top_k, idx, st);
}
int cbf_forward(const float*, const float*, const int32_t*, const float*, int, int, int, int, int, double, double,
                float*, cudaStream_t);
int controller_forward(const float*, const float*, const float*, const int32_t*, const float*, int, int, int, int,
                       int, double, double, float*, float*, float*, cudaStream_t);
int ttc_mask(const float*, const float*, const int32_t*, int, int, int, int, double, double, uint8_t*, cudaStream_t);
int ttc_dense(const float*, int, int, double, double, int32_t*, uint8_t*, cudaStream_t);
size_t scenario_workspace_bytes(int, int, int);
int scenario_generate(unsigned long long, int, int, int, double, float*, float*, int32_t*, void*, size_t, cudaStream_t);
size_t loss_scratch_bytes(int, int, int);
int loss_finalize(const void*, int, int, int, int, int, float*, cudaStream_t);
int loss_validate(const char*, const float*, const float*, const float*, const int32_t*, const float*, const float*, int,
                  int, int, int, const uint8_t*, const float*);
int loss_forward(const float*, const float*, const float*, const int32_t*, const float*, const float*, const float*,
                 int, int, int, int, double, double, double, double, uint8_t*, float*, void*, size_t, cudaStream_t);
int loss_backward(const float*, const float*, const float*, const float*, const float*, const uint8_t*, const float*,
                  const float*, int, int, int, double, double, float*, float*, float*, float*, cudaStream_t);
size_t cbf_backward_scratch_bytes(int, int, int);
size_t cbf_record_bytes(int, int, int);
size_t cbf_backward_record_scratch_bytes();
int cbf_forward_record(const float*, const float*, const int32_t*, const float*, int, int, int, int, double, double, float*, void*,
                       size_t, cudaStream_t);
int cbf_backward_record(const int32_t*, const float*, const float*, int, int, int, void*, size_t, float*, float*, void*, size_t,
                        cudaStream_t);
int cbf_backward(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, double,
                 double, float*, float*, void*, size_t, cudaStream_t);
size_t controller_backward_scratch_bytes(int, int, int);
bool controller_forward_is_tensor_core(int semantics);
int controller_backward(const float*, const float*, const float*, const int32_t*, const float*, const float*,
                        const float*, int, int, int, int, double, double, float*, float*, float*, void*, size_t,
                        cudaStream_t, float*, const uint8_t*);

static int check_dims(const char* fn, int B, int N, int M, int K) {
  MACBF_CHECK_ARG(B >= 1 && B <= 65535, "%s: B=%d out of range [1,65535]", fn, B);
  MACBF_CHECK_ARG(N >= 1, "%s: N=%d must be >= 1", fn, N);
  MACBF_CHECK_ARG(M >= 0, "%s: M=%d must be >= 0", fn, M);
  MACBF_CHECK_ARG(K >= 1 && K <= MACBF_MAX_K, "%s: K=%d out of range [1,%d]", fn, K, MACBF_MAX_K);
  MACBF_CHECK_ARG(K <= N + M, "%s: K=%d exceeds the candidate count N+M=%d", fn, K, N + M);
  MACBF_CHECK_ARG((long long)B * N * K < (1ll << 40), "%s: problem too large", fn);
  return 0;
}
#define REQ_PTR(fn, p) MACBF_CHECK_ARG((p) != nullptr, "%s: %s is NULL", fn, #p)
#define REQ_A16(fn, p) MACBF_CHECK_ARG((((uintptr_t)(p)) & 15) == 0, "%s: %s must be 16-byte aligned", fn, #p)

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// ---- side stream of the fused step ---------------------------------------------------------------------------------
// The neighbour search on S' (grid build + one thread per query: latency-bound, 14 warps per SM, no shared memory)
// only needs s_next, and CBF(S) only needs idx: the fused step runs the search on a second stream UNDER the tensor-core
// CBF(S) kernel, whose persistent CTAs leave a quarter of the register file and most issue slots idle.  Fork / join by
// events (capturable).  The stream and the two events are created once per (host thread, device) on first use and kept.
struct SideStream { cudaStream_t stream = nullptr; cudaEvent_t fork = nullptr, join = nullptr; };
static int side_stream_for_current_device(SideStream** out) {
  static thread_local SideStream side[64];        // per host thread: two threads never share fork / join events
  int dev = -1;
  MACBF_CHECK_CUDA(cudaGetDevice(&dev));
  MACBF_CHECK_ARG(dev >= 0 && dev < 64, "device ordinal %d not supported", dev);
  SideStream& s = side[dev];
  if (s.stream == nullptr) {
    MACBF_CHECK_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    MACBF_CHECK_CUDA(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
    MACBF_CHECK_CUDA(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  }
  *out = &s;
  return 0;
}


}  // namespace macbf

using namespace macbf;

extern "C" {

const char* macbf_last_error(void) { return g_err; }
int macbf_version(void) { return 100; }
long long macbf_launch_count(void) { return g_launches; }

int macbf_timing_begin(void* stream) {
  if (g_timing == nullptr) g_timing = new TimingState();
  g_timing->on = true;
  g_timing->stream = (cudaStream_t)stream;
  g_timing->used = 0;
  timing_mark("(begin)");
  return g_timing->used == 1 ? 0 : 2;
}

int macbf_timing_end(int max_marks, char* names, int name_stride, float* ms) {
  TimingState* t = g_timing;
  MACBF_CHECK_ARG(t != nullptr && t->on, "macbf_timing_end: macbf_timing_begin was not called on this thread");
  t->on = false;
  MACBF_CHECK_ARG(names != nullptr && ms != nullptr && name_stride >= 8, "macbf_timing_end: bad output buffers");
  if (t->used > 0) MACBF_CHECK_CUDA(cudaEventSynchronize(t->ev[t->used - 1]));
  int n = 0;
  for (int i = 1; i < t->used && n < max_marks; ++i, ++n) {
    float v = 0.f;
    MACBF_CHECK_CUDA(cudaEventElapsedTime(&v, t->ev[i - 1], t->ev[i]));
    ms[n] = v;
    strncpy(names + (size_t)n * name_stride, t->name[i], name_stride - 1);
    names[(size_t)n * name_stride + name_stride - 1] = 0;
  }
  return -n;   // <= 0: minus the number of marks returned; > 0: error
}
int macbf_neighbor_count(int N, int M, int top_k) { return (N + M) < top_k ? (N + M) : top_k; }

int macbf_neighbor_topk(const float* states, const float* obstacles, int B, int N, int M, int top_k, int32_t* idx,
                        void* stream) {
  const char* fn = "macbf_neighbor_topk";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  if (check_dims(fn, B, N, M, macbf_neighbor_count(N, M, top_k))) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return neighbor_topk(states, obstacles, B, N, M, top_k, idx, (cudaStream_t)stream);
}

int macbf_dense_topk(const float* x, int N, int C, int F, int k, int32_t* idx, void* stream) {
  const char* fn = "macbf_dense_topk";
  MACBF_CHECK_ARG(N >= 1 && C >= 1 && F >= 2, "%s: bad extents N=%d C=%d F=%d (F >= 2: dx, dy lead)", fn, N, C, F);
  MACBF_CHECK_ARG(k >= 1 && k <= MACBF_MAX_K && k < C, "%s: k=%d must be in [1,%d] and < C=%d", fn, k, MACBF_MAX_K, C);
  REQ_PTR(fn, x); REQ_PTR(fn, idx);
  return dense_topk(x, N, C, F, k, idx, (cudaStream_t)stream);
}

size_t macbf_neighbor_workspace_bytes(int B, int N, int M, int top_k) {
  return neighbor_grid_applicable(N, M, top_k) ? neighbor_grid_workspace_bytes(B, N, M) : 0;
}

int macbf_neighbor_topk_ws(const float* states, const float* obstacles, int B, int N, int M, int top_k, int32_t* idx,
                           void* workspace, size_t workspace_bytes, void* stream) {
  const char* fn = "macbf_neighbor_topk_ws";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  if (check_dims(fn, B, N, M, macbf_neighbor_count(N, M, top_k))) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(workspace != nullptr || macbf_neighbor_workspace_bytes(B, N, M, top_k) == 0,
                  "%s: workspace is NULL but %zu bytes are needed", fn, macbf_neighbor_workspace_bytes(B, N, M, top_k));
  return neighbor_any(states, obstacles, B, N, M, top_k, idx, workspace, workspace_bytes, (cudaStream_t)stream);
}

int macbf_neighbor_topk_part(const float* states, const float* obstacles, int B, int N, int M, int top_k, int part,
                             int nparts, int32_t* idx, void* workspace, size_t workspace_bytes, void* stream) {
  const char* fn = "macbf_neighbor_topk_part";
  MACBF_CHECK_ARG(nparts >= 1 && part >= 0 && part < nparts, "%s: part %d of %d", fn, part, nparts);
  tl_query_part = QueryPart{part, nparts};
  const int rc = macbf_neighbor_topk_ws(states, obstacles, B, N, M, top_k, idx, workspace, workspace_bytes, stream);
  tl_query_part = QueryPart{0, 1};
  return rc;
}

int macbf_cbf_forward(const float* states, const float* obstacles, const int32_t* idx, const float* cbf_params, int B,
                      int N, int M, int K, int semantics, double dist_min_thres, double obs_radius, float* h,
                      void* stream) {
  const char* fn = "macbf_cbf_forward";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_PTR(fn, cbf_params); REQ_PTR(fn, h); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(semantics == MACBF_LITERAL || semantics == MACBF_REPAIRED, "%s: bad semantics %d", fn, semantics);
  return cbf_forward(states, obstacles, idx, cbf_params, B, N, M, K, semantics, dist_min_thres, obs_radius, h,
                     (cudaStream_t)stream);
}

int macbf_controller_forward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                             const float* ctrl_params, int B, int N, int M, int K, int semantics, double obs_radius,
                             double dt, float* pool, float* u, float* s_next, void* stream) {
  const char* fn = "macbf_controller_forward";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, idx); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, pool);
  REQ_PTR(fn, u); REQ_A16(fn, states); REQ_A16(fn, pool);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(semantics == MACBF_LITERAL || semantics == MACBF_REPAIRED, "%s: bad semantics %d", fn, semantics);
  return controller_forward(states, goals, obstacles, idx, ctrl_params, B, N, M, K, semantics, obs_radius, dt, pool,
                            u, s_next, (cudaStream_t)stream);
}

int macbf_ttc_mask(const float* states, const float* obstacles, const int32_t* idx, int B, int N, int M, int K,
                   double dist_min_thres, double ttc, uint8_t* mask, void* stream) {
  const char* fn = "macbf_ttc_mask";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_PTR(fn, mask); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return ttc_mask(states, obstacles, idx, B, N, M, K, dist_min_thres, ttc, mask, (cudaStream_t)stream);
}

size_t macbf_loss_scratch_bytes(int B, int N, int K) { return loss_scratch_bytes(B, N, K); }

int macbf_loss_forward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                       const float* h, const float* h_next, const float* u, int B, int N, int M, int K,
                       double dist_min_thres, double ttc, double dt, double alpha, uint8_t* ttc_mask_out,
                       float* scalars, void* scratch, size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_loss_forward";
  if (loss_validate(fn, states, goals, obstacles, idx, h, u, B, N, M, K, ttc_mask_out, scalars)) return 1;
  REQ_PTR(fn, scratch);
  return loss_forward(states, goals, obstacles, idx, h, h_next, u, B, N, M, K, dist_min_thres, ttc, dt, alpha,
                      ttc_mask_out, scalars, scratch, scratch_bytes, (cudaStream_t)stream);
}

int macbf_loss_backward(const float* states, const float* goals, const float* h, const float* h_next, const float* u,
                        const uint8_t* ttc_mask_in, const float* scalars, const float* grad_scalars, int B, int N,
                        int K, double dt, double alpha, float* grad_h, float* grad_h_next, float* grad_u,
                        float* grad_states, void* stream) {
  const char* fn = "macbf_loss_backward";
  MACBF_CHECK_ARG(B >= 1 && B <= 65535 && N >= 1 && K >= 1 && K <= MACBF_MAX_K, "%s: bad extents B=%d N=%d K=%d", fn, B, N, K);
  REQ_PTR(fn, h); REQ_PTR(fn, ttc_mask_in); REQ_PTR(fn, scalars); REQ_PTR(fn, grad_scalars);
  MACBF_CHECK_ARG((grad_u == nullptr && grad_states == nullptr) || u == nullptr ||
                      (states != nullptr && goals != nullptr),
                  "%s: grad_u / grad_states need states and goals", fn);
  return loss_backward(states, goals, h, h_next, u, ttc_mask_in, scalars, grad_scalars, B, N, K, dt, alpha, grad_h,
                       grad_h_next, grad_u, grad_states, (cudaStream_t)stream);
}

size_t macbf_cbf_backward_scratch_bytes(int B, int N, int K) { return cbf_backward_scratch_bytes(B, N, K); }

int macbf_cbf_backward(const float* states, const float* obstacles, const int32_t* idx, const float* cbf_params,
                       const float* grad_h, int B, int N, int M, int K, double dist_min_thres, double obs_radius,
                       float* grad_params, float* grad_states, void* scratch, size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_cbf_backward";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_PTR(fn, cbf_params); REQ_PTR(fn, grad_h); REQ_PTR(fn, grad_params);
  REQ_PTR(fn, scratch); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return cbf_backward(states, obstacles, idx, cbf_params, grad_h, B, N, M, K, dist_min_thres, obs_radius, grad_params,
                      grad_states, scratch, scratch_bytes, (cudaStream_t)stream);
}

size_t macbf_cbf_record_bytes(int B, int N, int K) { return cbf_record_bytes(B, N, K); }
size_t macbf_cbf_backward_record_scratch_bytes(void) { return cbf_backward_record_scratch_bytes(); }

int macbf_cbf_forward_record(const float* states, const float* obstacles, const int32_t* idx, const float* cbf_params, int B,
                             int N, int M, int K, double dist_min_thres, double obs_radius, float* h, void* record,
                             size_t record_bytes, void* stream) {
  const char* fn = "macbf_cbf_forward_record";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, idx); REQ_PTR(fn, cbf_params); REQ_PTR(fn, h); REQ_PTR(fn, record); REQ_A16(fn, states);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return cbf_forward_record(states, obstacles, idx, cbf_params, B, N, M, K, dist_min_thres, obs_radius, h, record, record_bytes,
                            (cudaStream_t)stream);
}

int macbf_cbf_backward_record(const int32_t* idx, const float* cbf_params, const float* grad_h, int B, int N, int K,
                              void* record, size_t record_bytes, float* grad_params, float* grad_states, void* scratch,
                              size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_cbf_backward_record";
  if (check_dims(fn, B, N, 0, K)) return 1;
  REQ_PTR(fn, idx); REQ_PTR(fn, cbf_params); REQ_PTR(fn, grad_h); REQ_PTR(fn, record); REQ_PTR(fn, grad_params);
  REQ_PTR(fn, scratch); REQ_A16(fn, scratch);
  return cbf_backward_record(idx, cbf_params, grad_h, B, N, K, record, record_bytes, grad_params, grad_states, scratch,
                             scratch_bytes, (cudaStream_t)stream);
}

size_t macbf_controller_backward_scratch_bytes(int B, int N, int K) {
  return controller_backward_scratch_bytes(B, N, K);
}

int macbf_controller_backward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                              const float* ctrl_params, const float* grad_u, const float* grad_s_next, int B, int N,
                              int M, int K, double obs_radius, double dt, float* grad_params, float* grad_states,
                              float* grad_goals, void* scratch, size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_controller_backward";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, idx); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, grad_params);
  REQ_PTR(fn, scratch); REQ_A16(fn, states);
  MACBF_CHECK_ARG(grad_u != nullptr || grad_s_next != nullptr, "%s: no incoming gradient", fn);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return controller_backward(states, goals, obstacles, idx, ctrl_params, grad_u, grad_s_next, B, N, M, K, obs_radius,
                             dt, grad_params, grad_states, grad_goals, scratch, scratch_bytes, (cudaStream_t)stream, nullptr,
                             nullptr);
}

int macbf_controller_backward_saved(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                                    const float* ctrl_params, const float* grad_u, const float* grad_s_next, int B, int N,
                                    int M, int K, double obs_radius, double dt, float* pool_saved,
                                    const uint8_t* winners_saved, float* grad_params, float* grad_states,
                                    float* grad_goals, void* scratch, size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_controller_backward_saved";
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, idx); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, grad_params);
  REQ_PTR(fn, scratch); REQ_PTR(fn, pool_saved); REQ_PTR(fn, winners_saved); REQ_A16(fn, states); REQ_A16(fn, pool_saved);
  REQ_A16(fn, winners_saved);
  MACBF_CHECK_ARG(grad_u != nullptr || grad_s_next != nullptr, "%s: no incoming gradient", fn);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return controller_backward(states, goals, obstacles, idx, ctrl_params, grad_u, grad_s_next, B, N, M, K, obs_radius,
                             dt, grad_params, grad_states, grad_goals, scratch, scratch_bytes, (cudaStream_t)stream,
                             pool_saved, winners_saved);
}

// workspace: [pool (B*N*128 f32)] [loss scratch] [neighbour-grid scratch]
size_t macbf_step_workspace_bytes(int B, int N, int M, int top_k) {
  const int K = macbf_neighbor_count(N, M, top_k);
  return macbf_step_workspace_winners_offset(B, N, M, top_k) + align256((size_t)B * N * MACBF_POOL_DIM);
}

// graph scratch of the fused step: both grids at once where that applies (neighbor_topk_grid_pair)
static size_t step_graph_scratch_bytes(int B, int N, int M, int top_k) {
  const size_t one = macbf_neighbor_workspace_bytes(B, N, M, top_k);
  const size_t two = neighbor_grid_pair_applicable(N, M, top_k) ? neighbor_grid_pair_workspace_bytes(B, N, M) : 0;
  return one > two ? one : two;
}
size_t macbf_step_workspace_winners_offset(int B, int N, int M, int top_k) {
  const int K = macbf_neighbor_count(N, M, top_k);
  return align256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)) + align256(loss_scratch_bytes(B, N, K)) +
         align256(step_graph_scratch_bytes(B, N, M, top_k));
}

int macbf_step_forward(const float* states, const float* goals, const float* obstacles, const float* cbf_params,
                       const float* ctrl_params, int B, int N, int M, int top_k, int semantics, double dist_min_thres,
                       double obs_radius, double ttc, double dt, double alpha, int32_t* idx, int32_t* idx_next,
                       float* u, float* s_next, float* h, float* h_next, uint8_t* ttc_mask_out, float* scalars,
                       void* workspace, size_t workspace_bytes, void* stream) {
  return macbf_step_forward_ex(states, goals, obstacles, cbf_params, ctrl_params, B, N, M, top_k, semantics, dist_min_thres,
                               obs_radius, ttc, dt, alpha, 0, idx, idx_next, u, s_next, h, h_next, ttc_mask_out, scalars,
                               workspace, workspace_bytes, stream);
}

int macbf_step_forward_ex(const float* states, const float* goals, const float* obstacles, const float* cbf_params,
                          const float* ctrl_params, int B, int N, int M, int top_k, int semantics, double dist_min_thres,
                          double obs_radius, double ttc, double dt, double alpha, int flags, int32_t* idx,
                          int32_t* idx_next, float* u, float* s_next, float* h, float* h_next, uint8_t* ttc_mask_out,
                          float* scalars, void* workspace, size_t workspace_bytes, void* stream) {
  return macbf_step_forward_train(states, goals, obstacles, cbf_params, ctrl_params, B, N, M, top_k, semantics, dist_min_thres,
                                  obs_radius, ttc, dt, alpha, flags, idx, idx_next, u, s_next, h, h_next, ttc_mask_out, scalars,
                                  workspace, workspace_bytes, nullptr, nullptr, 0, stream);
}

int macbf_step_forward_train(const float* states, const float* goals, const float* obstacles, const float* cbf_params,
                             const float* ctrl_params, int B, int N, int M, int top_k, int semantics, double dist_min_thres,
                             double obs_radius, double ttc, double dt, double alpha, int flags, int32_t* idx,
                             int32_t* idx_next, float* u, float* s_next, float* h, float* h_next, uint8_t* ttc_mask_out,
                             float* scalars, void* workspace, size_t workspace_bytes, void* record, void* record_next,
                             size_t record_bytes, void* stream) {
  const char* fn = "macbf_step_forward";
  MACBF_CHECK_ARG((record == nullptr) == (record_next == nullptr), "%s: record and record_next go together", fn);
  MACBF_CHECK_ARG(record == nullptr || semantics == MACBF_REPAIRED, "%s: records need REPAIRED semantics", fn);
  // CBF forward, plain or recording (macbf_cbf_forward_record)
  auto cbf_fwd = [&](const float* s_, const int32_t* idx_, float* h_, void* rec_, int K_, cudaStream_t st_) -> int {
    if (rec_ != nullptr)
      return cbf_forward_record(s_, obstacles, idx_, cbf_params, B, N, M, K_, dist_min_thres, obs_radius, h_, rec_, record_bytes, st_);
    return cbf_forward(s_, obstacles, idx_, cbf_params, B, N, M, K_, semantics, dist_min_thres, obs_radius, h_, st_);
  };
  MACBF_CHECK_ARG((flags & ~(MACBF_STEP_REUSE_INDICES | MACBF_STEP_SAVE_WINNERS)) == 0, "%s: unknown flags 0x%x", fn, flags);
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  const int K = macbf_neighbor_count(N, M, top_k);
  if (check_dims(fn, B, N, M, K)) return 1;
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, cbf_params); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, idx);
  REQ_PTR(fn, idx_next); REQ_PTR(fn, u); REQ_PTR(fn, s_next); REQ_PTR(fn, h); REQ_PTR(fn, h_next);
  REQ_PTR(fn, ttc_mask_out); REQ_PTR(fn, scalars); REQ_PTR(fn, workspace);
  REQ_A16(fn, states); REQ_A16(fn, s_next); REQ_A16(fn, workspace);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(workspace_bytes >= macbf_step_workspace_bytes(B, N, M, top_k), "%s: workspace too small (%zu < %zu)",
                  fn, workspace_bytes, macbf_step_workspace_bytes(B, N, M, top_k));
  cudaStream_t st = (cudaStream_t)stream;
  float* pool = (float*)workspace;
  void* lscratch = (char*)workspace + align256((size_t)B * N * MACBF_POOL_DIM * sizeof(float));
  const size_t gbytes = step_graph_scratch_bytes(B, N, M, top_k);
  void* gscratch = gbytes ? (char*)lscratch + align256(loss_scratch_bytes(B, N, K)) : nullptr;
  int rc;
  // graph(S') needs positions only and p' = p + v dt does not depend on the controller (core.py:137 with core.py:74-86): where
  // the thread-per-query grid search applies, both graphs come from ONE build + ONE query launch here
  const bool pair = !(flags & MACBF_STEP_REUSE_INDICES) && gscratch != nullptr && neighbor_grid_pair_applicable(N, M, top_k);
  if (pair) rc = neighbor_topk_grid_pair(states, obstacles, B, N, M, K, dt, idx, idx_next, gscratch, gbytes, st);
  else rc = neighbor_any(states, obstacles, B, N, M, top_k, idx, gscratch, gbytes, st);
  if (rc) return rc;
  MACBF_CHECK_ARG(!(flags & MACBF_STEP_SAVE_WINNERS) || controller_forward_is_tensor_core(semantics),
                  "%s: MACBF_STEP_SAVE_WINNERS needs the tensor-core controller (REPAIRED semantics)", fn);
  if (flags & MACBF_STEP_SAVE_WINNERS)
    tl_save_winners = (uint8_t*)workspace + macbf_step_workspace_winners_offset(B, N, M, top_k);
  rc = controller_forward(states, goals, obstacles, idx, ctrl_params, B, N, M, K, semantics, obs_radius, dt, pool, u, s_next, st);
  tl_save_winners = nullptr;
  if (rc) return rc;
  if (flags & MACBF_STEP_REUSE_INDICES) {   // h' on S's graph (the upstream `indices=` argument the reference dropped)
    MACBF_CHECK_CUDA(cudaMemcpyAsync(idx_next, idx, (size_t)B * N * K * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    if ((rc = cbf_fwd(states, idx, h, record, K, st))) return rc;
  } else if (pair) {
    if ((rc = cbf_fwd(states, idx, h, record, K, st))) return rc;
  } else {
    // graph(S') (core.py:137) on the side stream, under CBF(S) on the caller's stream
    SideStream* side = nullptr;
    if ((rc = side_stream_for_current_device(&side))) return rc;
    MACBF_CHECK_CUDA(cudaEventRecord(side->fork, st));
    MACBF_CHECK_CUDA(cudaStreamWaitEvent(side->stream, side->fork, 0));
    if ((rc = neighbor_any(s_next, obstacles, B, N, M, top_k, idx_next, gscratch, gbytes, side->stream))) return rc;
    MACBF_CHECK_CUDA(cudaEventRecord(side->join, side->stream));
    if ((rc = cbf_fwd(states, idx, h, record, K, st))) return rc;
    MACBF_CHECK_CUDA(cudaStreamWaitEvent(st, side->join, 0));
  }
  if ((rc = cbf_fwd(s_next, idx_next, h_next, record_next, K, st))) return rc;
  return loss_forward(states, goals, obstacles, idx, h, h_next, u, B, N, M, K, dist_min_thres, ttc, dt, alpha,
                      ttc_mask_out, scalars, lscratch, loss_scratch_bytes(B, N, K), st);
}

// ---- row-sharded forward step of ONE env ---------------------------------------------------------------------------
struct RowScope {     // sets the thread-local row range for the launchers called inside the scope
  RowScope(int b, int e) { tl_row_range = RowRange{b, e}; }
  ~RowScope() { tl_row_range = RowRange{0, -1}; }
};
static int check_rows(const char* fn, int N, int row_begin, int row_end) {
  MACBF_CHECK_ARG(row_begin >= 0 && row_begin <= row_end && row_end <= N, "%s: row range [%d, %d) outside [0, %d)", fn,
                  row_begin, row_end, N);
  MACBF_CHECK_ARG(row_begin % MACBF_ROW_SHARD_ALIGN == 0 && (row_end == N || row_end % MACBF_ROW_SHARD_ALIGN == 0),
                  "%s: row range bounds must be multiples of %d agents (or N)", fn, MACBF_ROW_SHARD_ALIGN);
  return 0;
}

int macbf_step_forward_rows_a(const float* states, const float* goals, const float* obstacles, const float* ctrl_params,
                              int N, int M, int top_k, int semantics, double obs_radius, double dt, int row_begin,
                              int row_end, int32_t* idx, float* u, float* s_next, void* workspace, size_t workspace_bytes,
                              void* stream) {
  const char* fn = "macbf_step_forward_rows_a";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  const int K = macbf_neighbor_count(N, M, top_k);
  if (check_dims(fn, 1, N, M, K) || check_rows(fn, N, row_begin, row_end)) return 1;
  MACBF_CHECK_ARG(semantics == MACBF_REPAIRED, "%s: REPAIRED semantics only", fn);
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, idx); REQ_PTR(fn, u); REQ_PTR(fn, s_next);
  REQ_PTR(fn, workspace); REQ_A16(fn, states); REQ_A16(fn, s_next); REQ_A16(fn, workspace);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(workspace_bytes >= macbf_step_workspace_bytes(1, N, M, top_k), "%s: workspace too small (%zu < %zu)", fn,
                  workspace_bytes, macbf_step_workspace_bytes(1, N, M, top_k));
  if (row_end == row_begin) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  float* pool = (float*)workspace;
  void* lscratch = (char*)workspace + align256((size_t)N * MACBF_POOL_DIM * sizeof(float));
  const size_t gbytes = macbf_neighbor_workspace_bytes(1, N, M, top_k);
  void* gscratch = gbytes ? (char*)lscratch + align256(loss_scratch_bytes(1, N, K)) : nullptr;
  RowScope scope(row_begin, row_end);
  int rc;
  if ((rc = neighbor_any(states, obstacles, 1, N, M, top_k, idx, gscratch, gbytes, st))) return rc;
  return controller_forward(states, goals, obstacles, idx, ctrl_params, 1, N, M, K, semantics, obs_radius, dt, pool, u, s_next, st);
}

int macbf_step_forward_rows_b(const float* states, const float* s_next, const float* goals, const float* obstacles,
                              const float* cbf_params, int N, int M, int top_k, int semantics, double dist_min_thres,
                              double obs_radius, double ttc, double dt, double alpha, int row_begin, int row_end,
                              const int32_t* idx, int32_t* idx_next, const float* u, float* h, float* h_next,
                              uint8_t* ttc_mask_out, float* loss_partial, void* workspace, size_t workspace_bytes,
                              void* stream) {
  const char* fn = "macbf_step_forward_rows_b";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  const int K = macbf_neighbor_count(N, M, top_k);
  if (check_dims(fn, 1, N, M, K) || check_rows(fn, N, row_begin, row_end)) return 1;
  MACBF_CHECK_ARG(semantics == MACBF_REPAIRED, "%s: REPAIRED semantics only", fn);
  REQ_PTR(fn, states); REQ_PTR(fn, s_next); REQ_PTR(fn, goals); REQ_PTR(fn, cbf_params); REQ_PTR(fn, idx);
  REQ_PTR(fn, idx_next); REQ_PTR(fn, u); REQ_PTR(fn, h); REQ_PTR(fn, h_next); REQ_PTR(fn, ttc_mask_out);
  REQ_PTR(fn, loss_partial); REQ_PTR(fn, workspace); REQ_A16(fn, states); REQ_A16(fn, s_next); REQ_A16(fn, workspace);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(workspace_bytes >= macbf_step_workspace_bytes(1, N, M, top_k), "%s: workspace too small (%zu < %zu)", fn,
                  workspace_bytes, macbf_step_workspace_bytes(1, N, M, top_k));
  cudaStream_t st = (cudaStream_t)stream;
  MACBF_CHECK_CUDA(cudaMemsetAsync(loss_partial, 0, loss_scratch_bytes(1, N, K), st));   // rows of other ranges contribute 0
  if (row_end == row_begin) return 0;
  void* lscratch = (char*)workspace + align256((size_t)N * MACBF_POOL_DIM * sizeof(float));
  const size_t gbytes = macbf_neighbor_workspace_bytes(1, N, M, top_k);
  void* gscratch = gbytes ? (char*)lscratch + align256(loss_scratch_bytes(1, N, K)) : nullptr;
  RowScope scope(row_begin, row_end);
  int rc;
  if ((rc = neighbor_any(s_next, obstacles, 1, N, M, top_k, idx_next, gscratch, gbytes, st))) return rc;   // core.py:137
  if ((rc = cbf_forward(states, obstacles, idx, cbf_params, 1, N, M, K, semantics, dist_min_thres, obs_radius, h, st))) return rc;
  if ((rc = cbf_forward(s_next, obstacles, idx_next, cbf_params, 1, N, M, K, semantics, dist_min_thres, obs_radius, h_next, st)))
    return rc;
  return loss_forward(states, goals, obstacles, idx, h, h_next, u, 1, N, M, K, dist_min_thres, ttc, dt, alpha, ttc_mask_out,
                      nullptr, loss_partial, loss_scratch_bytes(1, N, K), st);
}

int macbf_step_forward_rows_c(const float* loss_partial_sum, int N, int K, float* scalars, void* stream) {
  const char* fn = "macbf_step_forward_rows_c";
  MACBF_CHECK_ARG(N >= 1 && K >= 1 && K <= MACBF_MAX_K, "%s: bad extents N=%d K=%d", fn, N, K);
  REQ_PTR(fn, loss_partial_sum); REQ_PTR(fn, scalars);
  return loss_finalize(loss_partial_sum, 1, N, K, 1, 1, scalars, (cudaStream_t)stream);
}

int macbf_ttc_dense(const float* states, int B, int N, double dist_min_sq, double ttc, int32_t* counts, uint8_t* mask,
                    void* stream) {
  const char* fn = "macbf_ttc_dense";
  MACBF_CHECK_ARG(B >= 1 && N >= 1, "%s: B=%d, N=%d must be >= 1", fn, B, N);
  MACBF_CHECK_ARG(B <= 65535, "%s: B=%d exceeds 65535", fn, B);
  REQ_PTR(fn, states); REQ_A16(fn, states);
  MACBF_CHECK_ARG(counts != nullptr || mask != nullptr, "%s: counts and mask are both NULL", fn);
  return ttc_dense(states, B, N, dist_min_sq, ttc, counts, mask, (cudaStream_t)stream);
}

size_t macbf_scenario_workspace_bytes(int B, int N, int n_candidates) { return scenario_workspace_bytes(B, N, n_candidates); }

int macbf_generate_scenario(unsigned long long seed, int B, int N, int n_candidates, double dist_min_thres, float* states,
                            float* goals, int32_t* accepted, void* workspace, size_t workspace_bytes, void* stream) {
  const char* fn = "macbf_generate_scenario";
  MACBF_CHECK_ARG(B >= 1 && N >= 1, "%s: B=%d, N=%d must be >= 1", fn, B, N);
  REQ_PTR(fn, states); REQ_PTR(fn, goals); REQ_PTR(fn, accepted); REQ_PTR(fn, workspace);
  REQ_A16(fn, states); REQ_A16(fn, workspace);
  return scenario_generate(seed, B, N, n_candidates, dist_min_thres, states, goals, accepted, workspace, workspace_bytes,
                           (cudaStream_t)stream);
}

size_t macbf_rollout_workspace_bytes(int B, int N, int M, int top_k) { return rollout_workspace_bytes(B, N, M, top_k); }

int macbf_rollout_forward(const float* states0, const float* goals, const float* obstacles, const float* ctrl_params,
                          const float* noise, int B, int N, int M, int top_k, int max_steps, int semantics, double obs_radius, double dt,
                          double dist_stop, float* traj_states, float* traj_actions, int32_t* steps_run,
                          void* workspace, size_t workspace_bytes, void* stream) {
  const char* fn = "macbf_rollout_forward";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  if (check_dims(fn, B, N, M, macbf_neighbor_count(N, M, top_k))) return 1;
  MACBF_CHECK_ARG(max_steps >= 1, "%s: max_steps=%d must be >= 1", fn, max_steps);
  REQ_PTR(fn, states0); REQ_PTR(fn, goals); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, traj_states); REQ_PTR(fn, traj_actions);
  REQ_PTR(fn, steps_run); REQ_PTR(fn, workspace); REQ_A16(fn, states0); REQ_A16(fn, traj_states); REQ_A16(fn, workspace);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  MACBF_CHECK_ARG(semantics == MACBF_LITERAL || semantics == MACBF_REPAIRED, "%s: bad semantics %d", fn, semantics);
  return rollout_forward(states0, goals, obstacles, ctrl_params, noise, B, N, M, top_k, max_steps, semantics, obs_radius, dt,
                         dist_stop, traj_states, traj_actions, steps_run, workspace, workspace_bytes, (cudaStream_t)stream);
}

int macbf_generate_obstacles(const int32_t* kinds, const double* centers, const double* sizes, int n_obstacles, int num,
                             double* points, void* stream) {
  const char* fn = "macbf_generate_obstacles";
  MACBF_CHECK_ARG(n_obstacles >= 1 && num >= 1 && (long long)n_obstacles * num < (1ll << 30), "%s: bad extents n_obstacles=%d num=%d",
                  fn, n_obstacles, num);
  REQ_PTR(fn, kinds); REQ_PTR(fn, centers); REQ_PTR(fn, sizes); REQ_PTR(fn, points);
  MACBF_CHECK_ARG(((((uintptr_t)centers) | ((uintptr_t)sizes) | ((uintptr_t)points)) & 7) == 0,
                  "%s: centers, sizes and points must be 8-byte aligned", fn);
  return obstacle_points(kinds, centers, sizes, n_obstacles, num, points, (cudaStream_t)stream);
}

size_t macbf_rollout_backward_scratch_bytes(int B, int N, int M, int top_k) {
  return rollout_backward_scratch_bytes(B, N, M, top_k);
}

int macbf_rollout_backward(const float* traj_states, const float* goals, const float* obstacles, const float* ctrl_params,
                           const float* grad_traj_states, const float* grad_traj_actions, int B, int N, int M, int top_k,
                           int steps, double obs_radius, double dt, float* grad_params, float* grad_states0,
                           float* grad_goals, void* scratch, size_t scratch_bytes, void* stream) {
  const char* fn = "macbf_rollout_backward";
  MACBF_CHECK_ARG(top_k >= 1 && top_k <= MACBF_MAX_K, "%s: top_k=%d out of range [1,%d]", fn, top_k, MACBF_MAX_K);
  if (check_dims(fn, B, N, M, macbf_neighbor_count(N, M, top_k))) return 1;
  MACBF_CHECK_ARG(steps >= 1, "%s: steps=%d must be >= 1", fn, steps);
  REQ_PTR(fn, traj_states); REQ_PTR(fn, goals); REQ_PTR(fn, ctrl_params); REQ_PTR(fn, grad_traj_actions);
  REQ_PTR(fn, grad_params); REQ_PTR(fn, scratch);
  REQ_A16(fn, traj_states); REQ_A16(fn, scratch);
  MACBF_CHECK_ARG(grad_traj_states == nullptr || (((uintptr_t)grad_traj_states) & 15) == 0,
                  "%s: grad_traj_states must be 16-byte aligned", fn);
  MACBF_CHECK_ARG(grad_states0 == nullptr || (((uintptr_t)grad_states0) & 15) == 0, "%s: grad_states0 must be 16-byte aligned", fn);
  MACBF_CHECK_ARG(M == 0 || obstacles != nullptr, "%s: obstacles is NULL but M=%d", fn, M);
  return rollout_backward(traj_states, goals, obstacles, ctrl_params, grad_traj_states, grad_traj_actions, B, N, M, top_k, steps,
                          obs_radius, dt, grad_params, grad_states0, grad_goals, scratch, scratch_bytes, (cudaStream_t)stream);
}

}  // extern "C"
