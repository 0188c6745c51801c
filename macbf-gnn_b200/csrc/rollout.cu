// Closed-loop rollout (train.py:59-81) without a host round trip per step: T x (neighbour graph ->
// controller -> Euler step) are enqueued back to back; a tiny kernel evaluates the reference's
// stop test (mean distance to goal < DIST_MIN_CHECK) per env on the device, latches the step count
// of every env, and raises a device flag once ALL envs have stopped - the kernels of the remaining
// steps see the flag and return immediately.
#include "common.cuh"

namespace macbf {

int neighbor_topk(const float*, const float*, int, int, int, int, int32_t*, cudaStream_t);
size_t neighbor_grid_workspace_bytes(int, int, int);
bool neighbor_grid_applicable(int, int, int);
int neighbor_topk_grid(const float*, const float*, int, int, int, int, int32_t*, void*, size_t, cudaStream_t);
int controller_forward(const float*, const float*, const float*, const int32_t*, const float*, int, int, int, int,
                       int, double, double, float*, float*, float*, cudaStream_t);

// ctl: [0] all_done flag, [1] number of envs still running, [2..2+B) done[b]
__global__ void rollout_init_kernel(int B, int max_steps, int* __restrict__ ctl, int32_t* __restrict__ steps_run) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t == 0) { ctl[0] = 0; ctl[1] = B; }
  if (t < B) { ctl[2 + t] = 0; steps_run[t] = max_steps; }
}

// one CTA per env: train.py:78-81 on the state AFTER step t
__global__ void __launch_bounds__(256)
rollout_stop_kernel(const float* __restrict__ s_next, const float* __restrict__ goals, int N, int t, float dist_stop,
                    int* __restrict__ ctl, int32_t* __restrict__ steps_run) {
  if (ctl[0] != 0) return;
  const int b = blockIdx.x;
  float acc = 0.f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    const float4 s = __ldg(reinterpret_cast<const float4*>(s_next) + (size_t)b * N + i);
    const float2 g = __ldg(reinterpret_cast<const float2*>(goals) + (size_t)b * N + i);
    const float dx = xsub(s.x, g.x), dy = xsub(s.y, g.y);
    acc += __fsqrt_rn(__fmaf_rn(dy, dy, xmul(dx, dx)));          // torch.norm over 2 elements
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[w];
    if (ctl[2 + b] == 0 && s / (float)N < dist_stop) {
      ctl[2 + b] = 1;
      steps_run[b] = t + 1;                                      // steps recorded for this env (train.py breaks after appending)
      if (atomicSub(&ctl[1], 1) == 1) {
        __threadfence();
        ctl[0] = 1;
      }
    }
  }
}

// train.py:66-68: a_t += noise_t (host-generated, zero rows where the coin said "no noise"), and the
// velocity half of s_{t+1} recomputed from the perturbed action with the same two ops.
__global__ void rollout_noise_kernel(const float* __restrict__ s_t, const float* __restrict__ noise, long long agents,
                                     float dt, float* __restrict__ a_t, float* __restrict__ s_next,
                                     const int* __restrict__ skip) {
  MACBF_SKIP_IF_SET(skip);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= agents) return;
  const float2 nz = __ldg(reinterpret_cast<const float2*>(noise) + i);
  if (nz.x == 0.f && nz.y == 0.f) return;
  float2 a = reinterpret_cast<float2*>(a_t)[i];
  a.x = xadd(a.x, nz.x); a.y = xadd(a.y, nz.y);
  reinterpret_cast<float2*>(a_t)[i] = a;
  const float4 s = __ldg(reinterpret_cast<const float4*>(s_t) + i);
  float4 n = reinterpret_cast<float4*>(s_next)[i];
  n.z = xadd(s.z, xmul(a.x, dt)); n.w = xadd(s.w, xmul(a.y, dt));
  reinterpret_cast<float4*>(s_next)[i] = n;
}

static inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

// workspace: [ctl ints][idx B*N*K][pool B*N*128 f32][grid scratch]
size_t rollout_workspace_bytes(int B, int N, int M, int top_k) {
  const int K = (N + M) < top_k ? (N + M) : top_k;
  size_t g = neighbor_grid_applicable(N, M, top_k) ? neighbor_grid_workspace_bytes(B, N, M) : 0;
  return a256((size_t)(2 + B) * sizeof(int)) + a256((size_t)B * N * K * sizeof(int32_t)) +
         a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)) + a256(g);
}

int rollout_forward(const float* states0, const float* goals, const float* obstacles, const float* ctrl_params,
                    const float* noise, int B,
                    int N, int M, int top_k, int max_steps, int semantics, double obs_radius, double dt, double dist_stop,
                    float* traj_states, float* traj_actions, int32_t* steps_run, void* workspace, size_t workspace_bytes,
                    cudaStream_t st) {
  const int K = (N + M) < top_k ? (N + M) : top_k;
  MACBF_CHECK_ARG(workspace_bytes >= rollout_workspace_bytes(B, N, M, top_k), "rollout_forward: workspace too small (%zu < %zu)",
                  workspace_bytes, rollout_workspace_bytes(B, N, M, top_k));
  char* w = (char*)workspace;
  int* ctl = (int*)w; w += a256((size_t)(2 + B) * sizeof(int));
  int32_t* idx = (int32_t*)w; w += a256((size_t)B * N * K * sizeof(int32_t));
  float* pool = (float*)w; w += a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float));
  void* gscratch = neighbor_grid_applicable(N, M, top_k) ? (void*)w : nullptr;
  const size_t gbytes = gscratch ? neighbor_grid_workspace_bytes(B, N, M) : 0;
  const size_t sbytes = (size_t)B * N * 4 * sizeof(float);

  rollout_init_kernel<<<(B + 255) / 256, 256, 0, st>>>(B, max_steps, ctl, steps_run);
  MACBF_CHECK_LAUNCH("rollout_init_kernel");
  MACBF_CHECK_CUDA(cudaMemcpyAsync(traj_states, states0, sbytes, cudaMemcpyDeviceToDevice, st));
  int rc = 0;
  tl_skip_flag = ctl;
  for (int t = 0; t < max_steps && rc == 0; ++t) {
    const float* s_t = traj_states + (size_t)t * B * N * 4;
    float* s_n = traj_states + (size_t)(t + 1) * B * N * 4;
    float* a_t = traj_actions + (size_t)t * B * N * 2;
    rc = gscratch ? neighbor_topk_grid(s_t, obstacles, B, N, M, K, idx, gscratch, gbytes, st)
                  : neighbor_topk(s_t, obstacles, B, N, M, top_k, idx, st);
    if (rc) break;
    rc = controller_forward(s_t, goals, obstacles, idx, ctrl_params, B, N, M, K, semantics, obs_radius, dt, pool, a_t, s_n, st);
    if (rc) break;
    if (noise != nullptr) {
      const long long agents = (long long)B * N;
      rollout_noise_kernel<<<(unsigned)((agents + 255) / 256), 256, 0, st>>>(s_t, noise + (size_t)t * agents * 2, agents,
                                                                            (float)dt, a_t, s_n, ctl);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) { set_error("rollout_noise_kernel: launch failed: %s", cudaGetErrorString(e)); rc = 2; break; }
      count_launch();
    }
    rollout_stop_kernel<<<B, 256, 0, st>>>(s_n, goals, N, t, (float)dist_stop, ctl, steps_run);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("rollout_stop_kernel: launch failed: %s", cudaGetErrorString(e)); rc = 2; break; }
    count_launch();
  }
  tl_skip_flag = nullptr;
  return rc;
}

// ---- backward through the rollout (the reference's trajectory is NOT detached, train.py:59-81: autograd runs back
// through every controller call and Euler step) ------------------------------------------------------------------
// Given dL/ds_t (t = 0 .. T-1, optional) and dL/da_t, with s_{t+1} = s_t + [v_t, a_t] dt and a_t = u_t (+ noise):
//   G_T = 0;  for t = T-1 .. 0:  graph(s_t);  controller backward at s_t with d/du = dL/da_t + dt G_{t+1}[v]
//   (macbf_controller_backward's grad_s_next argument) -> parameter gradients (accumulated), G_t = dL/ds_t + d u/d s_t
//   + the Euler step's direct terms.  The trajectory is the only thing kept from the forward; graphs and activations
//   are recomputed step by step, all enqueued back to back on the stream.
size_t controller_backward_scratch_bytes(int B, int N, int K);
int controller_backward(const float*, const float*, const float*, const int32_t*, const float*, const float*, const float*,
                        int, int, int, int, double, double, float*, float*, float*, void*, size_t, cudaStream_t, float*,
                        const uint8_t*);

// scratch: [idx B*N*K][G ping][G pong][grid scratch][controller backward scratch]
size_t rollout_backward_scratch_bytes(int B, int N, int M, int top_k) {
  const int K = (N + M) < top_k ? (N + M) : top_k;
  const size_t g = neighbor_grid_applicable(N, M, top_k) ? neighbor_grid_workspace_bytes(B, N, M) : 0;
  return a256((size_t)B * N * K * sizeof(int32_t)) + 2 * a256((size_t)B * N * 4 * sizeof(float)) + a256(g) +
         a256(controller_backward_scratch_bytes(B, N, K));
}

int rollout_backward(const float* traj_states, const float* goals, const float* obstacles, const float* ctrl_params,
                     const float* grad_traj_states, const float* grad_traj_actions, int B, int N, int M, int top_k, int steps,
                     double obs_radius, double dt, float* grad_params, float* grad_states0, float* grad_goals, void* scratch,
                     size_t scratch_bytes, cudaStream_t st) {
  const int K = (N + M) < top_k ? (N + M) : top_k;
  MACBF_CHECK_ARG(scratch_bytes >= rollout_backward_scratch_bytes(B, N, M, top_k), "rollout_backward: scratch too small (%zu < %zu)",
                  scratch_bytes, rollout_backward_scratch_bytes(B, N, M, top_k));
  const size_t sbytes = (size_t)B * N * 4 * sizeof(float);
  char* w = (char*)scratch;
  int32_t* idx = (int32_t*)w; w += a256((size_t)B * N * K * sizeof(int32_t));
  float* G[2];
  G[0] = (float*)w; w += a256(sbytes);
  G[1] = (float*)w; w += a256(sbytes);
  const size_t gbytes = neighbor_grid_applicable(N, M, top_k) ? neighbor_grid_workspace_bytes(B, N, M) : 0;
  void* gscratch = gbytes ? (void*)w : nullptr;
  w += a256(gbytes);
  void* cscratch = (void*)w;
  const size_t cbytes = controller_backward_scratch_bytes(B, N, K);
  int rc = 0;
  for (int t = steps - 1; t >= 0 && rc == 0; --t) {
    const float* s_t = traj_states + (size_t)t * B * N * 4;
    float* g_t = G[t & 1];
    const float* g_next = (t == steps - 1) ? nullptr : G[(t + 1) & 1];      // G_T = 0: the loss does not see s_T
    if (grad_traj_states != nullptr) {
      MACBF_CHECK_CUDA(cudaMemcpyAsync(g_t, grad_traj_states + (size_t)t * B * N * 4, sbytes, cudaMemcpyDeviceToDevice, st));
    } else {
      MACBF_CHECK_CUDA(cudaMemsetAsync(g_t, 0, sbytes, st));
    }
    rc = gscratch ? neighbor_topk_grid(s_t, obstacles, B, N, M, K, idx, gscratch, gbytes, st)
                  : neighbor_topk(s_t, obstacles, B, N, M, top_k, idx, st);
    if (rc) break;
    rc = controller_backward(s_t, goals, obstacles, idx, ctrl_params, grad_traj_actions + (size_t)t * B * N * 2, g_next, B, N,
                             M, K, obs_radius, dt, grad_params, g_t, grad_goals, cscratch, cbytes, st, nullptr, nullptr);
  }
  if (rc == 0 && grad_states0 != nullptr)
    MACBF_CHECK_CUDA(cudaMemcpyAsync(grad_states0, G[0], sbytes, cudaMemcpyDeviceToDevice, st));
  return rc;
}

}  // namespace macbf
