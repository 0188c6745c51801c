// K3b - backward of the controller (autograd of controller.py:31-63 + the Euler step of
// train.py:70 / core.py:133-134, REPAIRED semantics): entry point.  Recompute-based unless the training forward kept
// the pooled features and the arg-max slots:
//   1. ctrl_edge_tc (ctrl_tc.cu, forward with winners) rebuilds the pooled features,
//   2. ctrl_node_bwd_tc (ctrl_node_tc.cu / ctrl_node_bwd_tc.cu): node MLP forward + backward; emits d/dpool in place
//      of pool, d/dstates and d/dgoals of the (p-g, v) inputs and of the action,
//   3. ctrl_edge_bwd_tc (ctrl_edge_bwd_tc.cu): routes d/dpool to the arg-max slot of every (agent, channel)
//      (torch.max backward; first maximum), backward through 5->64->128,
//   4. reduce_partials_kernel: per-CTA weight-gradient slices -> grad_params, in CTA order.
#include "common.cuh"

namespace macbf {

int reduce_partials(const float* partial, int nslices, int P, float* out, cudaStream_t st);


static inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

bool ctrl_edge_bwd_tc_applicable(int K);
int ctrl_edge_bwd_tc_slices();
size_t ctrl_edge_bwd_tc_scratch_bytes(int B, int N, int K);
uint8_t* ctrl_edge_bwd_tc_argk(void* scratch);
int ctrl_edge_bwd_tc(const float*, const float*, const int32_t*, const float*, const float*, int, int, int, int, float*,
                     float*, void*, cudaStream_t, const uint8_t*);
int ctrl_edge_tc(const float* states, const float* obstacles, const int32_t* idx, const float* params, int B, int N,
                 int M, int K, double obs_radius, float* pool, cudaStream_t st, uint8_t* argk);
size_t ctrl_node_bwd_tc_scratch_bytes(int B, int N);
int ctrl_node_bwd_tc(const float*, const float*, const float*, const float*, const float*, int, int, double, float*, float*,
                     float*, float*, void*, cudaStream_t);
constexpr int CTRL_SLICES = 2 * NUM_SMS;   // the tensor-core edge backward runs two CTAs per SM
static inline size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

// scratch: [pool / dz : B*N*128 f32][partial : CTRL_SLICES * 34052 f32][winners + tile records of the tensor-core edge backward]
size_t controller_backward_scratch_bytes(int B, int N, int K) {
  return a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)) + a256((size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float)) +
         max_sz(ctrl_edge_bwd_tc_scratch_bytes(B, N, K), ((size_t)B * N * 128 + 4096) + ctrl_node_bwd_tc_scratch_bytes(B, N));   // node records sit behind the winners
}

int controller_backward(const float* states, const float* goals, const float* obstacles, const int32_t* idx,
                        const float* params, const float* grad_u, const float* grad_s_next, int B, int N, int M,
                        int K, double obs_radius, double dt, float* grad_params, float* grad_states,
                        float* grad_goals, void* scratch, size_t scratch_bytes, cudaStream_t st, float* pool_saved,
                        const uint8_t* argk_saved) {
  const long long agents = (long long)B * N;
  MACBF_CHECK_ARG(agents < (1ll << 31), "controller_backward: B*N must be < 2^31");
  const size_t need = controller_backward_scratch_bytes(B, N, K);
  MACBF_CHECK_ARG(scratch_bytes >= need, "controller_backward: scratch too small (%zu < %zu)", scratch_bytes, need);
  // pooled features and arg-max slots kept by the training forward (macbf_step_forward_ex with
  // MACBF_STEP_SAVE_WINNERS) make the edge-net recompute unnecessary; pool_saved is consumed (becomes d/dpool)
  const bool saved = pool_saved != nullptr && argk_saved != nullptr;
  float* pool = saved ? pool_saved : (float*)scratch;
  float* partial = (float*)((char*)scratch + a256((size_t)B * N * MACBF_POOL_DIM * sizeof(float)));
  MACBF_CHECK_CUDA(cudaMemsetAsync(partial, 0, (size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float), st));
  MACBF_CHECK_ARG(ctrl_edge_bwd_tc_applicable(K), "controller_backward: K=%d out of range", K);
  void* edge_scratch = (char*)partial + a256((size_t)CTRL_SLICES * MACBF_CTRL_PARAMS * sizeof(float));
  int rc = 0;
  if (!saved) rc = ctrl_edge_tc(states, obstacles, idx, params, B, N, M, K, obs_radius, pool, st, ctrl_edge_bwd_tc_argk(edge_scratch));
  if (rc) return rc;
  // the node records live where the edge backward's winners / records go afterwards - except the winners, which the
  // recomputed forward has already written: keep them (first B*N*128 bytes of edge_scratch, 1024-aligned)
  char* node_scratch = (char*)edge_scratch + (((size_t)agents * 128 + 2047) & ~(size_t)1023);
  rc = ctrl_node_bwd_tc(states, goals, params, grad_u, grad_s_next, B, N, dt, pool, partial, grad_states, grad_goals,
                        node_scratch, st);
  if (rc) return rc;
  rc = ctrl_edge_bwd_tc(states, obstacles, idx, params, pool, B, N, M, K, partial, grad_states, edge_scratch, st,
                        saved ? argk_saved : nullptr);
  if (rc) return rc;
  return reduce_partials(partial, CTRL_SLICES, MACBF_CTRL_PARAMS, grad_params, st);
}

}  // namespace macbf
