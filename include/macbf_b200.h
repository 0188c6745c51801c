/*
 * macbf_b200.h - C ABI of libmacbf_b200.so: the B200 (sm_100a) implementation of
 * macbf-gnn's per-step safety-certificate path.
 *
 * The reference (lucaballotta/macbf-gnn) has no FFI: its "operator API" is the Python call
 * surface of core.py / cbf.py / controller.py.  Each entry point below replaces the torch
 * op groups behind one of those functions; the reference lines are cited per function.
 * The Python drop-in modules (macbf-gnn_b200/{core,cbf,controller}.py) bind these through
 * ctypes; INTEGRATION.md shows the stub a maintainer of the reference would add.
 *
 * Conventions
 *  - All pointers are DEVICE pointers unless named host_*; fp32, row-major, contiguous.
 *  - Extents: B independent environments, N agents per env, M obstacle points per env
 *    (0 allowed; `obstacles` may then be NULL), K = neighbours per agent actually stored
 *    (K_eff = min(TOP_K, N + M), see macbf_neighbor_count).
 *  - states  [B,N,4] = (x, y, vx, vy);  goals [B,N,2];  obstacles [B,M,2];
 *    idx [B,N,K] int32, env-local candidate index in [0, N+M): j < N is agent j, j >= N is
 *    obstacle point j-N (candidate state (ox, oy, 0, 0), identity indicator 0).
 *  - `stream` is a cudaStream_t passed as void* (0 = default stream).  Calls enqueue work
 *    and return; nothing synchronises.
 *  - The library allocates nothing: outputs and scratch are caller-owned.
 *  - Return value: 0 on success, non-zero on error; macbf_last_error() gives the text
 *    (thread-local).  No exceptions cross the boundary.
 *  - `semantics`: MACBF_LITERAL reproduces the reference's forwards bug for bug
 *    (reshape-as-transpose cbf.py:40 / controller.py:43, arg-max-index pooling
 *    controller.py:46); MACBF_REPAIRED applies SURVEY Appendix A (F1, F2).
 *  - Parameters are flat fp32 buffers in state_dict order:
 *      CBF  (17089): cbf_net.{0,2,4,6}.{weight,bias}                (cbf.py:13-19)
 *      CTRL (34052): controller_centr_net.{0,2}.{weight,bias},
 *                    controller_dec_net.{0,2,4,6}.{weight,bias}      (controller.py:16-29)
 */
#ifndef MACBF_B200_H
#define MACBF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MACBF_LITERAL 0
#define MACBF_REPAIRED 1

#define MACBF_CBF_PARAMS 17089
#define MACBF_CTRL_PARAMS 34052
#define MACBF_POOL_DIM 128
#define MACBF_MAX_K 128
#define MACBF_NUM_SCALARS 12

/* ---- housekeeping ------------------------------------------------------------------ */
const char* macbf_last_error(void);
int macbf_version(void);
/* K_eff = min(top_k, N + M): core.py:236-237 returns all candidates when there are <= k. */
int macbf_neighbor_count(int N, int M, int top_k);

/* ---- A3: neighbour graph  (core.py:234-250 remove_distant_agents; key core.py:238) ----
 * idx[b,i,:] = the K candidates with the smallest key
 *   sqrt((dx^2 + 1e-6) + (dy^2 + 1e-6)), ascending, ties by ascending candidate index;
 * slot 0 is the agent itself.  If N + M <= top_k: idx[b,i,e] = e (no selection, index order).
 * The dense (N,N,C) tensor of cbf.py:23 / controller.py:33 / core.py:188 is never built. */
int macbf_neighbor_topk(const float* states, const float* obstacles, int B, int N, int M,
                        int top_k, int32_t* idx, void* stream);

/* Same result through an exact uniform-grid search (used when top_k <= 128 and the env has enough candidates,
 * i.e. whenever macbf_neighbor_workspace_bytes is non-zero; otherwise this is the brute-force kernel and needs
 * no workspace).  workspace: macbf_neighbor_workspace_bytes
 * bytes, 16-byte aligned, contents private. */
size_t macbf_neighbor_workspace_bytes(int B, int N, int M, int top_k);
int macbf_neighbor_topk_ws(const float* states, const float* obstacles, int B, int N, int M,
                           int top_k, int32_t* idx, void* workspace, size_t workspace_bytes,
                           void* stream);
/* Row-sharded graph of ONE large env over several GPUs: part `part` of `nparts` computes a contiguous
 * share of the queries (of the cell-sorted order on the grid path, so the shares are spatially compact; of the agent
 * index otherwise) and leaves every other row of idx untouched.  Start from a zeroed idx on every rank and sum the
 * arrays (one all-reduce of B*N*K int32) to obtain the full graph everywhere; positions are replicated. */
int macbf_neighbor_topk_part(const float* states, const float* obstacles, int B, int N, int M, int top_k,
                             int part, int nparts, int32_t* idx, void* workspace, size_t workspace_bytes,
                             void* stream);

/* The reference's own call shape: core.remove_distant_agents(x[N,C,F], k) on an already
 * materialised pair tensor (core.py:234-250); keys from x[i,j,0:2].  idx [N,k] int32 (column
 * indices, ascending key, ties by index).  Requires k < C (core.py:236-237 returns x otherwise). */
int macbf_dense_topk(const float* x, int N, int C, int F, int k, int32_t* idx, void* stream);

/* ---- A2,A4,A5: CBF head  (cbf.py:21-45) ------------------------------------------------
 * h[b,i,k] = mask * MLP_cbf([dx,dy,dvx,dvy,eye, d - dist_min_thres]),
 * d = sqrt((dx^2+1e-4)+(dy^2+1e-4)), mask = d <= obs_radius.  h is [B,N,K] (the reference's
 * (N,1,K) with the unit dim dropped). */
int macbf_cbf_forward(const float* states, const float* obstacles, const int32_t* idx,
                      const float* cbf_params, int B, int N, int M, int K, int semantics,
                      double dist_min_thres, double obs_radius, float* h, void* stream);

/* ---- A6,A7,A8: controller + Euler step  (controller.py:31-63, core.py:74-86, train.py:70)
 * pool [B,N,128]: masked max over the K edge embeddings (LITERAL: arg-max slot as float);
 * u [B,N,2]: action; s_next [B,N,4] = states + [v, u] * dt (may be NULL). */
int macbf_controller_forward(const float* states, const float* goals, const float* obstacles,
                             const int32_t* idx, const float* ctrl_params, int B, int N, int M,
                             int K, int semantics, double obs_radius, double dt, float* pool,
                             float* u, float* s_next, void* stream);

/* ---- A9: time-to-collision mask  (core.py:187-209) -------------------------------------
 * mask[b,i,k] in {0,1}. */
int macbf_ttc_mask(const float* states, const float* obstacles, const int32_t* idx, int B, int N,
                   int M, int K, double dist_min_thres, double ttc, uint8_t* mask, void* stream);

/* ---- A0: scenario generation on the device  (core.py:45-71 generate_data) -------------------
 * The reference's sequential dart throwing (starts in a square of side sqrt(max(1,N/8)), goals within
 * +-0.5 of the starts, all pairs farther apart than dist_min_thres, unplaced rows repelling from the
 * origin) resolved in parallel on a counter-based random stream (seed, env, index); one CTA per env.
 * n_candidates >= N start candidates are drawn per env (2N + 256 is ample at the reference's density);
 * accepted [B] receives how many were accepted: if accepted[b] < N, env b is incomplete - call again with
 * more candidates.  states [B,N,4] (zero velocity), goals [B,N,2].  The stream is NOT numpy's MT19937. */
size_t macbf_scenario_workspace_bytes(int B, int N, int n_candidates);
int macbf_generate_scenario(unsigned long long seed, int B, int N, int n_candidates,
                            double dist_min_thres, float* states, float* goals, int32_t* accepted,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- evaluation-time dense check  (core.py:212-231 ttc_dangerous_mask_np; train.py:74-75) ----
 * For every ordered pair (i, j) of an env: within sqrt(dist_min_sq) or colliding within ttc, in the
 * reference's numpy mixed float32/float64 arithmetic (bit-exact mask).  dist_min_sq is the host's
 * DIST_MIN_CHECK ** 2 (a double).  counts [B,N] = number of dangerous j per row (safety rate of
 * train.py:74 = 1 - counts/N); mask [B,N,N] in {0,1}; either may be NULL, not both. */
int macbf_ttc_dense(const float* states, int B, int N, double dist_min_sq, double ttc,
                    int32_t* counts, uint8_t* mask, void* stream);

/* ---- A10,A11,A12,A13: losses  (core.py:89-184, train.py:93-98) --------------------------
 * scalars [B,12] per env: 0 loss_dang, 1 loss_safe, 2 acc_dang, 3 acc_safe,
 *   4 loss_dang_deriv, 5 loss_safe_deriv, 6 acc_dang_deriv, 7 acc_safe_deriv, 8 loss_action,
 *   9 total = 10*(2*l0 + l1 + 2*l4 + l5 + 0.01*l8), 10 n_dang, 11 n_safe.
 * h_next may be NULL (derivative terms = 0), u/goals may be NULL (action term = 0).
 * ttc_mask [B,N,K] is written (A9 is evaluated inside).  scratch: macbf_loss_scratch_bytes.
 * Alignment (checked; a violation is an error code, nothing is launched): states 16 bytes, goals / u /
 * obstacles 8 bytes. */
size_t macbf_loss_scratch_bytes(int B, int N, int K);
int macbf_loss_forward(const float* states, const float* goals, const float* obstacles,
                       const int32_t* idx, const float* h, const float* h_next, const float* u,
                       int B, int N, int M, int K, double dist_min_thres, double ttc, double dt,
                       double alpha, uint8_t* ttc_mask, float* scalars, void* scratch,
                       size_t scratch_bytes, void* stream);

/* ---- losses without any allocation (no counterpart in the reference) ----------------------
 * Same computation and outputs as macbf_loss_forward, for callers that cannot obtain device
 * scratch: hosts whose allocator is frozen while a step is recorded or replayed, real-time loops
 * that must not enter the driver's allocator, or C / Fortran simulators that call the library
 * without owning a device allocator.  The partial sums live in a fixed 256 KiB array inside the
 * library (present on every GPU that loads it, until the process exits), so the problem must
 * satisfy macbf_loss_scratch_bytes(B,N,K) <= 4 * MACBF_LOSS_NOALLOC_MAX_FLOATS; larger problems
 * are rejected with an error and nothing is launched.  The first call on each device looks up
 * the array's device address (cudaGetSymbolAddress, with that device current); make that first
 * call before starting any stream capture.  All calls on a device share the one array: it keeps
 * the previous call's partial sums (every slot read is rewritten first, nothing needs clearing),
 * and two calls whose kernels overlap on the GPU (different streams, no ordering) race on it.
 * The array is looked up for the CURRENT device: the buffers and the stream must belong to it (a states
 * pointer owned by another device is rejected).  Same argument and alignment checks as macbf_loss_forward. */
#define MACBF_LOSS_NOALLOC_MAX_FLOATS 65536
int macbf_loss_forward_noalloc(const float* states, const float* goals, const float* obstacles,
                               const int32_t* idx, const float* h, const float* h_next, const float* u,
                               int B, int N, int M, int K, double dist_min_thres, double ttc, double dt,
                               double alpha, uint8_t* ttc_mask, float* scalars, void* stream);

/* Seeds of the backward pass: given d(objective)/d(scalars) [B,12] (only entries 0,1,4,5,8 are
 * differentiable; the rest are ignored) produce dL/dh, dL/dh_next [B,N,K], dL/du [B,N,2] and the
 * action-loss part of dL/dstates [B,N,4] (a_ref depends on the states, core.py:177-178); any
 * output may be NULL, all are overwritten.  Uses ttc_mask and scalars[.,10:12] of the forward. */
int macbf_loss_backward(const float* states, const float* goals, const float* h, const float* h_next,
                        const float* u, const uint8_t* ttc_mask, const float* scalars,
                        const float* grad_scalars, int B, int N, int K, double dt, double alpha,
                        float* grad_h, float* grad_h_next, float* grad_u, float* grad_states,
                        void* stream);

/* ---- backward of A5 (autograd of cbf.py:21-45) ------------------------------------------
 * grad_params [17089] is ACCUMULATED into (caller zeroes it); grad_states [B,N,4] is
 * accumulated into when non-NULL (caller zeroes it).  Activations are recomputed from
 * states + idx; nothing from the forward needs to be kept.  REPAIRED semantics only.
 * Scratch holds the per-CTA gradient slices and the tile records the tensor-core kernels hand to each
 * other (bf16 operand images, 141 KB per 128 edges, at most 8192 tiles = 1.1 GB at a time; larger
 * problems are processed in chunks), so ask for its size - it is not small. */
size_t macbf_cbf_backward_scratch_bytes(int B, int N, int K);
int macbf_cbf_backward(const float* states, const float* obstacles, const int32_t* idx,
                       const float* cbf_params, const float* grad_h, int B, int N, int M, int K,
                       double dist_min_thres, double obs_radius, float* grad_params,
                       float* grad_states, void* scratch, size_t scratch_bytes, void* stream);

/* Training steps: instead of discarding the forward pass and recomputing it in macbf_cbf_backward, the forward can WRITE
 * the tile records (macbf_cbf_forward_record: h as macbf_cbf_forward, REPAIRED semantics, tensor-core path) and the
 * backward then starts at the delta chain (macbf_cbf_backward_record: dL/dh -> grad_params accumulated, grad_states
 * accumulated when non-NULL, as macbf_cbf_backward).  One CBF forward per pass less; the price is memory: the record
 * buffer holds ALL tiles of the problem, macbf_cbf_record_bytes = 173 KB per 128 edge rows (1.16 GB for 64 x 1024
 * agents, K = 12), and stays live between the two calls.  The backward only adds to the record (dL/dh-dependent
 * parts), so one record serves any number of backward calls.  scratch: macbf_cbf_backward_record_scratch_bytes()
 * (the per-CTA gradient slices). */
size_t macbf_cbf_record_bytes(int B, int N, int K);
size_t macbf_cbf_backward_record_scratch_bytes(void);
int macbf_cbf_forward_record(const float* states, const float* obstacles, const int32_t* idx,
                             const float* cbf_params, int B, int N, int M, int K, double dist_min_thres,
                             double obs_radius, float* h, void* record, size_t record_bytes, void* stream);
int macbf_cbf_backward_record(const int32_t* idx, const float* cbf_params, const float* grad_h, int B, int N,
                              int K, void* record, size_t record_bytes, float* grad_params, float* grad_states,
                              void* scratch, size_t scratch_bytes, void* stream);

/* ---- backward of A6-A8 (autograd of controller.py:31-63 + Euler step) --------------------
 * grad_u [B,N,2] and (optional) grad_s_next [B,N,4] are the incoming gradients; grad_params
 * [34052] and (optional) grad_states [B,N,4], grad_goals [B,N,2] are accumulated into.
 * Scratch: pooled features / d-pool (512 B per agent), gradient slices, arg-max slots (128 B per agent)
 * and tile records of the edge backward (96 KB per 128 edges, at most 16384 tiles at a time). */
size_t macbf_controller_backward_scratch_bytes(int B, int N, int K);
int macbf_controller_backward(const float* states, const float* goals, const float* obstacles,
                              const int32_t* idx, const float* ctrl_params, const float* grad_u,
                              const float* grad_s_next, int B, int N, int M, int K,
                              double obs_radius, double dt, float* grad_params, float* grad_states,
                              float* grad_goals, void* scratch, size_t scratch_bytes, void* stream);
/* Same, for a training step whose forward kept the pooled features and arg-max slots (macbf_step_forward_ex with
 * MACBF_STEP_SAVE_WINNERS: pool at offset 0 and the slots at macbf_step_workspace_winners_offset() of its
 * workspace): the edge-net recompute is skipped.  pool_saved is CONSUMED (it is overwritten by d/dpool). */
int macbf_controller_backward_saved(const float* states, const float* goals, const float* obstacles,
                                    const int32_t* idx, const float* ctrl_params, const float* grad_u,
                                    const float* grad_s_next, int B, int N, int M, int K, double obs_radius,
                                    double dt, float* pool_saved, const uint8_t* winners_saved,
                                    float* grad_params, float* grad_states, float* grad_goals, void* scratch,
                                    size_t scratch_bytes, void* stream);

/* ---- the whole per-step path in one call (what bench.py times) ---------------------------
 * graph(S) -> u, S' -> graph(S') -> h(S), h(S') -> TTC mask -> scalars.  Workspace layout is
 * private; size it with macbf_step_workspace_bytes.  Outputs u, s_next, h, h_next, idx,
 * idx_next, ttc_mask, scalars as documented above (idx_next/h_next/ttc_mask may not be NULL).
 * graph(S') needs positions only and p' = p + v dt does not depend on the controller, so where the thread-per-query grid
 * search applies both graphs come from one build + one query launch at the start of the call.  Otherwise the neighbour
 * search on S' runs on a second, library-owned stream under CBF(S) (fork / join by events on `stream`, so the call stays
 * ordered on `stream` and can be captured); that stream and its two events are created on the calling thread's first
 * such call per device and kept (the one exception to "the library holds no resources"). */
size_t macbf_step_workspace_bytes(int B, int N, int M, int top_k);
/* flags for macbf_step_forward_ex.  MACBF_STEP_REUSE_INDICES: h' = CBF(S') is evaluated on the neighbour
 * sets of S instead of a fresh top-k on S' (the `indices=` argument of the code the reference derives
 * from; the reference itself recomputes, core.py:137, which stays the default): slot k of h and h' then
 * refer to the same neighbour and one of the two graph searches is saved.  idx_next receives a copy of idx. */
#define MACBF_STEP_REUSE_INDICES 1
/* MACBF_STEP_SAVE_WINNERS: the controller forward also stores, per (agent, pooled channel), the neighbour slot that
 * won the max (uint8, 255 = the maximum is 0) at macbf_step_workspace_winners_offset() of the workspace, for
 * macbf_controller_backward_saved.  Tensor-core (REPAIRED) path only. */
#define MACBF_STEP_SAVE_WINNERS 2
size_t macbf_step_workspace_winners_offset(int B, int N, int M, int top_k);
int macbf_step_forward(const float* states, const float* goals, const float* obstacles,
                       const float* cbf_params, const float* ctrl_params, int B, int N, int M,
                       int top_k, int semantics, double dist_min_thres, double obs_radius, double ttc,
                       double dt, double alpha, int32_t* idx, int32_t* idx_next, float* u,
                       float* s_next, float* h, float* h_next, uint8_t* ttc_mask, float* scalars,
                       void* workspace, size_t workspace_bytes, void* stream);
int macbf_step_forward_ex(const float* states, const float* goals, const float* obstacles,
                       const float* cbf_params, const float* ctrl_params, int B, int N, int M,
                       int top_k, int semantics, double dist_min_thres, double obs_radius, double ttc,
                       double dt, double alpha, int flags, int32_t* idx, int32_t* idx_next, float* u,
                       float* s_next, float* h, float* h_next, uint8_t* ttc_mask, float* scalars,
                       void* workspace, size_t workspace_bytes, void* stream);

/* macbf_step_forward_ex for a training step: with record / record_next non-NULL (each macbf_cbf_record_bytes(B, N, K)
 * bytes, K = macbf_neighbor_count) the two CBF passes are recording forwards (macbf_cbf_forward_record) and
 * macbf_cbf_backward_record(idx, .., record) / (idx_next, .., record_next) replace the two macbf_cbf_backward calls.
 * Both NULL: identical to macbf_step_forward_ex. */
int macbf_step_forward_train(const float* states, const float* goals, const float* obstacles,
                             const float* cbf_params, const float* ctrl_params, int B, int N, int M,
                             int top_k, int semantics, double dist_min_thres, double obs_radius, double ttc,
                             double dt, double alpha, int flags, int32_t* idx, int32_t* idx_next, float* u,
                             float* s_next, float* h, float* h_next, uint8_t* ttc_mask, float* scalars,
                             void* workspace, size_t workspace_bytes, void* record, void* record_next,
                             size_t record_bytes, void* stream);

/* ---- A0: obstacle point sets on the device (core.py:7-42 generate_obstacle_circle / generate_obstacle_rectangle) ----
 * kinds [n] (0 = circle, 1 = rectangle), centers [n,2], sizes [n,2] ((radius, unused) or the side lengths (a, b)),
 * all on the device; points [n * num, 2] float64 receives `num` points per obstacle in the reference's order (circle:
 * angles 2 pi i / num; rectangle: clockwise from the top-left corner with the reference's per-side counts), computed
 * in float64 in numpy's operation order.  Cast to float32 for the `obstacles` argument of the other entry points. */
int macbf_generate_obstacles(const int32_t* kinds, const double* centers, const double* sizes, int n_obstacles,
                             int num, double* points, void* stream);

/* ---- closed-loop rollout (train.py:59-81) with the stop test on the device ------------------
 * For t = 0 .. max_steps-1: record s_t; u_t = Controller(s_t, goals); s_{t+1} = s_t + [v,u_t] dt; record
 * u_t; an env stops after the first t with mean_i |p_{t+1} - g| < dist_stop (train.py:78-81).
 * traj_states [max_steps+1, B,N,4] (slot 0 = states0), traj_actions [max_steps, B,N,2],
 * steps_run [B] = number of recorded steps per env (slots beyond it hold the continued simulation of
 * that env while others were still running, or are untouched once every env has stopped).
 * noise: NULL, or [max_steps, B,N,2] added to u_t before the Euler step (train.py:66-68; the host draws
 * it, all-zero rows = no noise at that step).
 * No host synchronisation: once all envs have stopped the remaining kernels return at once. */
size_t macbf_rollout_workspace_bytes(int B, int N, int M, int top_k);
int macbf_rollout_forward(const float* states0, const float* goals, const float* obstacles,
                          const float* ctrl_params, const float* noise, int B, int N, int M, int top_k,
                          int max_steps,
                          int semantics, double obs_radius, double dt, double dist_stop,
                          float* traj_states, float* traj_actions, int32_t* steps_run, void* workspace,
                          size_t workspace_bytes, void* stream);

/* Backward through the rollout.  The reference does not detach its trajectory (train.py:59-81 builds the visited states
 * and actions inside autograd, train.py:88-100 backpropagates the losses through all of them), so the gradient of a
 * loss L(s_0 .. s_{T-1}, a_0 .. a_{T-1}) runs back through every controller call and Euler step:
 *   G_T = 0;  t = T-1 .. 0:  graph(s_t), controller backward at s_t with d/du = dL/da_t + dt G_{t+1}[v],
 *   G_t = dL/ds_t + (d u_t / d s_t)^T d/du + the Euler step's direct terms.
 * traj_states [>= steps, B,N,4] as written by macbf_rollout_forward; grad_traj_states [steps, B,N,4] (may be NULL = 0)
 * and grad_traj_actions [steps, B,N,2] are dL/ds_t, dL/da_t and must be ZERO in the slots of an env beyond its
 * steps_run (they are not part of its trajectory).  grad_params [34052] and grad_goals [B,N,2] (optional) are
 * accumulated into; grad_states0 [B,N,4] (optional) is overwritten with dL/ds_0.  Graphs and activations are
 * recomputed per step; nothing but the trajectory is kept from the forward.  REPAIRED semantics.  No host sync. */
size_t macbf_rollout_backward_scratch_bytes(int B, int N, int M, int top_k);
int macbf_rollout_backward(const float* traj_states, const float* goals, const float* obstacles,
                           const float* ctrl_params, const float* grad_traj_states,
                           const float* grad_traj_actions, int B, int N, int M, int top_k, int steps,
                           double obs_radius, double dt, float* grad_params, float* grad_states0,
                           float* grad_goals, void* scratch, size_t scratch_bytes, void* stream);

/* ---- 8f-4: forward step of ONE large env row-sharded over several GPUs (no counterpart in the reference: its dense
 * (N,N,C) tensors of cbf.py:23-30 / controller.py:33 / core.py:188 do not exist at the sizes where this matters) --------
 * The step of macbf_step_forward for B = 1, REPAIRED semantics, restricted to the agents [row_begin, row_end) as
 * queries; every agent (and obstacle point) stays a candidate, so `states` (and, for _b, the complete `s_next`) are the
 * full arrays on every rank, and all output arrays are full-size with only the rows of the range written.  Range
 * bounds are multiples of MACBF_ROW_SHARD_ALIGN agents (or N), which keeps tiles and reduction chunks whole.
 *   _a  graph(S) rows -> idx; controller -> u, s_next rows                          (core.py:234-250, controller.py:31-63)
 *       then the ranks exchange their s_next rows (one all-gather of 16 B per agent)
 *   _b  graph(S') rows -> idx_next; CBF(S), CBF(S') -> h, h_next rows; TTC mask rows; the loss reduction's partial
 *       sums of the rows into loss_partial (macbf_loss_scratch_bytes(1, N, K) bytes, zeroed here first)
 *       then the ranks add their loss_partial arrays (one all-reduce, sum)
 *   _c  summed partial sums -> scalars[12]                                          (core.py:89-184, train.py:93-98)
 * Every row's values are those of the unsharded step bit for bit; the scalars differ from it only by the order of the
 * fp32 additions of the action term.  workspace: macbf_step_workspace_bytes(1, N, M, top_k).  Forward only. */
#define MACBF_ROW_SHARD_ALIGN 1024
int macbf_step_forward_rows_a(const float* states, const float* goals, const float* obstacles,
                              const float* ctrl_params, int N, int M, int top_k, int semantics, double obs_radius,
                              double dt, int row_begin, int row_end, int32_t* idx, float* u, float* s_next,
                              void* workspace, size_t workspace_bytes, void* stream);
int macbf_step_forward_rows_b(const float* states, const float* s_next, const float* goals, const float* obstacles,
                              const float* cbf_params, int N, int M, int top_k, int semantics, double dist_min_thres,
                              double obs_radius, double ttc, double dt, double alpha, int row_begin, int row_end,
                              const int32_t* idx, int32_t* idx_next, const float* u, float* h, float* h_next,
                              uint8_t* ttc_mask, float* loss_partial, void* workspace, size_t workspace_bytes,
                              void* stream);
int macbf_step_forward_rows_c(const float* loss_partial_sum, int N, int K, float* scalars, void* stream);

/* Kernel launches issued by this library on the calling thread since load (bench.py reports
 * the per-step delta as gpu_launches). */
long long macbf_launch_count(void);

/* Per-launch device timing, a measurement aid (no counterpart in the reference; bench.py's live per-kernel
 * breakdown).  Between macbf_timing_begin(stream) and macbf_timing_end every kernel this library launches from the
 * calling thread is followed by a CUDA event on `stream` (pass the stream the calls use).  macbf_timing_end waits
 * for the last event and returns, per launch in launch order, its name (NUL-terminated, name_stride bytes apiece)
 * and the milliseconds since the previous mark - on one stream with back-to-back launches that is the kernel's
 * duration plus whatever other work the caller put on the stream in between.  Returns minus the number of marks
 * written (0 = none), or a positive error code.  At most 4096 launches are recorded; events are created on first
 * use and kept.  Not for use inside a timed region: every mark costs an event record. */
int macbf_timing_begin(void* stream);
int macbf_timing_end(int max_marks, char* names, int name_stride, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* MACBF_B200_H */
