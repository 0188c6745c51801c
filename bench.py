#!/usr/bin/env python
"""Benchmark of the per-step safety-certificate path (BASELINE.json metric: agent-steps/s).

A "step" is one pass of the hot path over one batch of synthetic swarms:
    graph(S) -> u, S' -> graph(S') -> h(S), h(S') -> TTC mask -> 9 loss scalars      (forward step)
  + seeds -> CBF/controller backward -> parameter gradients + NCCL all-reduce (N>1)    (training step)
Default workload = BASELINE.json configs[1]: 1024 agents x 64 envs, top-k 12, per GPU.  The headline
(`value`, `e2e`, `roofline`) is the forward step that config names; the same JSON line carries a `train`
block - the training step of the same workload, timed in the same process, with the gradient
all-reduce inside the timed region - and a live per-kernel breakdown of both steps.
One process per GPU; envs are independent, so ranks share nothing on the forward path
("scaling": "weak": every rank runs the same per-GPU batch) and exactly one collective when training.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU algorithm (oracle port) on host cores

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "macbf-gnn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

METRIC = "agent_steps_per_sec"
UNIT = "agent-steps/s"
CBF_EDGE_FLOP, CTRL_EDGE_FLOP, NODE_FLOP = 33664, 17024, 50176


def flops_per_agent(K, mode):
    """CBF edge 33,664, controller edge 17,024, node 50,176 flop."""
    fwd = K * (2 * CBF_EDGE_FLOP + CTRL_EDGE_FLOP) + NODE_FLOP      # two CBF passes (h and h') + controller
    return fwd if mode == "forward" else 3 * fwd


def bytes_per_agent(K, mode):
    """Algorithmic HBM bytes per agent-step: s,g,u,s' + h,h',idx,idx' + mask."""
    fwd = 16 + 8 + 8 + 16 + 4 * K * 4 + K
    return fwd if mode == "forward" else fwd + 36 * K + 16


# Algorithmic work per launch of the kernels a step is made of: (bound, per-edge, per-agent).
# tensor: flop; hbm: bytes.  Used for the live per-kernel table and for the `roofline` objects.
KERNEL_WORK = {
    "cbf_forward_tc_kernel": ("tensor", CBF_EDGE_FLOP, 0),
    "cbf_forward_tc_kernel<dump>": ("tensor", CBF_EDGE_FLOP, 0),
    "cbf_forward_tc_kernel<record>": ("tensor", CBF_EDGE_FLOP, 0),     # training: the forward pass that writes the tile records
    "ctrl_edge_tc_kernel": ("tensor", CTRL_EDGE_FLOP, 0),
    "ctrl_node_tc_kernel": ("tensor", 0, NODE_FLOP),
    "ctrl_node_tc_kernel<dump>": ("tensor", 0, NODE_FLOP),
    # delta chain dA2 = dZ3 W3, dA1 = dZ2 W2 (2 x 16,384) + first/last layer pieces (dZ3, dW1, d/dstates)
    "cbf_dgrad_tc_kernel": ("tensor", 2 * (64 * 128 + 128 * 64) + 2 * 64 + 2 * 2 * 6 * 64, 0),
    # dW3 = A2^T dZ3, dW2 = dZ2^T A1
    "wgrad_cbf_ring_kernel": ("tensor", 2 * (128 * 64 + 64 * 128), 0),
    "wgrad_tc_kernel<ctrl>": ("tensor", 2 * 64 * 128, 0),
    "wgrad_tc_kernel<node>": ("tensor", 0, 2 * (64 * 128 + 128 * 64 + 128 * 64)),
    "ctrl_edge_bwd_tc_kernel": ("tensor", 2 * 128 * 64 + 2 * 2 * 5 * 64, 0),
    "ctrl_node_dgrad_tc_kernel": ("tensor", 0, 2 * (64 * 128 + 128 * 64 + 64 * 132) + 2 * 4 * 64),
    "knn_grid_tpq_kernel": ("hbm", 0, 0), "grid_build_kernel": ("hbm", 0, 0),       # filled in per K below
    "loss_partial_kernel": ("hbm", 4 + 4 + 4 + 1, 16 + 8 + 8),
    "loss_backward_kernel": ("hbm", 4 + 4 + 1 + 4 + 4, 16 + 8 + 8 + 8),
}


def load_json(path):
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f)
    return None


def load_peaks():
    p = load_json(os.path.join(ROOT, "MEASURED_PEAKS.json"))
    if p:
        out = {"hbm_gbs": p.get("hbm_gbs", 6650.0), "bf16_tflops": p.get("bf16_tflops", 1590.0),
               "bf16_tflops_sustained": p.get("bf16_tflops_sustained", 1400.0), "source": "MEASURED_PEAKS.json"}
    else:
        out = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
               "source": "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"}
    # tcgen05 kind::tf32 dense rate of this GPU, measured by tools/peak_probe.cu (committed output); without it the
    # ratio tf32 = bf16 / 2 is an assumption and is reported as such
    probe = load_json(os.path.join(ROOT, "profiles", "r2_peak_probe.json"))
    if probe and probe.get("tf32_mma_tflops"):
        out["tf32_tflops"] = float(probe["tf32_mma_tflops"])
        out["tf32_source"] = "profiles/r2_peak_probe.json (tools/peak_probe.cu: tcgen05.mma kind::tf32 on all SMs)"
    else:
        out["tf32_tflops"] = out["bf16_tflops"] / 2.0
        out["tf32_source"] = "assumed bf16/2 (no probe output committed)"
    return out


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def make_inputs(B, N, M, seed):
    from oracle import macbf_oracle as O   # synthetic-swarm generator shared with the CPU arm (inputs only)
    s, g = O.synthetic_swarm(B, N, seed)
    obs = O.synthetic_obstacles(B, N, M // 24, M // 24, seed) if M > 0 else None
    return s, g, obs


def cpu_arm_step(O, p_cbf, p_ctrl, s, g, obs, k, mode, envs):
    """One bounded sample of the workload on the host cores: `envs` envs through the oracle."""
    for b in range(envs):
        O.full_step(p_cbf, p_ctrl, s[b], g[b], None if obs is None else obs[b], k, O.REPAIRED,
                    with_grad=(mode == "train"))


def cpu_baseline_run(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, mode, envs, seconds, N):
    torch.set_num_threads(os.cpu_count() or 1)
    cpu_arm_step(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, mode, 1)      # warm-up
    t0 = time.perf_counter()
    n_done = 0
    while True:
        cpu_arm_step(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, mode, envs)
        n_done += envs
        if time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return {"value": n_done * N / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n_done} envs x {N} agents of the same workload ({mode} step) in {dt:.1f} s, {envs} envs per pass "
                      f"(oracle port of the reference's CPU algorithm, torch {torch.__version__} CPU "
                      f"{torch.backends.cpu.get_cpu_capability()}; envs are independent, so agent-steps/s does not "
                      f"depend on how many of them are timed)"}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port of core.py/cbf.py/controller.py with the
    Appendix-A repairs; /root/reference itself is Python that does not import and is absent on the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import macbf_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    B, N, k = args.envs, args.agents, args.top_k
    sample_envs = max(1, min(B, args.cpu_envs))
    s, g, obs = make_inputs(sample_envs, N, args.obstacles, 1234 + 2000)
    p_cbf, p_ctrl = O.init_params(5)
    for _ in range(args.warmup):
        cpu_arm_step(O, p_cbf, p_ctrl, s, g, obs, k, args.mode, sample_envs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_arm_step(O, p_cbf, p_ctrl, s, g, obs, k, args.mode, sample_envs)
    dt = time.perf_counter() - t0
    value = sample_envs * N * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "envs_timed_per_step": sample_envs,
                         "sample": f"{sample_envs} of {B} envs x {N} agents per step (envs are independent; "
                                   f"throughput per agent-step does not depend on the env count), torch {torch.__version__} "
                                   f"{torch.backends.cpu.get_cpu_capability()}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(args):
    if (args.agents, args.envs, args.top_k, args.obstacles) == (1024, 64, 12, 0):
        tag = "BASELINE configs[1]"
    elif args.agents == 4096 and args.envs == 1 and args.obstacles > 0:
        tag = "BASELINE configs[2]"
    elif args.agents == 16384:
        tag = "BASELINE configs[3] (per-GPU share of the 256 envs)"
    elif args.agents == 65536 and args.envs == 1:
        tag = "BASELINE configs[4] (one point of the density sweep)"
    else:
        tag = "custom"
    return {"workload": f"{tag}: {args.agents} agents x {args.envs} envs per GPU, top-k {args.top_k}, "
                        f"{args.obstacles} obstacle points, {args.mode} step "
                        "(graph+controller+Euler+graph'+CBF(S)+CBF(S')+TTC+losses)",
            "agents": args.agents, "envs_per_gpu": args.envs, "top_k": args.top_k, "obstacles": args.obstacles,
            "mode": args.mode, "semantics": "REPAIRED", "l2": "flushed between timed steps (256 MiB write)",
            **({"variant": "reuse_indices (h' on S's graph; NOT the reference's default path)"}
               if getattr(args, "reuse_indices", False) else {})}


class Arm:
    """One measured path (the forward step or the training step) with `n_slots` independent sets of device buffers:
    slot 0 serves the device-resident timing, both serve the pipelined end-to-end loop."""

    def __init__(self, mode, args, dev, world, cf, tf, obs_d, n_slots=2):
        from macbf_b200 import ops
        self.mode, self.args, self.cf, self.tf, self.obs_d, self.world = mode, args, cf, tf, obs_d, world
        B, N, M, k = args.envs, args.agents, args.obstacles, args.top_k
        if mode == "train":
            from macbf_b200 import step as step_mod
            self.trainers = [step_mod.TrainStep(B, N, M, k, dev, world_size=world) for _ in range(n_slots)]
            self.bufs = [t.buf for t in self.trainers]
        else:
            self.trainers = None
            self.bufs = [ops.StepBuffers(B, N, M, k, dev) for _ in range(n_slots)]
        self.ops = ops

    def run(self, slot, s, g, all_reduce=True):
        if self.trainers is None:
            self.ops.step_forward(s, g, self.cf, self.tf, self.bufs[slot], self.obs_d, reuse_indices=self.args.reuse_indices)
        else:
            self.trainers[slot].forward_backward(s, g, self.cf, self.tf, self.obs_d, all_reduce=all_reduce,
                                                 reuse_indices=self.args.reuse_indices)


def measure_device(arm, s_d, g_d, steps, warmup, flush, barrier, dist, world, dev, local_rank, ops):
    """`value`: inputs resident in HBM, one CUDA-event pair per step on the launching stream, L2 flushed (untimed)
    between steps, max over ranks.  Training at N>1: the gradient all-reduce is inside the event pair."""
    for _ in range(warmup):
        arm.run(0, s_d, g_d)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ops.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    probes = None
    if arm.trainers is not None:
        probes = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    for i, (a, b) in enumerate(evs):
        flush.zero_()                       # evict the previous step's working set from L2 (untimed)
        if probes is not None:
            arm.trainers[0].ev_backward_done, arm.trainers[0].ev_sync_done = probes[i]
        a.record()
        arm.run(0, s_d, g_d)
        b.record()
    barrier()
    if probes is not None:
        arm.trainers[0].ev_backward_done = arm.trainers[0].ev_sync_done = None
    launches = ops.launch_count() - l0
    clocks = sampler.stop()
    mine = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([mine], dtype=torch.float64, device=dev)
    per_rank = [mine / steps]
    if world > 1:
        gathered = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        per_rank = [float(x.item()) / steps for x in gathered]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    out = {"total_ms": total_ms, "ms_per_step": total_ms / steps, "per_rank_ms_per_step": per_rank,
           "launches": launches, "clocks": clocks}
    if probes is not None:
        # from the last backward kernel to the end of the gradient sync: the part of the all-reduce that is NOT hidden
        # under the controller backward (the CBF half is issued before it)
        ex = sum(p.elapsed_time(q) for p, q in probes) / steps
        t2 = torch.tensor([ex], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        out["allreduce_exposed_ms"] = float(t2.item())
    return out


def measure_e2e(arm, s_pin, g_pin, steps, warmup, flush, barrier, dist, world, dev):
    """`e2e`: the same step through the public API with HOST buffers.  Every step copies its inputs from pinned host
    memory (one block [states | goals]) and reads its result block [scalars | u] back; the loop is software-pipelined
    over two buffer sets and three streams (H2D of step i+1 and D2H of step i-1 run under the kernels of step i).
    The clock is the device time from the first H2D to the last D2H, minus the L2-flush kernels (each bracketed by
    its own event pair on the compute stream, which is the critical path)."""
    ops = arm.ops
    B, N = arm.args.envs, arm.args.agents
    nslots = len(arm.bufs)
    comp = torch.cuda.current_stream()
    s_h2d, s_d2h = torch.cuda.Stream(), torch.cuda.Stream()
    in_pin, in_dev, views, out_pin = [], [], [], []
    for j in range(nslots):
        ip = torch.empty((B * N * 6,), dtype=torch.float32).pin_memory()
        ip[:B * N * 4].copy_(s_pin.reshape(-1))
        ip[B * N * 4:].copy_(g_pin.reshape(-1))
        idv = torch.empty((B * N * 6,), dtype=torch.float32, device=dev)
        in_pin.append(ip)
        in_dev.append(idv)
        views.append((idv[:B * N * 4].view(B, N, 4), idv[B * N * 4:].view(B, N, 2)))
        out_pin.append(torch.empty_like(arm.bufs[j].result, device="cpu").pin_memory())

    def loop(n, timed):
        ev_h2d = [torch.cuda.Event() for _ in range(nslots)]
        ev_comp = [torch.cuda.Event() for _ in range(nslots)]
        ev_d2h = [torch.cuda.Event() for _ in range(nslots)]
        fl = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)] if timed else None
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record(comp)
        s_h2d.wait_event(start)
        s_d2h.wait_event(start)
        for i in range(n):
            j = i % nslots
            with torch.cuda.stream(s_h2d):
                if i >= nslots:
                    s_h2d.wait_event(ev_comp[j])          # the step that last read this input slot is done
                in_dev[j].copy_(in_pin[j], non_blocking=True)
                ev_h2d[j].record(s_h2d)
            if timed:
                fl[i][0].record(comp)
            flush.zero_()
            if timed:
                fl[i][1].record(comp)
            comp.wait_event(ev_h2d[j])
            if i >= nslots:
                comp.wait_event(ev_d2h[j])                # this slot's previous result has been read out
            arm.run(j, *views[j])
            ev_comp[j].record(comp)
            with torch.cuda.stream(s_d2h):
                s_d2h.wait_event(ev_comp[j])
                out_pin[j].copy_(arm.bufs[j].result, non_blocking=True)
                ev_d2h[j].record(s_d2h)
        for j in range(nslots):
            comp.wait_event(ev_d2h[j])
        end.record(comp)
        end.synchronize()
        if not timed:
            return 0.0
        return start.elapsed_time(end) - sum(a.elapsed_time(b) for a, b in fl)

    loop(max(warmup, nslots), False)
    barrier()
    ms = loop(steps, True)
    barrier()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    h2d = in_pin[0].numel() * 4
    d2h = out_pin[0].numel() * 4
    return {"value": world * B * N * steps / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h, "ms_per_step": ms / steps,
            "how": "pipelined over 2 buffer sets / 3 streams; device time first H2D -> last D2H minus the in-stream "
                   "L2-flush kernels; pinned host blocks [states|goals] in, [scalars|u] out, every step"}


def kernel_breakdown(arm, s_d, g_d, flush, reps, edges, agents, peaks):
    """Live per-kernel table of one step: CUDA events after every library launch (macbf_timing_begin/_end), averaged
    over `reps` steps with the L2 flushed before each.  ms = time since the previous launch mark on the stream."""
    ops = arm.ops
    agg, order = {}, []
    for _ in range(reps):
        flush.zero_()
        torch.cuda.synchronize()
        ops.timing_begin()
        arm.run(0, s_d, g_d, all_reduce=False)     # rank 0 only: no collective here (its cost is `allreduce_exposed_ms`)
        for name, ms in ops.timing_end():
            if name not in agg:
                agg[name] = [0, 0.0]
                order.append(name)
            agg[name][0] += 1
            agg[name][1] += ms
    total = sum(v[1] for v in agg.values()) / reps
    table = []
    for name in order:
        n, ms = agg[name]
        row = {"kernel": name, "launches_per_step": n / reps, "ms_per_launch": ms / n, "share": (ms / reps) / total}
        w = KERNEL_WORK.get(name)
        if w and (w[1] or w[2]):
            per_launch = w[1] * edges + w[2] * agents          # every launch covers the whole batch
            if w[0] == "tensor":
                row["tflops"] = per_launch / (ms / n * 1e-3) / 1e12
                row["frac_of_bf16_peak_sustained"] = row["tflops"] / peaks["bf16_tflops_sustained"]
            else:
                row["gbs"] = per_launch / (ms / n * 1e-3) / 1e9
                row["frac_of_hbm_peak"] = row["gbs"] / peaks["hbm_gbs"]
        table.append(row)
    return {"ms_per_step_sum": total, "kernels": table,
            "how": "CUDA event after every library launch on the launching stream; one more event per launch than "
                   "the timed steps have, so the sum is slightly above ms_per_step"}


def run_rowshard(args):
    """--row-shard: ONE large env (BASELINE configs[4], the density sweep's swarm) with its query rows sharded over the
    ranks (macbf_b200/rowshard.py): positions replicated, two small collectives per step (all-gather of s_next, all-reduce
    of the loss partial sums), both inside the timed region.  Strong scaling: the total work is fixed."""
    from macbf_b200 import ops, rowshard
    from oracle import macbf_oracle as O
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    N, k, M = args.agents, args.top_k, args.obstacles
    s_h, g_h, obs_h = make_inputs(1, N, M, 1234 + 5000 + k)          # the same swarm on every rank
    p_cbf, p_ctrl = O.init_params(5)
    cf = O.flatten_params(p_cbf, O.CBF_KEYS).to(dev)
    tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).to(dev)
    s_d, g_d = s_h.to(dev), g_h.to(dev)
    obs_d = obs_h.to(dev) if obs_h is not None else None
    step = rowshard.RowShardedStep(N, M, k, dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step.forward(s_d, g_d, cf, tf, obs_d)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ops.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in evs:
        flush.zero_()
        a.record()
        step.forward(s_d, g_d, cf, tf, obs_d)
        b.record()
    barrier()
    launches = ops.launch_count() - l0
    clocks = sampler.stop()
    mine = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([mine], dtype=torch.float64, device=dev)
    per_rank = [mine / args.steps]
    if world > 1:
        gathered = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        per_rank = [float(x.item()) / args.steps for x in gathered]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    # parity on hardware: every rank's rows against the unsharded step computed locally on rank 0's GPU
    parity = None
    own = {n: getattr(step.buf, n)[0, step.lo:step.hi].clone() for n in ("idx", "idx_next", "u", "h", "h_next")}
    scal = step.buf.scalars.clone()
    ref = ops.StepBuffers(1, N, M, k, dev)
    ops.step_forward(s_d, g_d, cf, tf, ref, obs_d)
    ok = all(torch.equal(own[n], getattr(ref, n)[0, step.lo:step.hi]) for n in own)
    ok = ok and bool(((scal - ref.scalars).abs() <= 1e-6 * ref.scalars.abs().clamp_min(1.0)).all())
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    parity = bool(int(flag.item()))
    if rank == 0:
        value = N * args.steps / (total_ms * 1e-3)
        cfg = workload_config(args)
        cfg["workload"] = (f"BASELINE configs[4] (one point of the density sweep): ONE env of {N} agents, top-k {k}, forward step, "
                           f"query rows sharded over {world} GPU(s)")
        cfg["parallelism"] = f"row-sharded x{world} (positions replicated; all-gather of s_next + all-reduce of loss partial sums per step)"
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": cfg, "gpu_launches": launches, "clocks": clocks,
                "per_rank_ms_per_step": per_rank, "rows_per_rank": [hi - lo for lo, hi in step.ranges],
                "rows_equal_unsharded_step": parity,
                "e2e": None, "roofline": None, "cpu_baseline": None}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="forward", choices=["forward", "train"],
                    help="which step the headline (value/e2e/roofline) describes; with 'forward' (default) the "
                         "training step is timed too and reported in the `train` block")
    ap.add_argument("--agents", type=int, default=1024)
    ap.add_argument("--envs", type=int, default=64, help="envs per GPU")
    ap.add_argument("--top-k", type=int, default=12)
    ap.add_argument("--obstacles", type=int, default=0, help="obstacle points per env (multiple of 24)")
    ap.add_argument("--cpu-envs", type=int, default=4, help="envs per CPU-arm step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train-block", action="store_true")
    ap.add_argument("--row-shard", action="store_true",
                    help="one env (--envs 1) with its query rows sharded over the ranks (SURVEY 8f-4); strong scaling")
    ap.add_argument("--reuse-indices", action="store_true",
                    help="opt-in variant (not the reference's default): h' on S's neighbour sets, one graph search per step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return
    if args.row_shard:
        assert args.envs == 1, "--row-shard is for ONE large env (--envs 1)"
        run_rowshard(args)
        return

    from macbf_b200 import ops
    from oracle import macbf_oracle as O  # parameter init + inputs shared with the CPU arm; checker only

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, N, k, M = args.envs, args.agents, args.top_k, args.obstacles
    K = ops.neighbor_count(N, M, k)
    edges, agents = B * N * K, B * N
    KERNEL_WORK["knn_grid_tpq_kernel"] = ("hbm", 4, 16)       # idx out per edge, state in per agent
    peaks = load_peaks()

    seed0 = 1234 + 2000
    s_h, g_h, obs_h = make_inputs(B, N, M, seed0 + rank)
    p_cbf, p_ctrl = O.init_params(5)
    cf = O.flatten_params(p_cbf, O.CBF_KEYS).to(dev)
    tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).to(dev)
    s_pin, g_pin = s_h.pin_memory(), g_h.pin_memory()
    s_d, g_d = s_h.to(dev), g_h.to(dev)
    obs_d = obs_h.to(dev) if obs_h is not None else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(mode, with_cpu):
        arm = Arm(mode, args, dev, world, cf, tf, obs_d)
        d = measure_device(arm, s_d, g_d, args.steps, args.warmup, flush, barrier, dist, world, dev, local_rank, ops)
        value = world * B * N * args.steps / (d["total_ms"] * 1e-3)
        e2e = measure_e2e(arm, s_pin, g_pin, args.steps, args.warmup, flush, barrier, dist, world, dev)
        res = {"value": value, "ms_per_step": d["ms_per_step"], "per_rank_ms_per_step": d["per_rank_ms_per_step"],
               "e2e": e2e, "gpu_launches": d["launches"], "clocks": d["clocks"]}
        if "allreduce_exposed_ms" in d:
            res["allreduce_exposed_ms"] = d["allreduce_exposed_ms"]
            res["allreduce"] = ("NCCL all-reduce(sum) of the flat gradient inside the timed region: [0,17089) issued "
                                "before the controller backward (overlapped), [17089,51141) after it"
                                if world > 1 else "single GPU: no collective")
        grad_parity = None
        if mode == "train" and world > 1:
            # on-hardware multi-GPU gradient parity: rank 0 recomputes every rank's shard locally
            # with the global-mean seed, sums them, and compares with the all-reduced buffer of the last timed step
            reduced = arm.trainers[0].grad.clone()
            if rank == 0:
                acc = torch.zeros_like(reduced)
                for r in range(world):
                    sr, gr, orr = make_inputs(B, N, M, seed0 + r)
                    _, g_loc = arm.trainers[1].forward_backward(sr.to(dev), gr.to(dev), cf, tf,
                                                                orr.to(dev) if orr is not None else None,
                                                                all_reduce=False, reuse_indices=args.reuse_indices)
                    acc += g_loc
                grad_parity = float((acc - reduced).abs().max() / acc.abs().max().clamp_min(1e-30))
            barrier()
        if grad_parity is not None:
            res["grad_parity_max_rel"] = grad_parity
            res["grad_parity_how"] = (f"max|allreduce(grad) - sum of the {world} shards recomputed on rank 0| / max|grad| "
                                      "(scale-relative; differs only by the order of the fp32 sums)")
        if rank == 0:
            res["kernel_breakdown"] = kernel_breakdown(arm, s_d, g_d, flush, 5, edges, agents, peaks)
            res["step_tflops"] = (value / world) * flops_per_agent(K, mode) / 1e12
            res["step_hbm_frac"] = (value / world) * bytes_per_agent(K, mode) / (peaks["hbm_gbs"] * 1e9)
            if with_cpu and not args.no_cpu_baseline:
                res["cpu_baseline"] = cpu_baseline_run(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, mode,
                                                       max(1, min(B, args.cpu_envs)), 10.0 if mode == "forward" else 6.0, N)
        barrier()
        del arm
        torch.cuda.empty_cache()
        return res

    head = measure(args.mode, True)
    train = None
    if args.mode == "forward" and not args.no_train_block:
        train = measure("train", True)

    # ---------------- roofline of the dominant kernel (CBF edge MLP), timed alone with CUDA events ----------------
    roofline = None
    if rank == 0:
        buf = ops.StepBuffers(B, N, M, k, dev)
        ops.step_forward(s_d, g_d, cf, tf, buf, obs_d)
        reps = max(5, min(20, args.steps))
        for _ in range(3):
            ops.cbf_forward(s_d, buf.idx, cf, obs_d)
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kms = 0.0
        for _ in range(reps):
            flush.zero_()
            k0.record()
            h_tmp = ops.cbf_forward(s_d, buf.idx, cf, obs_d)
            k1.record()
            k1.synchronize()
            kms += k0.elapsed_time(k1)
        kms /= reps
        del h_tmp
        flop = edges * float(CBF_EDGE_FLOP)
        achieved = flop / (kms * 1e-3) / 1e12
        peak = peaks["bf16_tflops"]
        prof = load_json(os.path.join(ROOT, "profiles", "r2_ncu_summary.json")) or {}
        same_workload = prof.get("workload") == [N, B, k, M]
        kprof = (prof.get("kernels", {}).get("cbf_forward_tc_kernel", {}) if same_workload else {})
        share = next((r["share"] for r in head["kernel_breakdown"]["kernels"] if r["kernel"] == "cbf_forward_tc_kernel"),
                     None)
        roofline = {
            "kernel": "cbf_forward_tc_kernel (fused CBF edge MLP 6-64-128-64-1 on tcgen05/TMEM, 3xTF32; 2 launches per "
                      "forward step)",
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture; only
            # quoted when this run's workload is the profiled one
            "traffic": kprof.get("dram_bytes"),
            "traffic_source": ("profiles/r2_ncu_summary.json" if kprof else
                               "null: no ncu capture committed for this workload"),
            "algorithmic_bytes_per_launch": edges * 8 + agents * 16,
            "peak_source": f"bf16 burst, {peaks['source']} (kernel timed alone)",
            "ms_per_launch": kms, "edges_per_launch": edges, "flop_per_edge": CBF_EDGE_FLOP,
            "share_of_step_live": share,
            "note": "fp32 results within 1e-5 are required, so every MMA runs as a 3xTF32 split: the tensor pipe "
                    "executes 3x the algorithmic flop of the two 64x128 layers at the TF32 rate",
            "tf32_mma_tflops_executed": 3 * achieved * (2 * (64 * 128 + 128 * 64)) / float(CBF_EDGE_FLOP),
            "tf32_peak_tflops": peaks["tf32_tflops"], "tf32_peak_source": peaks["tf32_source"],
            "frac_of_3xtf32_ceiling": achieved / (peaks["tf32_tflops"] / 3.0),
            "tensor_pipe_active_pct_ncu": kprof.get("tensor_active_pct"),
            "hbm_gbs_achieved": (edges * 8 + agents * 16) / (kms * 1e-3) / 1e9, "hbm_peak_gbs": peaks["hbm_gbs"],
            "step_hbm_frac": head["step_hbm_frac"], "step_tflops": head["step_tflops"],
        }
        if train is not None or args.mode == "train":
            blk = train if train is not None else head
            # slowest kernel of the backward pass, from the live table (timed inside the step -> sustained peak)
            bwd = [r for r in blk["kernel_breakdown"]["kernels"]
                   if any(t in r["kernel"] for t in ("dgrad", "wgrad", "<dump>", "<record>", "bwd", "backward")) and "tflops" in r]
            if bwd:
                top = max(bwd, key=lambda r: r["ms_per_launch"] * r["launches_per_step"])
                blk["roofline"] = {
                    "kernel": top["kernel"], "bound": "tensor", "achieved": top["tflops"],
                    "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                    "frac": top["tflops"] / peaks["bf16_tflops_sustained"],
                    "frac_of_3xtf32_ceiling": (None if "wgrad" in top["kernel"] else
                                               top["tflops"] / (peaks["tf32_tflops"] / 3.0)),
                    "ms_per_launch": top["ms_per_launch"], "launches_per_step": top["launches_per_step"],
                    "traffic": ((prof.get("kernels", {}).get(top["kernel"], {}).get("dram_bytes")) if same_workload else None),
                    "peak_source": f"bf16 sustained, {peaks['source']} (kernel timed inside the training step)",
                    "note": "algorithmic flop of this kernel per launch (bench.py KERNEL_WORK) / its mean time between "
                            "launch marks; the wgrad kernels run bf16x3 products (hi/lo operand images), the others 3xTF32",
                }

    if rank == 0:
        cpu = head.pop("cpu_baseline", None)
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "clocks": head["clocks"],
            "per_rank_ms_per_step": head["per_rank_ms_per_step"],
            "roofline": roofline, "cpu_baseline": cpu,
            "vs_cpu_baseline": {"device": head["value"] / cpu["value"], "e2e": head["e2e"]["value"] / cpu["value"]} if cpu else None,
            "kernel_breakdown": head["kernel_breakdown"],
        }
        for key in ("allreduce_exposed_ms", "allreduce", "grad_parity_max_rel", "grad_parity_how"):
            if key in head:
                line[key] = head[key]
        if train is not None:
            train["unit"] = UNIT
            train["workload"] = "the same workload, training step: forward + loss seeds + CBF backward on (S', graph') " \
                                "and (S, graph) + controller backward + gradient all-reduce (train.py:88-103)"
            line["train"] = train
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
