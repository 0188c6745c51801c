#!/usr/bin/env python
"""Benchmark of the per-step safety-certificate path (BASELINE.json metric: agent-steps/s).

A "step" is one pass of the hot path over one batch of synthetic swarms:
    graph(S) -> u, S' -> graph(S') -> h(S), h(S') -> TTC mask -> 9 loss scalars      (--mode forward)
  + seeds -> CBF/controller backward -> parameter gradients (+ NCCL all-reduce, N>1)   (--mode train)
Default workload = BASELINE.json configs[1]: 1024 agents x 64 envs, top-k 12, forward, per GPU.
One process per GPU; envs are independent, so ranks share nothing on the forward path
("scaling": "weak": every rank runs the same per-GPU batch).

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


def flops_per_agent(K, mode):
    """CBF edge 33,664, controller edge 17,024, node 50,176 flop."""
    fwd = K * (2 * 33664 + 17024) + 50176          # two CBF passes (h and h') + controller
    return fwd if mode == "forward" else 3 * fwd


def bytes_per_agent(K, mode):
    """Algorithmic HBM bytes per agent-step: s,g,u,s' + h,h',idx,idx' + mask."""
    fwd = 16 + 8 + 8 + 16 + 4 * K * 4 + K
    return fwd if mode == "forward" else fwd + 36 * K + 16


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p.get("hbm_gbs", 6650.0), "bf16_tflops": p.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", 1400.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


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
            time.sleep(0.02)

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
        "config": workload_config(args, sample_envs),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"{sample_envs} of {B} envs x {N} agents per step (envs are independent; "
                                   f"throughput per agent-step does not depend on the env count), torch {torch.__version__} "
                                   f"{torch.backends.cpu.get_cpu_capability()}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(args, envs_timed=None):
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
               if getattr(args, "reuse_indices", False) else {}),
            **({"envs_timed_on_cpu": envs_timed} if envs_timed is not None else {})}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="forward", choices=["forward", "train"])
    ap.add_argument("--agents", type=int, default=1024)
    ap.add_argument("--envs", type=int, default=64, help="envs per GPU")
    ap.add_argument("--top-k", type=int, default=12)
    ap.add_argument("--obstacles", type=int, default=0, help="obstacle points per env (multiple of 24)")
    ap.add_argument("--cpu-envs", type=int, default=4, help="envs per CPU-arm step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reuse-indices", action="store_true",
                    help="opt-in variant (not the reference's default): h' on S's neighbour sets, one graph search per step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return

    from macbf_b200 import ops
    from oracle import macbf_oracle as O  # parameter init + inputs shared with the CPU arm; checker only

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, N, k, M = args.envs, args.agents, args.top_k, args.obstacles
    K = ops.neighbor_count(N, M, k)

    s_h, g_h, obs_h = make_inputs(B, N, M, 1234 + 2000 + rank)
    p_cbf, p_ctrl = O.init_params(5)
    cf = O.flatten_params(p_cbf, O.CBF_KEYS).to(dev)
    tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).to(dev)
    s_pin, g_pin = s_h.pin_memory(), g_h.pin_memory()
    s_d, g_d = s_h.to(dev), g_h.to(dev)
    obs_d = obs_h.to(dev) if obs_h is not None else None
    buf = ops.StepBuffers(B, N, M, k, dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    trainer = None
    if args.mode == "train":
        from macbf_b200 import step as step_mod
        trainer = step_mod.TrainStep(B, N, M, k, dev, world_size=world)

    def one_step(s, g):
        if trainer is None:
            ops.step_forward(s, g, cf, tf, buf, obs_d, reuse_indices=args.reuse_indices)
        else:
            trainer.forward_backward(s, g, cf, tf, obs_d, reuse_indices=args.reuse_indices)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing (value) ----------------
    for _ in range(args.warmup):
        one_step(s_d, g_d)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ops.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in evs:
        flush.zero_()                       # evict the previous step's working set from L2 (untimed)
        a.record()
        one_step(s_d, g_d)
        b.record()
    barrier()
    launches = ops.launch_count() - l0
    clocks = sampler.stop()
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * B * N * args.steps / (total_ms * 1e-3)

    # ---------------- end-to-end through the public API with host buffers (e2e) ----------------
    # host side: one pinned input block [states | goals] and one pinned result block [scalars | u], so a step costs
    # one H2D and one D2H copy; the device views below alias the blocks
    src = buf if trainer is None else trainer.buf
    in_pin = torch.empty((B * N * 6,), dtype=torch.float32).pin_memory()
    in_pin[:B * N * 4].copy_(s_pin.reshape(-1))
    in_pin[B * N * 4:].copy_(g_pin.reshape(-1))
    in_dev = torch.empty((B * N * 6,), dtype=torch.float32, device=dev)
    s_in = in_dev[:B * N * 4].view(B, N, 4)
    g_in = in_dev[B * N * 4:].view(B, N, 2)
    out_pin = torch.empty_like(src.result, device="cpu").pin_memory()
    out_scal = out_pin[:B * ops.NUM_SCALARS].view(B, ops.NUM_SCALARS)
    out_u = out_pin[B * ops.NUM_SCALARS:].view(B, N, 2)

    def e2e_step():
        in_dev.copy_(in_pin, non_blocking=True)
        one_step(s_in, g_in)
        out_pin.copy_(src.result, non_blocking=True)

    for _ in range(args.warmup):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2e_ms = 0.0
    for _ in range(args.steps):
        flush.zero_()
        e0.record()
        e2e_step()
        e1.record()
        e1.synchronize()
        e2e_ms += e0.elapsed_time(e1)
    barrier()
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * B * N * args.steps / (float(t.item()) * 1e-3)
    h2d = s_pin.numel() * 4 + g_pin.numel() * 4
    d2h = out_scal.numel() * 4 + out_u.numel() * 4

    # ---------------- roofline of the dominant kernel (CBF edge MLP), timed live with CUDA events ----------------
    roofline = None
    cpu_baseline = None
    if rank == 0:
        peaks = load_peaks()
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
        edges = B * N * K
        flop = edges * 33664.0
        achieved = flop / (kms * 1e-3) / 1e12
        peak = peaks["bf16_tflops"]
        roofline = {
            "kernel": "cbf_forward_tc_kernel (fused CBF edge MLP 6-64-128-64-1 on tcgen05/TMEM, 3xTF32; 2 launches "
                      "per step, 47% of the step's GPU time in profiles/r1_launches_forward_c2.txt)",
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            # dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full (profiles/r1_ncu_full_forward.txt);
            # algorithmic bytes per launch = edges * 8 (idx in, h out) + states: the 3 MB of h stay in L2
            "traffic": 4311808,
            "peak_source": f"MEASURED_PEAKS.json bf16 burst ({peaks['source']})",
            "ms_per_launch": kms, "edges_per_launch": edges, "flop_per_edge": 33664,
            "note": "fp32 results within 1e-5 are required, so every MMA runs as a 3xTF32 split: the tensor pipe "
                    "executes 3x the algorithmic flop at the TF32 rate (half the bf16 rate); an fp32-exact kernel "
                    "can therefore reach at most 1/6 of the bf16 peak used as the denominator",
            "tf32_mma_tflops_executed": 3 * achieved * (2 * (64 * 128 + 128 * 64)) / 33664.0,
            "frac_of_3xtf32_ceiling": achieved / (peak / 6.0),
            "tensor_pipe_active_pct_ncu_r1": 75.4,
            "hbm_gbs_achieved": edges * (4 + 4) / (kms * 1e-3) / 1e9, "hbm_peak_gbs": peaks["hbm_gbs"],
            "step_hbm_frac": (value / world) * bytes_per_agent(K, args.mode) / (peaks["hbm_gbs"] * 1e9),
            "step_tflops": (value / world) * flops_per_agent(K, args.mode) / 1e12,
        }
        del h_tmp
        if not args.no_cpu_baseline:
            torch.set_num_threads(os.cpu_count() or 1)
            envs = max(1, min(B, args.cpu_envs))
            cpu_arm_step(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, args.mode, 1)      # warm-up
            t0 = time.perf_counter()
            n_done = 0
            while True:
                cpu_arm_step(O, p_cbf, p_ctrl, s_h, g_h, obs_h, k, args.mode, envs)
                n_done += envs
                if time.perf_counter() - t0 > 10.0:
                    break
            dt = time.perf_counter() - t0
            cpu_baseline = {"value": n_done * N / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                            "sample": f"{n_done} envs x {N} agents of the same workload in {dt:.1f} s "
                                      f"(oracle port, torch {torch.__version__} CPU {torch.backends.cpu.get_cpu_capability()})"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
