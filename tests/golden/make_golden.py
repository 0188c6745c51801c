"""Generate the committed golden fixtures under tests/golden/.

Run in the BUILD container only (needs /root/reference):  python tests/golden/make_golden.py

Two families:
* ``ref_*.npz``  - outputs of the UNMODIFIED reference modules (imported through
  oracle/ref_loader.py import stubs) on seeded inputs: LITERAL CBF.forward
  (cbf.py:21-45), Controller.forward (controller.py:31-63), remove_distant_agents
  (core.py:234-250), dynamics (core.py:74-86), ttc_dangerous_mask_np (core.py:212-231),
  generate_data (core.py:45-71).  These pin the oracle restatement.
* ``oracle_*.npz`` - outputs of oracle/macbf_oracle.py (REPAIRED full step incl. parameter
  gradients) so the GPU tests do not depend on the CPU torch build of the GPU box.
* ``ref_repaired_*.npz`` - outputs of the reference source with the Appendix-A repairs applied as the
  committed list of exact text substitutions in oracle/repairs.py (each must match the stated number of
  times) and executed under the same stubs: controller / CBF forwards, TTC mask (core.py:187-209), the 9
  loss scalars (core.py:89-184), the total (train.py:93,98) and its autograd parameter gradients.  These pin
  the REPAIRED restatement of the oracle for the rows whose reference code crashes as shipped.
While generating, the script asserts that the oracle's LITERAL restatement reproduces the
reference bit for bit, and that its REPAIRED restatement reproduces the patched reference
(masks bit for bit, floats within 1e-6 scale-relative).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import macbf_oracle as O  # noqa: E402
from oracle import ref_loader  # noqa: E402
from oracle import repairs  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def sd_to_np(sd):
    return {k: v.detach().numpy().copy() for k, v in sd.items()}


def ref_case(name, core, cbf_mod, ctrl_mod, states, goals, seed):
    torch.manual_seed(seed)
    cbf_net = cbf_mod.CBF(4)
    ctrl_net = ctrl_mod.Controller(4)
    p_cbf = {k: v.detach() for k, v in cbf_net.state_dict().items()}
    p_ctrl = {k: v.detach() for k, v in ctrl_net.state_dict().items()}
    with torch.no_grad():
        h = cbf_net(states)
        u = ctrl_net(states, goals)
        x = states.unsqueeze(1) - states.unsqueeze(0)
        x = torch.cat([x, torch.eye(states.shape[0]).unsqueeze(2)], dim=2)
        xk, ind = core.remove_distant_agents(x, 12)
        dsdt = core.dynamics(states, u)
        ttc_np = core.ttc_dangerous_mask_np(states.numpy())
    # pin: oracle LITERAL == reference, bit for bit
    h_o = O.cbf_forward(p_cbf, states, O.LITERAL)
    u_o = O.controller_forward(p_ctrl, states, goals, O.LITERAL)
    idx_o = O.neighbor_indices(states, 12)
    assert torch.equal(h_o, h), name
    assert torch.equal(u_o, u), name
    assert torch.equal(idx_o.reshape(-1), ind[:, 1]), name
    assert np.array_equal(O.ttc_dangerous_mask_dense_np(states.numpy()), ttc_np), name
    assert torch.equal(O.dynamics(states, u), dsdt)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        states=states.numpy(), goals=goals.numpy(), h=h.numpy(), u=u.numpy(),
        idx=ind.numpy(), xk=xk.numpy(), dsdt=dsdt.numpy(), ttc_np=ttc_np,
        **{"cbf/" + k: v for k, v in sd_to_np(p_cbf).items()},
        **{"ctrl/" + k: v for k, v in sd_to_np(p_ctrl).items()},
    )
    print("wrote", name, "h", tuple(h.shape), "u", tuple(u.shape))


def repaired_case(name, states, goals, p_cbf, p_ctrl, k, out):
    """Run the mechanically repaired reference (oracle/repairs.py) on the inputs of an oracle case and pin the
    oracle's REPAIRED outputs `out` (O.full_step) to it."""
    mods = repairs.load_repaired()
    core, cbf_mod, ctrl_mod, _ = mods
    repairs.set_top_k(mods, k)
    cbf_net, ctrl_net = cbf_mod.CBF(4), ctrl_mod.Controller(4)
    cbf_net.load_state_dict(p_cbf)
    ctrl_net.load_state_dict(p_ctrl)
    a = ctrl_net(states, goals)                                               # train.py:64
    h = cbf_net(states)                                                       # train.py:88
    lb = core.loss_barrier(h, states)                                         # train.py:89
    ld = core.loss_derivatives(states, a, h, cbf_net)                         # train.py:90-91
    la = core.loss_actions(states, goals, a)                                  # train.py:92
    total = 10 * sum([2 * lb[0], lb[1], 2 * ld[0], ld[1], 0.01 * la])         # train.py:93,98 (repair T1)
    total.backward()                                                          # train.py:103
    mask = core.ttc_dangerous_mask(states)
    scalars = torch.stack([torch.as_tensor(x, dtype=torch.float32).detach().reshape(()) for x in (*lb, *ld, la)])
    grad_cbf = torch.cat([cbf_net.state_dict(keep_vars=True)[n].grad.reshape(-1) for n in O.CBF_KEYS])
    grad_ctrl = torch.cat([ctrl_net.state_dict(keep_vars=True)[n].grad.reshape(-1) for n in O.CTRL_KEYS])

    def close(x, y, what):
        x, y = x.detach().double(), y.detach().double()
        err = float((x - y).abs().max() / y.abs().max().clamp_min(1e-30))
        assert err <= 1e-6, (name, what, err)
        return err
    assert torch.equal(mask.reshape(-1), out["ttc_mask"].reshape(-1)), name
    errs = {w: close(x, y, w) for w, x, y in (
        ("u", out["u"], a), ("h", out["h"], h), ("scalars", out["scalars"], scalars), ("total", out["total"], total),
        ("grad_cbf", out["grad_cbf"], grad_cbf), ("grad_ctrl", out["grad_ctrl"], grad_ctrl))}
    np.savez_compressed(
        os.path.join(OUT, "ref_repaired_" + name + ".npz"),
        k=np.int32(k), u=a.detach().numpy(), h=h.detach().numpy(), ttc_mask=mask.numpy(), scalars=scalars.numpy(),
        total=total.detach().numpy(), grad_cbf=grad_cbf.numpy(), grad_ctrl=grad_ctrl.numpy())
    print("wrote ref_repaired_" + name, "oracle vs patched reference:", {w: f"{e:.1e}" for w, e in errs.items()})


def oracle_case(name, n, m_circ, m_rect, k, seed, vel_std=0.3, generate_data=None):
    p_cbf, p_ctrl = O.init_params(seed)
    if generate_data is not None:
        np.random.seed(0)
        s, g = generate_data(n, 0.07)
        states, goals = torch.from_numpy(s), torch.from_numpy(g)
        # give the agents a velocity so the TTC branch is exercised (reference starts at v=0)
        gen = torch.Generator().manual_seed(seed)
        states[:, 2:] = torch.randn(n, 2, generator=gen) * vel_std
    else:
        s, g = O.synthetic_swarm(1, n, seed, vel_std)
        states, goals = s[0], g[0]
    obstacles = None
    if m_circ + m_rect > 0:
        obstacles = O.synthetic_obstacles(1, n, m_circ, m_rect, seed)[0]
    out = O.full_step(p_cbf, p_ctrl, states, goals, obstacles, k, O.REPAIRED, with_grad=True)
    idx = O.neighbor_indices(states, k, obstacles)
    idx_next = O.neighbor_indices(out["s_next"], k, obstacles)
    dec = O.decidable_rows(states, k, obstacles) if idx.shape[1] == k and states.shape[0] + (0 if obstacles is None else obstacles.shape[0]) > k else torch.ones(n, dtype=torch.bool)
    # LITERAL forwards too (restatement), for the CUDA LITERAL mode
    lit = {}
    if states.shape[0] + (0 if obstacles is None else obstacles.shape[0]) > k:
        lit["h_literal"] = O.cbf_forward(p_cbf, states, O.LITERAL, obstacles, k).numpy()
        lit["u_literal"] = O.controller_forward(p_ctrl, states, goals, O.LITERAL, obstacles, k).numpy()
        lit["pool_literal"] = O.controller_pool(p_ctrl, states, O.LITERAL, obstacles, k).numpy()
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        states=states.numpy(), goals=goals.numpy(),
        obstacles=(obstacles.numpy() if obstacles is not None else np.zeros((0, 2), np.float32)),
        k=np.int32(k), idx=idx.numpy().astype(np.int32), idx_next=idx_next.numpy().astype(np.int32),
        decidable=dec.numpy(),
        u=out["u"].numpy(), h=out["h"].numpy(), h_next=out["h_next"].numpy(), s_next=out["s_next"].numpy(),
        ttc_mask=out["ttc_mask"].numpy(), scalars=out["scalars"].numpy(), total=out["total"].numpy(),
        grad_cbf=out["grad_cbf"].numpy(), grad_ctrl=out["grad_ctrl"].numpy(),
        pool=O.controller_pool(p_ctrl, states, O.REPAIRED, obstacles, k).numpy(),
        cbf_flat=O.flatten_params(p_cbf, O.CBF_KEYS).numpy(), ctrl_flat=O.flatten_params(p_ctrl, O.CTRL_KEYS).numpy(),
        **lit,
    )
    print("wrote", name, "scalars", out["scalars"].numpy().round(5), "total", float(out["total"]),
          "n_dang", int(out["ttc_mask"].sum()), "decidable", float(dec.float().mean()))
    if obstacles is None:       # obstacles are an extension without reference semantics
        repaired_case(name.replace("oracle_", ""), states, goals, p_cbf, p_ctrl, k, out)


sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from helpers import OBSTACLE_SPECS  # noqa: E402  (tests/helpers.py)


def obstacle_case(core):
    """core.py:7-42 of the unmodified reference on OBSTACLE_SPECS (12 points each, and 7 / 30 for the first two)."""
    out = {}
    for i, (kind, center, size) in enumerate(OBSTACLE_SPECS):
        fn = core.generate_obstacle_circle if kind == "circle" else core.generate_obstacle_rectangle
        out[f"p{i}"] = fn(center, size)
    out["circle_num7"] = core.generate_obstacle_circle((1.0, 2.0), 0.3, num=7)
    out["rect_num30"] = core.generate_obstacle_rectangle((0.0, 0.0), (0.6, 0.4), num=30)
    np.savez_compressed(os.path.join(OUT, "ref_obstacles.npz"), **out)
    print("ref_obstacles:", {k: v.shape for k, v in out.items()})


def main():
    if "--obstacles-only" in sys.argv:
        obstacle_case(ref_loader.load()[0])
        return
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    print("torch", torch.__version__, "capability", torch.backends.cpu.get_cpu_capability())
    core, cbf_mod, ctrl_mod, _ = ref_loader.load()

    # --- reference family ---
    np.random.seed(0)
    s, g = core.generate_data(32, 0.07)           # config 1 input (train.py:51), v = 0
    ref_case("ref_c1_n32", core, cbf_mod, ctrl_mod, torch.from_numpy(s), torch.from_numpy(g), seed=0)
    s, g = O.synthetic_swarm(1, 200, 11)
    ref_case("ref_n200_vel", core, cbf_mod, ctrl_mod, s[0], g[0], seed=1)
    s, g = O.synthetic_swarm(1, 13, 12)            # smallest N for which the reference's top-k path runs
    ref_case("ref_n13_vel", core, cbf_mod, ctrl_mod, s[0], g[0], seed=2)
    np.random.seed(3)
    s, g = core.generate_data(64, 0.07)
    np.savez_compressed(os.path.join(OUT, "ref_generate_data_n64_seed3.npz"), states=s, goals=g)

    obstacle_case(core)

    # --- oracle family (REPAIRED, with gradients) ---
    oracle_case("oracle_c1_n32", 32, 0, 0, 12, seed=100, generate_data=core.generate_data)
    oracle_case("oracle_n96_obs", 96, 2, 2, 12, seed=101)
    oracle_case("oracle_n8_small", 8, 0, 0, 12, seed=102)          # N <= K: no top-k, K_eff = N
    oracle_case("oracle_n300_k16", 300, 0, 0, 16, seed=103)
    oracle_case("oracle_n257_k40", 257, 1, 1, 40, seed=104)        # two register slots per lane in the warp top-k
    oracle_case("oracle_n1024", 1024, 0, 0, 12, seed=105)


if __name__ == "__main__":
    main()
