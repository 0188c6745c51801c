"""Drop-in for the reference's controller.py: ``Controller(in_dim).forward(states[N,4], goals[N,2])
-> a[N,2]`` with the same ``state_dict`` keys (``controller_centr_net.{0,2}`` Conv1d,
``controller_dec_net.{0,2,4,6}`` Linear) and initialisation order.

The forward (controller.py:31-63) and its autograd run in libmacbf_b200.so: top-k graph, edge
MLP 5->64->128 with the sensing-radius mask and max pooling fused per agent, node MLP
132->64->128->64->4, gains and the PD-like action.  states [B,N,4] / goals [B,N,2] are accepted
as an extension and give a [B,N,2].
"""
import torch
import torch.nn as nn

import core
from config import *  # noqa: F401,F403
from macbf_b200 import functional as _F
from macbf_b200 import ops as _ops


class Controller(nn.Module):

    def __init__(self, in_dim):
        super().__init__()
        if in_dim != 4:
            raise ValueError("the CUDA path is specialised for the double-integrator state (in_dim = 4)")
        # parameter containers, created in the reference's order (controller.py:16-29)
        self.controller_centr_net = nn.Sequential(
            nn.Conv1d(in_dim + 1, 64, (1,)), nn.ReLU(),
            nn.Conv1d(64, 128, (1,)), nn.ReLU())
        self.controller_dec_net = nn.Sequential(
            nn.Linear(128 + 4, 64), nn.ReLU(),
            nn.Linear(64, 128), nn.ReLU(),
            nn.Linear(128, 64), nn.ReLU(),
            nn.Linear(64, 4))

    def _params(self):
        ps = [p for i in (0, 2) for p in (self.controller_centr_net[i].weight, self.controller_centr_net[i].bias)]
        ps += [p for i in (0, 2, 4, 6) for p in (self.controller_dec_net[i].weight, self.controller_dec_net[i].bias)]
        return ps

    def forward(self, states, goals, indices=None):
        sb, squeeze = _F._batched(states, 4)
        gb, _ = _F._batched(goals, 2)
        obs = core._obstacles_for(sb)
        idx = indices if indices is not None else _ops.neighbor_topk(sb.detach(), TOP_K, obs)
        a = _F.ControllerFunction.apply(sb.contiguous(), gb.contiguous(), obs, idx, core.SEMANTICS, *self._params())
        return a[0] if squeeze else a
