"""Drop-in for the reference's cbf.py: ``CBF(in_dim).forward(states[N,4]) -> h[N,1,K]`` with the
same ``state_dict`` keys (``cbf_net.{0,2,4,6}.{weight,bias}``, Conv1d weights (out,in,1)) and the
same parameter initialisation order, so a reference checkpoint or seed carries over.

The forward (cbf.py:21-45) and its autograd run in libmacbf_b200.so: neighbour search, feature
assembly, radius mask and the 6->64->128->64->1 edge MLP are one fused pass per call; the dense
(N,N,6) tensor of cbf.py:23-30 is never built.  A leading env dim is accepted as an extension:
states [B,N,4] -> h [B,N,1,K].
"""
import torch
import torch.nn as nn

import core
from config import *  # noqa: F401,F403
from macbf_b200 import functional as _F
from macbf_b200 import ops as _ops


class CBF(nn.Module):

    def __init__(self, in_dim: int):
        super().__init__()
        if in_dim != 4:
            raise ValueError("the CUDA path is specialised for the double-integrator state (in_dim = 4)")
        widths = [in_dim + 2, 64, 128, 64, 1]
        layers = []
        for i in range(4):
            layers.append(nn.Conv1d(widths[i], widths[i + 1], (1,)))   # parameter container only
            if i < 3:
                layers.append(nn.ReLU())
        self.cbf_net = nn.Sequential(*layers)

    def _params(self):
        return [p for i in (0, 2, 4, 6) for p in (self.cbf_net[i].weight, self.cbf_net[i].bias)]

    def forward(self, states, indices=None):
        sb, squeeze = _F._batched(states, 4)
        obs = core._obstacles_for(sb)
        idx = indices if indices is not None else _ops.neighbor_topk(sb.detach(), TOP_K, obs)
        h = _F.CBFFunction.apply_auto(sb.contiguous(), obs, idx, core.SEMANTICS, *self._params())
        h = h.unsqueeze(2)                      # [B,N,1,K]: the reference's channels-first (N,1,K)
        return h[0] if squeeze else h
