"""Pure-PyTorch CPU port of HashGrid + MLP + L2 + Adam.  BASELINE/TEST INFRASTRUCTURE ONLY (bench.py cpu_baseline).

BASELINE.json: "a pure-PyTorch CPU implementation of the same HashGrid+MLP, timed on the box's host cores ... a
reported baseline, not the optimisation target".  Same index math as the reference (common_device.h:690-718,
825-868; grid.h:94-163), fp32 arithmetic, autograd for the backward pass, torch.optim.Adam for the update."""
from __future__ import annotations

import torch

from . import tcnn_oracle as O

_PRIMES = [1, 2654435761, 805459861, 3674653429, 2097192037, 1434869437, 2165219737]
_M32 = 0xFFFFFFFF


class HashGridMLP(torch.nn.Module):
    def __init__(self, grid: O.GridConfig, mlp: O.MlpConfig, params_fp32):
        super().__init__()
        self.g, self.m = grid, mlp
        p = torch.as_tensor(params_fp32, dtype=torch.float32)
        ws, pos = [], 0
        for (r, c) in mlp.shapes():
            ws.append(torch.nn.Parameter(p[pos:pos + r * c].reshape(r, c).clone()))
            pos += r * c
        self.weights = torch.nn.ParameterList(ws)
        self.table = torch.nn.Parameter(p[pos:].reshape(-1, grid.n_features_per_level).clone())

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        g = self.g
        D = g.n_dims
        feats = []
        for l in range(g.n_levels):
            scale, res, size, base = float(g.scales[l]), int(g.resolutions[l]), g.level_size(l), int(g.offsets[l])
            p = x * scale + 0.5
            fl = torch.floor(p)
            frac = p - fl
            cell = fl.to(torch.int64) & _M32
            # level-constant part of grid_index
            strides, stride = [], 1
            for d in range(D):
                if stride > size:
                    break
                strides.append(stride)
                stride = (stride * res) & _M32
            hashed = size < stride
            acc = 0
            for c in range(1 << D):
                w = torch.ones(x.shape[0], dtype=torch.float32)
                idx = torch.zeros(x.shape[0], dtype=torch.int64)
                for d in range(D):
                    bit = (c >> d) & 1
                    w = w * (frac[:, d] if bit else (1 - frac[:, d]))
                    pg = (cell[:, d] + bit) & _M32
                    if hashed:
                        idx = idx ^ ((pg * _PRIMES[d]) & _M32)
                    elif d < len(strides):
                        idx = (idx + pg * strides[d]) & _M32
                idx = idx % size + base
                acc = acc + w[:, None] * self.table[idx]
            feats.append(acc)
        enc = torch.cat(feats, dim=1)
        pad = self.m.n_input - enc.shape[1]
        return torch.nn.functional.pad(enc, [0, pad]) if pad else enc

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = self.encode(x)
        for w in self.weights[:-1]:
            h = torch.relu(h @ w.T)
        return (h @ self.weights[-1].T)[:, : self.m.n_output]


def make_trainer(grid: O.GridConfig, mlp: O.MlpConfig, params_fp32, adam: O.AdamConfig):
    model = HashGridMLP(grid, mlp, params_fp32)
    opt = torch.optim.Adam(model.parameters(), lr=adam.learning_rate, betas=(adam.beta1, adam.beta2), eps=adam.epsilon)

    def step(x: torch.Tensor, target: torch.Tensor) -> float:
        opt.zero_grad(set_to_none=True)
        loss = ((model(x) - target) ** 2).mean()
        loss.backward()
        opt.step()
        return float(loss.detach())

    return model, step
