"""Wide&Deep CTR model (Criteo-shaped: 13 dense features, 26 sparse slots).

wide:  sum_s w_wide[id_s] + dense . w_d
deep:  concat(emb[id_1..id_26], dense) -> 400 -> 400 -> 400 -> 1   (ReLU)
logit = wide + deep; loss = binary cross-entropy with logits.

The sparse parameters (w_wide, emb) are a `ShardedEmbedding` (parallel/ps.py);
this module holds the dense part only.  Reference parity: the workload of
deploy/examples/wide_and_deep*.yaml (Paddle's demo-wide-and-deep image); the
reference contains no model code.
"""
from __future__ import annotations

import torch
import torch.nn as nn

DENSE_FEATURES = 13
SPARSE_SLOTS = 26


class WideDeepDense(nn.Module):
    def __init__(self, emb_dim: int = 16, hidden=(400, 400, 400), slots: int = SPARSE_SLOTS,
                 dense: int = DENSE_FEATURES, fused: bool = False):
        super().__init__()
        self.in_dim = slots * emb_dim + dense
        dims = [self.in_dim] + list(hidden)
        if fused:
            from ..ops.gemm import FusedLinear

            # hidden layers on the tcgen05 kernel with bias + ReLU fused into the epilogue
            self.layers = nn.ModuleList([FusedLinear(a, b, relu=True) for a, b in zip(dims[:-1], dims[1:])])
        else:
            self.layers = nn.ModuleList([nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:])])
        self.fused = fused
        self.out = nn.Linear(dims[-1], 1)
        self.wide_dense = nn.Linear(dense, 1)
        for l in list(self.layers) + [self.out, self.wide_dense]:
            nn.init.normal_(l.weight, std=1.0 / (l.in_features ** 0.5))
            nn.init.zeros_(l.bias)

    def forward(self, deep_emb: torch.Tensor, wide_sum: torch.Tensor, dense: torch.Tensor) -> torch.Tensor:
        """deep_emb [B, slots*emb_dim], wide_sum [B] fp32, dense [B, 13] -> logits [B] fp32."""
        x = torch.cat([deep_emb, dense.to(deep_emb.dtype)], dim=1)
        for l in self.layers:
            x = l(x) if self.fused else torch.relu(l(x))
        deep = self.out(x).squeeze(1).float()
        wide = wide_sum + self.wide_dense(dense.to(self.wide_dense.weight.dtype)).squeeze(1).float()
        return deep + wide


def bce_with_logits(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    return torch.nn.functional.binary_cross_entropy_with_logits(logits, labels.float())
