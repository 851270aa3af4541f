"""Wide&Deep CTR model (Criteo-shaped: 13 dense features, 26 sparse slots).

wide:  sum_s w_wide[id_s] + dense . w_d
deep:  concat(emb[id_1..id_26], dense) -> 400 -> 400 -> 400 -> 1   (ReLU)
logit = wide + deep; loss = binary cross-entropy with logits.

The sparse parameters (w_wide, emb) are a `ShardedEmbedding` (parallel/ps.py);
this module holds the dense part only.  With `fused=True` the hidden tower is
`ops.mlp.FusedMLP` (tcgen05 fused-linear kernels forward and backward) and the
concatenated input is zero-padded to a multiple of 8 columns so that every
operand row is 16-byte aligned for TMA.  Reference parity: the workload of
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
        self.fused = fused
        self.raw_in = slots * emb_dim + dense
        self.in_dim = (self.raw_in + 7) // 8 * 8 if fused else self.raw_in
        dims = [self.in_dim] + list(hidden)
        if fused:
            from ..ops.mlp import FusedMLP

            self.tower = FusedMLP(dims)
        else:
            self.tower = _TorchMLP(dims)
        self.out = nn.Linear(dims[-1], 1)
        self.wide_dense = nn.Linear(dense, 1)
        for l in list(self.tower.layers) + [self.out, self.wide_dense]:
            nn.init.normal_(l.weight, std=1.0 / (l.in_features ** 0.5))
            nn.init.zeros_(l.bias)

    @property
    def layers(self):
        return self.tower.layers

    def _tower_input(self, deep_emb: torch.Tensor, dense: torch.Tensor) -> torch.Tensor:
        if deep_emb.shape[1] == self.in_dim and self.in_dim != self.raw_in - dense.shape[1]:
            return deep_emb  # the lookup kernel already wrote [embeddings | dense | zero pad] (ShardedEmbedding.set_output_layout)
        parts = [deep_emb, dense.to(deep_emb.dtype)]
        if self.in_dim != self.raw_in:
            parts.append(torch.zeros((deep_emb.shape[0], self.in_dim - self.raw_in), dtype=deep_emb.dtype,
                                     device=deep_emb.device))
        return torch.cat(parts, dim=1)

    def loss(self, deep_emb: torch.Tensor, wide_sum: torch.Tensor, dense: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Mean BCE-with-logits of forward(...) against labels.  On CUDA bf16 with the fused tower the output
        layer, the wide part and the loss run as one kernel (ops/wd_head.py), forward and backward."""
        from ..ops.wd_head import head_bce_loss

        x_in = self._tower_input(deep_emb, dense)
        fused_head = (self.fused and x_in.is_cuda and x_in.dtype == torch.bfloat16
                      and self.out.weight.dtype == torch.bfloat16 and self.out.in_features % 8 == 0)
        if not fused_head:
            return bce_with_logits(self.forward(deep_emb, wide_sum, dense), labels)
        h = self.tower(x_in, top_grad_masked=True)
        return head_bce_loss(h, self.out.weight, self.out.bias, wide_sum, dense, self.wide_dense.weight,
                             self.wide_dense.bias, labels)

    def forward(self, deep_emb: torch.Tensor, wide_sum: torch.Tensor, dense: torch.Tensor) -> torch.Tensor:
        """deep_emb [B, slots*emb_dim], wide_sum [B] fp32, dense [B, 13] -> logits [B] fp32."""
        x = self.tower(self._tower_input(deep_emb, dense))
        deep = self.out(x).squeeze(1).float()
        wide = wide_sum + self.wide_dense(dense.to(self.wide_dense.weight.dtype)).squeeze(1).float()
        return deep + wide


class _TorchMLP(nn.Module):
    def __init__(self, dims):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:])])

    def forward(self, x):
        for l in self.layers:
            x = torch.relu(l(x))
        return x


def bce_with_logits(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    return torch.nn.functional.binary_cross_entropy_with_logits(logits, labels.float())
