"""Building blocks with the reference's class names and call surface.

``MLP``, ``Head``, ``MultiHeadAttention`` and ``Block`` mirror reference
``src/models/mlp.py:5-67``, ``attention.py:6-96`` and ``transformer_block.py:6-61``: same
constructor arguments, same ``forward`` / ``forward_embedding`` / ``project_embedding``
methods, same math.  The differences are structural: attention keeps ONE packed
``wqkv`` (3d, d) weight instead of 3*H small Linears and no ``tril`` buffer (the causal mask
is implicit in the kernel / SDPA), and heads run batched instead of in a Python loop.

These ``forward`` methods are the plain-PyTorch statement of the math.  They are what runs
on CPU, what the GPU numerics tests use as the same-layout oracle, and what a user gets when
calling a sub-module directly.  The training / generation hot path on a B200 does not go
through them: ``Transformer`` dispatches whole-model work to ``runtime.engine`` which drives
the hand-written sm_100a kernels over the same flat parameter buffer.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class MLP(nn.Module):
    """Linear(d,4d) -> ReLU -> Linear(4d,d), both with bias (reference mlp.py:24-26)."""

    def __init__(self, n_embed: int) -> None:
        super().__init__()
        self.hidden = nn.Linear(n_embed, 4 * n_embed)
        self.relu = nn.ReLU()
        self.proj = nn.Linear(4 * n_embed, n_embed)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.project_embedding(self.forward_embedding(x))

    def forward_embedding(self, x: torch.Tensor) -> torch.Tensor:
        """Post-ReLU hidden activation (reference mlp.py:43-53)."""
        return self.relu(self.hidden(x))

    def project_embedding(self, x: torch.Tensor) -> torch.Tensor:
        """Down projection of a hidden activation (reference mlp.py:55-66)."""
        return self.proj(x)


class Head(nn.Module):
    """One causal attention head with its own key/query/value Linears.

    Kept for API parity with reference attention.py:6-58 (users can build and call a single
    head).  ``Transformer`` does not instantiate it; its heads live packed in
    ``MultiHeadAttention.wqkv``.  No ``tril`` buffer is registered: the mask is implicit.
    """

    def __init__(self, head_size: int, n_embed: int, context_length: int) -> None:
        super().__init__()
        self.key = nn.Linear(n_embed, head_size, bias=False)
        self.query = nn.Linear(n_embed, head_size, bias=False)
        self.value = nn.Linear(n_embed, head_size, bias=False)
        self.context_length = context_length

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        k, q, v = self.key(x), self.query(x), self.value(x)
        return F.scaled_dot_product_attention(q, k, v, is_causal=True)


class MultiHeadAttention(nn.Module):
    """Causal multi-head attention, packed QKV, no output projection (reference
    attention.py:60-96: heads are concatenated and returned as they are)."""

    def __init__(self, n_head: int, n_embed: int, context_length: int) -> None:
        super().__init__()
        assert n_embed % n_head == 0
        self.n_head, self.n_embed, self.context_length = n_head, n_embed, context_length
        self.head_size = n_embed // n_head
        w = torch.empty(3 * n_embed, n_embed)
        for h in range(n_head):                      # reference RNG order: key, query, value
            for third in (1, 0, 2):                  # packed thirds are [query; key; value]
                r0 = third * n_embed + h * self.head_size
                nn.init.kaiming_uniform_(w[r0:r0 + self.head_size], a=math.sqrt(5))
        self.wqkv = nn.Parameter(w)

    def head_weights(self, h: int):
        """(key, query, value) weights of head ``h`` as views, each (head_size, n_embed)."""
        d, hs = self.n_embed, self.head_size
        w = self.wqkv
        return (w[d + h * hs:d + (h + 1) * hs], w[h * hs:(h + 1) * hs],
                w[2 * d + h * hs:2 * d + (h + 1) * hs])

    def qkv(self, x: torch.Tensor):
        B, T, _ = x.shape
        qkv = F.linear(x, self.wqkv)                                    # (B,T,3d)
        q, k, v = qkv.view(B, T, 3, self.n_head, self.head_size).permute(2, 0, 3, 1, 4)
        return q, k, v                                                  # each (B,H,T,hs)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, T, _ = x.shape
        q, k, v = self.qkv(x)
        o = F.scaled_dot_product_attention(q, k, v, is_causal=True)     # scale 1/sqrt(hs)
        return o.transpose(1, 2).reshape(B, T, self.n_embed)            # head-major columns


class Block(nn.Module):
    """Pre-LN block: x + attn(ln1(x)); x + mlp(ln2(x)) (reference transformer_block.py:44-46)."""

    def __init__(self, n_head: int, n_embed: int, context_length: int) -> None:
        super().__init__()
        self.ln1 = nn.LayerNorm(n_embed)
        self.attn = MultiHeadAttention(n_head, n_embed, context_length)
        self.ln2 = nn.LayerNorm(n_embed)
        self.mlp = MLP(n_embed)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x + self.attn(self.ln1(x))
        x = x + self.mlp(self.ln2(x))
        return x

    def forward_embedding(self, x: torch.Tensor):
        """(post-ReLU MLP hidden (B,T,4d), residual (B,T,d)) - reference :48-61."""
        res = x + self.attn(self.ln1(x))
        return self.mlp.forward_embedding(self.ln2(res)), res
