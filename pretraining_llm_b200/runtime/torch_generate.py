"""Plain-PyTorch generation with a KV cache (CPU path and GPU oracle).

Semantics of reference ``Transformer.generate`` (``src/models/transformer.py:96-114``); see
``Transformer.generate`` for the cache/window discussion.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


def sample_next(logits: torch.Tensor, temperature: float, top_k: Optional[int],
                generator: Optional[torch.Generator]) -> torch.Tensor:
    """softmax -> multinomial(1), the reference's sampler (transformer.py:110-112)."""
    logits = logits.float()
    if temperature != 1.0:
        logits = logits / temperature
    if top_k is not None:
        kth = torch.topk(logits, min(top_k, logits.size(-1)), dim=-1).values[..., -1:]
        logits = logits.masked_fill(logits < kth, float("-inf"))
    probs = F.softmax(logits, dim=-1)
    return torch.multinomial(probs, num_samples=1, generator=generator)


def _step_with_cache(model, tok: torch.Tensor, pos0: int, cache) -> torch.Tensor:
    """Run tokens ``tok`` (B,t) at positions pos0.. through the model, appending to ``cache``
    (list of [k, v] per layer, each (B,H,len,hs)); return last-position logits (B,V)."""
    B, t = tok.shape
    pos = torch.arange(pos0, pos0 + t, device=tok.device)
    x = model.token_embed(tok) + model.position_embed(pos)
    for li, block in enumerate(model.attn_blocks):
        h = block.ln1(x)
        q, k, v = block.attn.qkv(h)
        if cache[li] is None:
            cache[li] = [k, v]
        else:
            cache[li][0] = torch.cat([cache[li][0], k], dim=2)
            cache[li][1] = torch.cat([cache[li][1], v], dim=2)
        K, V = cache[li]
        if t == 1:
            o = F.scaled_dot_product_attention(q, K, V)
        else:  # prefill chunk against the whole cache: causal w.r.t. absolute positions
            L = K.shape[2]
            mask = (torch.arange(L, device=tok.device)[None, :]
                    <= (pos0 + torch.arange(t, device=tok.device))[:, None])
            o = F.scaled_dot_product_attention(q, K, V, attn_mask=mask)
        x = x + o.transpose(1, 2).reshape(B, t, -1)
        x = x + block.mlp(block.ln2(x))
    x = model.layer_norm(x[:, -1:, :])
    return model.lm_head(x)[:, 0, :]


@torch.no_grad()
def generate_torch(model, idx: torch.Tensor, max_new_tokens: int, *, temperature: float = 1.0,
                   top_k: Optional[int] = None, use_cache: bool = True,
                   generator: Optional[torch.Generator] = None,
                   return_logits: bool = False):
    ctx = model.context_length
    cache = [None] * len(model.attn_blocks)
    cached_len = 0
    trace = []
    for _ in range(max_new_tokens):
        T = idx.shape[1]
        if use_cache and T <= ctx:
            logits = _step_with_cache(model, idx[:, cached_len:], cached_len, cache)
            cached_len = T
        else:
            # window overflow (or cache disabled): what the reference does every step
            idx_cond = idx[:, -ctx:]
            x = model._pre_attn_pass(idx_cond)
            for block in model.attn_blocks:
                x = block(x)
            logits = model.lm_head(model.layer_norm(x[:, -1:, :]))[:, 0, :]
        if return_logits:
            trace.append(logits.float().clone())
        nxt = sample_next(logits, temperature, top_k, generator)
        idx = torch.cat((idx, nxt.to(idx.dtype)), dim=1)
    return (idx, trace) if return_logits else idx
