"""``Transformer`` - the reference's model class re-built over a flat packed parameter store.

Call surface of reference ``src/models/transformer.py:6-114``: constructor
``Transformer(n_head, n_embed, context_length, vocab_size, N_BLOCKS)`` (``n_blocks=`` is also
accepted - the reference's own training script passes that spelling and crashes),
``forward(idx, targets=None) -> (logits, loss)``, ``forward_embedding(idx)``,
``generate(idx, max_new_tokens)``, and a ``state_dict()`` / ``load_state_dict()`` that speak the
reference's checkpoint layout (per-head key/query/value, ``tril`` and ``pos_idxs`` buffers).

Storage: all parameters are views into ONE flat fp32 buffer (``models/layout.py``); on a
B200 the whole-model forward/backward, the optimizer and generation run in
``runtime.engine.Engine`` over that buffer with hand-written sm_100a kernels.  The module
tree (``attn_blocks[i].ln1/attn/ln2/mlp`` ...) stays usable piece by piece and is the CPU path.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .layout import ParamLayout
from .modules import Block


class Transformer(nn.Module):
    def __init__(self, n_head: int, n_embed: int, context_length: int, vocab_size: int,
                 N_BLOCKS: Optional[int] = None, *, n_blocks: Optional[int] = None) -> None:
        super().__init__()
        if N_BLOCKS is None:
            N_BLOCKS = n_blocks
        if N_BLOCKS is None:
            raise TypeError("Transformer() needs N_BLOCKS (or n_blocks)")
        self.n_head, self.n_embed, self.vocab_size = n_head, n_embed, vocab_size
        self.context_length = context_length
        self.N_BLOCKS = N_BLOCKS
        self.layout = ParamLayout(n_head, n_embed, context_length, vocab_size, N_BLOCKS)
        # construction order == the reference's, so the same torch seed gives the same weights
        self.token_embed = nn.Embedding(vocab_size, n_embed)
        self.position_embed = nn.Embedding(context_length, n_embed)
        self.attn_blocks = nn.ModuleList([Block(n_head, n_embed, context_length)
                                          for _ in range(N_BLOCKS)])
        self.layer_norm = nn.LayerNorm(n_embed)
        self.lm_head = nn.Linear(n_embed, vocab_size)
        self._flat: Optional[torch.Tensor] = None
        self._engine = None
        self.use_native = True      # set False to force the plain-PyTorch path on CUDA
        self._flatten()

    # ------------------------------------------------------------------ flat storage
    def _flatten(self) -> None:
        """(Re)build the flat buffer and re-point every Parameter at its slice."""
        params = dict(super().named_parameters())
        any_p = next(iter(params.values()))
        flat = torch.zeros(self.layout.total, dtype=any_p.dtype, device=any_p.device)
        with torch.no_grad():
            for name, p in params.items():
                v = self.layout.view(flat, name)
                v.copy_(p.data)
                p.data = v
        self._flat = flat
        self._engine = None

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        self._flatten()
        return out

    @property
    def flat_params(self) -> torch.Tensor:
        return self._flat

    def num_parameters(self) -> int:
        return self.layout.num_params()

    # ------------------------------------------------------------------ engine dispatch
    def _native_ok(self) -> bool:
        if self.use_native and self._engine is not None and getattr(self._engine, "store", None) is not None:
            return True          # sharded data parallelism: the parameters live in the engine's shards, not in _flat
        if not self.use_native or self._flat is None or self._flat.device.type != "cuda":
            return False
        if self._flat.dtype != torch.float32:
            return False
        from ..ops import lib
        return lib.available()

    def _engine_has_full_params(self) -> bool:
        return self._flat is not None and (self._engine is None or getattr(self._engine, "store", None) is None)

    @property
    def engine(self):
        if self._engine is None:
            from ..runtime.engine import Engine
            self._engine = Engine(self)
        return self._engine

    # ------------------------------------------------------------------ forward
    def _pre_attn_pass(self, idx: torch.Tensor) -> torch.Tensor:
        B, T = idx.shape
        if T > self.context_length:
            raise IndexError(f"sequence length {T} exceeds context_length {self.context_length}")
        pos = torch.arange(T, device=idx.device)
        return self.token_embed(idx) + self.position_embed(pos)

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None
                ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """(logits (B,T,V), loss or None) - reference transformer.py:56-78.

        On the native path under bf16 training the logits are the bf16 tensor the fused LM-head
        kernel wrote (the reference's autocast logits are bf16 too) and ``loss`` carries a single
        autograd node whose backward runs the engine's hand-written backward schedule."""
        if self._native_ok():
            return self.engine.module_forward(idx, targets)
        x = self._pre_attn_pass(idx)
        for block in self.attn_blocks:
            x = block(x)
        x = self.layer_norm(x)
        logits = self.lm_head(x)
        loss = None
        if targets is not None:
            B, T, C = logits.shape
            loss = F.cross_entropy(logits.view(B * T, C), targets.reshape(B * T).long())
        return logits, loss

    def forward_embedding(self, idx: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """(post-ReLU hidden of the LAST block (B,T,4d), its residual input (B,T,d)).

        Equals the reference (transformer.py:80-94) for N_BLOCKS == 1.  For deeper models the
        reference feeds a (B,T,4d) tensor into the next block and raises; here
        blocks 0..L-2 run normally and the last block reports its embedding.  On a CUDA model the sweep runs on the
        native kernels (bf16 compute, fp32 copies returned) unless gradients are wanted."""
        if self._native_ok() and not torch.is_grad_enabled() and self._engine_has_full_params():
            return self.engine.forward_embedding(idx)
        x = self._pre_attn_pass(idx)
        residual = x
        for i, block in enumerate(self.attn_blocks):
            if i + 1 < len(self.attn_blocks):
                x = block(x)
            else:
                x, residual = block.forward_embedding(x)
        return x, residual

    # ------------------------------------------------------------------ generation
    @torch.no_grad()
    def generate(self, idx: torch.Tensor, max_new_tokens: int, *, temperature: float = 1.0,
                 top_k: Optional[int] = None, use_cache: bool = True,
                 generator: Optional[torch.Generator] = None) -> torch.Tensor:
        """Autoregressive sampling, reference semantics (transformer.py:96-114): softmax of the
        last position's logits, one multinomial draw per row, append.  Defaults reproduce the
        reference (temperature 1, no top-k).

        While ``len <= context_length`` a KV cache makes each step O(len) instead of a full
        re-forward; once the window overflows, the reference re-bases learned absolute positions
        every step (``idx[:, -context_length:]``) which invalidates any cache, so the method
        falls back to a windowed full forward per token exactly like the reference."""
        if self._native_ok():
            return self.engine.generate(idx, max_new_tokens, temperature=temperature,
                                        top_k=top_k, use_cache=use_cache, generator=generator)
        from ..runtime.torch_generate import generate_torch
        return generate_torch(self, idx, max_new_tokens, temperature=temperature, top_k=top_k,
                              use_cache=use_cache, generator=generator)

    # ------------------------------------------------------------------ checkpoint layout
    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        """Reference-layout state_dict - loads with the reference's strict
        ``load_state_dict``.  ``packed_state_dict()`` gives the native layout."""
        if args:  # legacy positional (destination, prefix, keep_vars)
            destination = args[0]
            prefix = args[1] if len(args) > 1 else prefix
        if self._flat is None:
            raise RuntimeError("the parameters are sharded across ranks (dp_mode='zero'): call "
                               "engine.dp.gather_full_state() on every rank first (Trainer.save_checkpoint does)")
        sd = self.layout.to_reference_state_dict(self._flat, clone=not keep_vars)
        out = destination if destination is not None else OrderedDict()
        for k, v in sd.items():
            out[prefix + k] = v
        return out

    def packed_state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((n, self.layout.view(self._flat, n).detach().clone())
                           for n in self.layout.slots)

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        if self._flat is None:
            raise RuntimeError("the parameters are sharded across ranks (dp_mode='zero'): call "
                               "engine.dp.materialize() on every rank first (Trainer.load_checkpoint does)")
        missing, unexpected = self.layout.from_reference_state_dict(self._flat, state_dict, strict)
        if self._engine is not None:
            self._engine.params_changed()
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)


def strip_checkpoint_prefixes(sd):
    return ParamLayout.strip_prefixes(sd)
