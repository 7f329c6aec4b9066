"""Flat packed parameter layout and the translation to/from the reference checkpoint layout.

The reference stores one ``(hs, d)`` Linear per head for each of key/query/value plus a
``(ctx, ctx)`` ``tril`` buffer per head (reference ``src/models/attention.py:29-33``); this
framework stores ONE flat fp32 buffer whose slices are the packed tensors the kernels want:

    token_embed.weight (V,d) | position_embed.weight (ctx,d) |
    per block i:  ln1.weight ln1.bias | attn.wqkv (3d,d) = [Wq; Wk; Wv], head h = rows
                  [h*hs,(h+1)*hs) of each third | ln2.weight ln2.bias |
                  mlp.hidden.weight (4d,d) mlp.hidden.bias | mlp.proj.weight (d,4d) mlp.proj.bias |
    layer_norm.weight layer_norm.bias | lm_head.weight (V,d) lm_head.bias (V)

Every slice starts on a multiple of ``ALIGN`` elements so that the bf16 shadow copy of a
slice is 256-byte aligned (TMA descriptors need 16 B; vector loads and the per-rank shard
boundaries of the fused reduce+AdamW kernel like more).

``to_reference_state_dict`` / ``from_reference_state_dict`` implement the contract of
SURVEY 2.6 (key names, order, shapes, dtypes, the ``pos_idxs`` and per-head ``tril``
buffers, ``_orig_mod.`` / ``module.`` prefixes - defect D11), and the same for the torch
``AdamW.state_dict()`` layout of reference ``scripts/train_transformer.py:106-109,126``.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, List, Tuple

import torch

ALIGN = 128  # elements


@dataclass(frozen=True)
class Slot:
    name: str
    shape: Tuple[int, ...]
    offset: int  # element offset in the flat buffer
    layer: int   # -1 embeddings, 0..L-1 blocks, L = final LN + LM head

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


class ParamLayout:
    """Name -> slice map of the flat parameter buffer for one model shape."""

    def __init__(self, n_head: int, n_embed: int, context_length: int, vocab_size: int,
                 n_blocks: int):
        assert n_embed % n_head == 0
        self.H, self.d, self.ctx, self.V, self.L = n_head, n_embed, context_length, vocab_size, n_blocks
        self.hs = n_embed // n_head
        d, V, ctx = self.d, self.V, self.ctx
        self.slots: "OrderedDict[str, Slot]" = OrderedDict()
        self._cursor = 0
        self.layer_ranges: List[Tuple[int, int]] = []  # [start, end) element range per bucket

        def add(name, shape, layer):
            off = self._cursor
            s = Slot(name, tuple(shape), off, layer)
            self.slots[name] = s
            self._cursor = (off + s.numel + ALIGN - 1) // ALIGN * ALIGN

        start = self._cursor
        add("token_embed.weight", (V, d), -1)
        add("position_embed.weight", (ctx, d), -1)
        self.embed_range = (start, self._cursor)
        for i in range(n_blocks):
            start = self._cursor
            p = f"attn_blocks.{i}."
            add(p + "ln1.weight", (d,), i)
            add(p + "ln1.bias", (d,), i)
            add(p + "attn.wqkv", (3 * d, d), i)
            add(p + "ln2.weight", (d,), i)
            add(p + "ln2.bias", (d,), i)
            add(p + "mlp.hidden.weight", (4 * d, d), i)
            add(p + "mlp.hidden.bias", (4 * d,), i)
            add(p + "mlp.proj.weight", (d, 4 * d), i)
            add(p + "mlp.proj.bias", (d,), i)
            self.layer_ranges.append((start, self._cursor))
        start = self._cursor
        add("layer_norm.weight", (d,), n_blocks)
        add("layer_norm.bias", (d,), n_blocks)
        add("lm_head.weight", (V, d), n_blocks)
        add("lm_head.bias", (V,), n_blocks)
        self.head_range = (start, self._cursor)
        self.total = self._cursor

    # ------------------------------------------------------------------ basic helpers
    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        s = self.slots[name]
        return flat[s.offset:s.offset + s.numel].view(s.shape)

    def num_params(self) -> int:
        return sum(s.numel for s in self.slots.values())

    def buckets(self) -> List[Tuple[str, int, int]]:
        """Gradient buckets in the order backward finishes them: LM head, blocks L-1..0,
        embeddings.  Each is a contiguous [start, end) element range of the flat buffer."""
        out = [("head", *self.head_range)]
        for i in reversed(range(self.L)):
            out.append((f"block{i}", *self.layer_ranges[i]))
        out.append(("embed", *self.embed_range))
        return out

    # ------------------------------------------------------------------ reference layout
    def reference_param_names(self) -> List[str]:
        """Parameter names in the reference's ``model.parameters()`` order."""
        names = ["token_embed.weight", "position_embed.weight"]
        for i in range(self.L):
            p = f"attn_blocks.{i}."
            names += [p + "ln1.weight", p + "ln1.bias"]
            for h in range(self.H):
                q = p + f"attn.heads.{h}."
                names += [q + "key.weight", q + "query.weight", q + "value.weight"]
            names += [p + "ln2.weight", p + "ln2.bias", p + "mlp.hidden.weight",
                      p + "mlp.hidden.bias", p + "mlp.proj.weight", p + "mlp.proj.bias"]
        names += ["layer_norm.weight", "layer_norm.bias", "lm_head.weight", "lm_head.bias"]
        return names

    def reference_keys(self) -> List[str]:
        """All state_dict keys (parameters and buffers) in the reference's registration order."""
        keys = ["pos_idxs", "token_embed.weight", "position_embed.weight"]
        for i in range(self.L):
            p = f"attn_blocks.{i}."
            keys += [p + "ln1.weight", p + "ln1.bias"]
            for h in range(self.H):
                q = p + f"attn.heads.{h}."
                keys += [q + "tril", q + "key.weight", q + "query.weight", q + "value.weight"]  # buffers first
            keys += [p + "ln2.weight", p + "ln2.bias", p + "mlp.hidden.weight",
                     p + "mlp.hidden.bias", p + "mlp.proj.weight", p + "mlp.proj.bias"]
        keys += ["layer_norm.weight", "layer_norm.bias", "lm_head.weight", "lm_head.bias"]
        return keys

    def _ref_source(self, ref_name: str) -> Tuple[str, int, int]:
        """reference param name -> (packed slot name, first row, n rows); rows=-1 = whole."""
        if ".attn.heads." in ref_name:
            pre, rest = ref_name.split(".attn.heads.")
            h, which, _ = rest.split(".")
            third = {"query": 0, "key": 1, "value": 2}[which]
            return pre + ".attn.wqkv", third * self.d + int(h) * self.hs, self.hs
        return ref_name, 0, -1

    def ref_view(self, flat: torch.Tensor, ref_name: str) -> torch.Tensor:
        """The slice of ``flat`` holding the reference parameter ``ref_name`` (a view)."""
        slot, r0, n = self._ref_source(ref_name)
        t = self.view(flat, slot)
        return t if n < 0 else t[r0:r0 + n]

    def to_reference_state_dict(self, flat: torch.Tensor, *, include_buffers: bool = True,
                                clone: bool = True) -> "OrderedDict[str, torch.Tensor]":
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        dev = flat.device
        tril = None
        if include_buffers:
            # one shared tensor for every head's key: torch.save stores shared storage once
            # (never H*L copies of a ctx*ctx fp32 mask)
            tril = torch.tril(torch.ones(self.ctx, self.ctx, device=dev))
        for k in self.reference_keys():
            if k == "pos_idxs":
                if include_buffers:
                    out[k] = torch.arange(self.ctx, device=dev)
            elif k.endswith(".tril"):
                if include_buffers:
                    out[k] = tril
            else:
                v = self.ref_view(flat, k).detach()
                out[k] = v.clone() if clone else v
        return out

    @staticmethod
    def strip_prefixes(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Drop ``module.`` (DDP) and ``_orig_mod.`` (torch.compile) prefixes (defect D11)."""
        out = OrderedDict()
        for k, v in sd.items():
            changed = True
            while changed:
                changed = False
                for pre in ("module.", "_orig_mod."):
                    if k.startswith(pre):
                        k = k[len(pre):]
                        changed = True
            out[k] = v
        return out

    def from_reference_state_dict(self, flat: torch.Tensor, sd: Dict[str, torch.Tensor],
                                  strict: bool = True) -> Tuple[List[str], List[str]]:
        """Copy a reference-layout (or packed-layout) state_dict into ``flat``.
        Returns (missing_keys, unexpected_keys) like ``nn.Module.load_state_dict``."""
        sd = self.strip_prefixes(sd)
        packed = any(k.endswith("attn.wqkv") for k in sd)
        expected = list(self.slots) if packed else self.reference_param_names()
        buffers = set() if packed else {k for k in self.reference_keys() if k not in set(expected)}
        missing, seen = [], set()
        with torch.no_grad():
            for name in expected:
                if name not in sd:
                    missing.append(name)
                    continue
                seen.add(name)
                dst = self.view(flat, name) if packed else self.ref_view(flat, name)
                src = sd[name]
                if tuple(src.shape) != tuple(dst.shape):
                    raise RuntimeError(f"size mismatch for {name}: checkpoint {tuple(src.shape)} "
                                       f"vs model {tuple(dst.shape)}")
                dst.copy_(src)
        unexpected = [k for k in sd if k not in seen and k not in buffers]
        if strict:
            # buffers are derivable (arange / tril): the reference's strict load wants them,
            # ours tolerates their absence but checks shapes when present
            if missing or unexpected:
                raise RuntimeError(f"Error(s) in loading state_dict: missing keys {missing[:8]}"
                                   f"{'...' if len(missing) > 8 else ''}, unexpected keys "
                                   f"{unexpected[:8]}{'...' if len(unexpected) > 8 else ''}")
        return missing, unexpected

    # ------------------------------------------------------------------ optimizer state
    def to_reference_optimizer_state(self, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
                                     step: int, lr: float, betas=(0.9, 0.999), eps=1e-8,
                                     weight_decay=0.01) -> dict:
        """torch ``AdamW.state_dict()`` layout indexed by the reference's parameter order."""
        names = self.reference_param_names()
        state = {}
        for i, n in enumerate(names):
            state[i] = {
                "step": torch.tensor(float(step)),
                "exp_avg": self.ref_view(exp_avg, n).detach().clone(),
                "exp_avg_sq": self.ref_view(exp_avg_sq, n).detach().clone(),
            }
        group = {"lr": lr, "betas": tuple(betas), "eps": eps, "weight_decay": weight_decay,
                 "amsgrad": False, "maximize": False, "foreach": None, "capturable": False,
                 "differentiable": False, "fused": None, "decoupled_weight_decay": True,
                 "params": list(range(len(names)))}
        return {"state": state if step > 0 else {}, "param_groups": [group]}

    def from_reference_optimizer_state(self, osd: dict, exp_avg: torch.Tensor,
                                       exp_avg_sq: torch.Tensor) -> Tuple[int, dict]:
        """Inverse of the above.  Returns (step, param_group_hyperparams)."""
        names = self.reference_param_names()
        step = 0
        with torch.no_grad():
            for i, n in enumerate(names):
                st = osd["state"].get(i)
                if st is None:
                    continue
                self.ref_view(exp_avg, n).copy_(st["exp_avg"])
                self.ref_view(exp_avg_sq, n).copy_(st["exp_avg_sq"])
                step = int(float(st["step"]))
        return step, dict(osd["param_groups"][0])


def iter_init_order(layout: ParamLayout) -> Iterable[str]:
    """Reference parameter names in the order its constructors draw random numbers
    (SURVEY 7.3 item 7): embeddings, then per block per head key/query/value, MLP hidden w/b,
    proj w/b, finally lm_head w/b.  LayerNorms draw nothing."""
    yield "token_embed.weight"
    yield "position_embed.weight"
    for i in range(layout.L):
        p = f"attn_blocks.{i}."
        for h in range(layout.H):
            for which in ("key", "query", "value"):
                yield p + f"attn.heads.{h}.{which}.weight"
        yield p + "mlp.hidden.weight"
        yield p + "mlp.hidden.bias"
        yield p + "mlp.proj.weight"
        yield p + "mlp.proj.bias"
    yield "lm_head.weight"
    yield "lm_head.bias"
