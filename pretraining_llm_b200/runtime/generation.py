"""KV-cache generation on the native kernels.

Semantics of the reference's ``Transformer.generate`` (reference
``src/models/transformer.py:96-114``): at every step sample one token per row from
``softmax(logits[:, -1])`` and append it.  The reference recomputes the whole window for every
token in fp32; here

  * prefill runs the training forward kernels once over the prompt (flash attention) while the
    K/V rows of every layer are appended to a ``[L, B, H, ctx, hs]`` bf16 cache, and only the LAST
    position goes through the LM head;
  * each further token is ONE CUDA-graph replay of the decode step (embed -> L x [LN, QKV GEMM,
    cache append, decode attention, LN, MLP GEMMs] -> LN -> LM head -> sampler -> scatter), with
    the position read from device memory so the same graph serves every step and nothing
    synchronises with the host;
  * once ``len > context_length`` the reference re-bases the learned absolute positions every
    step (``idx[:, -context_length:]``), which invalidates any cache: from there on each token is
    a windowed full forward, exactly like the reference.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from .. import ops

BF16, F32 = torch.bfloat16, torch.float32


class _DecodeBuffers:
    def __init__(self, eng, B: int):
        lay = eng.layout
        d, V = lay.d, lay.V
        dev = eng.device
        e = lambda *s, dt=BF16: torch.empty(*s, device=dev, dtype=dt)
        self.B = B
        self.ra, self.rb = e(B, d, dt=F32), e(B, d, dt=F32)
        self.h = e(B, d)
        self.qkv = e(B, 3 * d)
        self.att = e(B, d)
        self.act = e(B, 4 * d)
        self.mlp_out = e(B, d)
        self.hf = e(B, d)
        self.Vp = (V + 7) // 8 * 8
        self.logits = e(B, self.Vp)
        self.next_tok = torch.zeros(B, 1, device=dev, dtype=torch.int64)
        self.pos = torch.zeros(1, device=dev, dtype=torch.int32)
        self.kc = torch.zeros(lay.L, B, lay.H, lay.ctx, lay.hs, device=dev, dtype=BF16)
        self.vc = torch.zeros_like(self.kc)
        self.seq = torch.zeros(B, lay.ctx + 1, device=dev, dtype=torch.int64)
        # Decode attention: one CTA per (row, head) when that alone fills the GPU (B * H >= 2 x 148: measured 1.77 ms per
        # token step at batch 32 on the 1.3B shape against 2.14 - 2.40 ms with the split kernel, whose 8-key blocks keep
        # fewer bytes in flight per round trip); the split-KV kernel when the batch is small and B * H CTAs would leave
        # most SMs idle however long the context is.  B200_DECODE_SPLITS=n forces n splits (0 = the one-CTA kernel).
        want = os.environ.get("B200_DECODE_SPLITS")
        if want is not None:
            self.splits = int(want)
        else:
            self.splits = 0 if B * lay.H >= 2 * 148 else max(1, min(8, (4 * 148 + B * lay.H - 1) // (B * lay.H)))
        self.att_scratch = torch.empty(B * lay.H * max(self.splits, 1) * (lay.hs + 2), device=dev, dtype=F32)
        self.att_tickets = torch.zeros(B * lay.H, device=dev, dtype=torch.int32)


class Generator:
    def __init__(self, eng):
        self.eng = eng
        self.bufs = {}

    def _buffers(self, B: int) -> _DecodeBuffers:
        b = self.bufs.get(B)
        if b is None:
            self.bufs.clear()
            b = self.bufs[B] = _DecodeBuffers(self.eng, B)
        return b

    # ------------------------------------------------------------------ pieces
    def _gemm(self, db, x, w, out, bias=None, relu=False) -> None:
        """Decode-sized GEMM: weight-streaming kernel when the batch fits one MMA operand, tile kernel otherwise."""
        if db is not None and x.shape[0] <= 128:
            ops.gemm_skinny(x, w, out, bias=bias, relu=relu)
        else:
            ops.gemm(x, w, out=out, bias=bias, relu=relu)

    def _head(self, hf: torch.Tensor, logits: torch.Tensor, db=None) -> None:
        eng = self.eng
        V = eng.layout.V
        self._gemm(db, hf, eng.W("lm_head.weight"), logits[:, :V], bias=eng.P("lm_head.bias"))

    def _prefill(self, idx: torch.Tensor, db: Optional[_DecodeBuffers], logits: torch.Tensor) -> None:
        """Forward over the prompt; fills the KV cache (if db) and the last-position logits."""
        eng = self.eng
        B, T = idx.shape
        w = eng.workspace(B, T, False)
        eng.sync_weights()
        w.idx.copy_(idx)
        ops.embed_fwd(w.idx, eng.P("token_embed.weight"), eng.P("position_embed.weight"), out=w.resid_a[0])
        eng._blocks_forward(w, kv=(db.kc, db.vc) if db is not None else None, pos0=0)
        hf_last = db.hf if db is not None else torch.empty(B, eng.layout.d, device=eng.device, dtype=BF16)
        ops.gather_last_rows(w.hf, hf_last, B, T)
        self._head(hf_last, logits, db)

    def _decode_step(self, db: _DecodeBuffers, temperature: float, seed: int, sample: bool = True) -> None:
        """One token for every row: reads db.next_tok / db.pos, writes db.logits, db.next_tok, db.seq."""
        eng = self.eng
        lay = eng.layout
        B, L, H, hs, ctx = db.B, lay.L, lay.H, lay.hs, lay.ctx
        ops.embed_fwd(db.next_tok, eng.P("token_embed.weight"), eng.P("position_embed.weight"), out=db.ra,
                      pos_dev=db.pos)
        for l in range(L):
            pre = f"attn_blocks.{l}."
            if l == 0:
                ops.ln_fwd(db.ra, eng.P(pre + "ln1.weight"), eng.P(pre + "ln1.bias"), out=db.h, want_stats=False)
            else:
                ops.ln_fwd(db.rb, eng.P(pre + "ln1.weight"), eng.P(pre + "ln1.bias"), branch=db.mlp_out,
                           x_out=db.ra, out=db.h, want_stats=False)
            self._gemm(db, db.h, eng.W(pre + "attn.wqkv"), db.qkv)
            if db.splits > 0:
                ops.attn_decode_split(db.qkv, db.kc[l], db.vc[l], db.att, B, H, hs, ctx, db.att_scratch, db.att_tickets,
                                      db.splits, pos_dev=db.pos, append=True)
            else:
                ops.attn_decode(db.qkv, db.kc[l], db.vc[l], db.att, B, H, hs, ctx, pos_dev=db.pos, append=True)
            ops.ln_fwd(db.ra, eng.P(pre + "ln2.weight"), eng.P(pre + "ln2.bias"), branch=db.att, x_out=db.rb,
                       out=db.h, want_stats=False)
            self._gemm(db, db.h, eng.W(pre + "mlp.hidden.weight"), db.act, bias=eng.P(pre + "mlp.hidden.bias"), relu=True)
            self._gemm(db, db.act, eng.W(pre + "mlp.proj.weight"), db.mlp_out, bias=eng.P(pre + "mlp.proj.bias"))
        ops.ln_fwd(db.rb, eng.P("layer_norm.weight"), eng.P("layer_norm.bias"), branch=db.mlp_out, x_out=db.ra,
                   out=db.hf, want_stats=False)
        self._head(db.hf, db.logits, db)
        if sample:
            ops.sample(db.logits, db.next_tok, lay.V, temperature=temperature, seed=seed, step=1, pos_dev=db.pos)
            ops.scatter_tokens(db.next_tok, db.seq, pos_dev=db.pos)
            ops.incr_int(db.pos, 1)

    def _sample_host_topk(self, logits: torch.Tensor, out_tok: torch.Tensor, V: int, temperature: float,
                          top_k: Optional[int], seed: int, step: int) -> None:
        if top_k is not None:
            lv = logits[:, :V]
            kth = torch.topk(lv.float(), min(top_k, V), dim=-1).values[:, -1:]
            lv.masked_fill_(lv.float() < kth, float("-inf"))
        ops.sample(logits, out_tok, V, temperature=temperature, seed=seed, step=step)

    # ------------------------------------------------------------------ public
    @torch.no_grad()
    def generate(self, idx: torch.Tensor, max_new_tokens: int, *, temperature: float = 1.0,
                 top_k: Optional[int] = None, use_cache: bool = True,
                 generator: Optional[torch.Generator] = None, return_logits: bool = False,
                 use_graph: bool = True):
        eng = self.eng
        lay = eng.layout
        ctx, V = lay.ctx, lay.V
        idx = idx.to(eng.device).long().contiguous()
        B, T0 = idx.shape
        seed = int(torch.randint(0, 2 ** 62, (1,), generator=generator).item())
        out = torch.empty(B, T0 + max_new_tokens, device=eng.device, dtype=torch.int64)
        out[:, :T0] = idx
        trace = []
        produced = 0
        if max_new_tokens <= 0:
            return (out, trace) if return_logits else out
        n_cached = min(max_new_tokens, ctx - T0 + 1) if (use_cache and T0 <= ctx) else 0
        if n_cached > 0:
            db = self._buffers(B)
            db.seq[:, :T0] = idx
            self._prefill(idx, db, db.logits)
            if return_logits:
                trace.append(db.logits[:, :V].float().clone())
            self._sample_host_topk(db.logits, db.next_tok, V, temperature, top_k, seed, step=T0)
            ops.scatter_tokens(db.next_tok, db.seq, col=T0)
            db.pos.fill_(T0)
            produced = 1
            graph = None
            eager = top_k is not None or return_logits or not use_graph
            while produced < n_cached:
                if eager:
                    self._decode_step(db, temperature, seed, sample=False)
                    if return_logits:
                        trace.append(db.logits[:, :V].float().clone())
                    pos = T0 + produced - 1          # position just processed
                    self._sample_host_topk(db.logits, db.next_tok, V, temperature, top_k, seed, step=pos + 1)
                    ops.scatter_tokens(db.next_tok, db.seq, col=pos + 1)
                    ops.incr_int(db.pos, 1)
                elif graph is None and produced >= 2:
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        self._decode_step(db, temperature, seed)
                    graph.replay()
                elif graph is not None:
                    graph.replay()
                else:
                    self._decode_step(db, temperature, seed)   # warm-up eager step before capture
                produced += 1
            out[:, T0:T0 + produced] = db.seq[:, T0:T0 + produced]
        # window overflow (or cache disabled): the reference's per-token windowed full forward
        logits = torch.empty(B, (V + 7) // 8 * 8, device=eng.device, dtype=BF16)
        nxt = torch.empty(B, 1, device=eng.device, dtype=torch.int64)
        while produced < max_new_tokens:
            cur = T0 + produced
            cond = out[:, max(0, cur - ctx):cur].contiguous()
            self._prefill(cond, None, logits)
            if return_logits:
                trace.append(logits[:, :V].float().clone())
            self._sample_host_topk(logits, nxt, V, temperature, top_k, seed, step=cur)
            out[:, cur] = nxt[:, 0]
            produced += 1
        return (out, trace) if return_logits else out
