"""Engine: the static kernel schedule that runs a ``Transformer`` on a B200.

One ``Engine`` owns, for one model:
  * the flat fp32 master parameters (shared with the nn.Module's Parameters), a flat bf16 shadow
    copy the GEMMs read, a flat fp32 gradient buffer and the flat AdamW moments;
  * per-(B, T) workspaces with every activation the backward pass needs, allocated once;
  * the forward / backward / optimizer launch sequences (hand-written sm_100a kernels only - no
    autograd, no cuBLAS, no Triton on this path) and CUDA graphs of whole training steps.

Forward per block (reference src/models/transformer_block.py:44-46 + attention.py + mlp.py):
    h1 = LN1(r_a)                       ln_fwd            (r_a = residual stream, fp32)
    qkv = h1 @ Wqkv^T                   gemm              (packed [Wq; Wk; Wv])
    att = causal_attention(qkv)         attn_fwd          (flash, no T x T tensor)
    r_b = r_a + att ; h2 = LN2(r_b)     ln_fwd (fused residual add)
    act = relu(h2 @ W1^T + b1)          gemm (+bias +ReLU epilogue)
    mo  = act @ W2^T + b2               gemm (+bias epilogue)
    r_a' = r_b + mo ; h1' = LN1'(r_a')  ln_fwd of the NEXT block (or the final LayerNorm)
LM head (reference transformer.py:70-77): logits = hf @ Wlm^T + b with softmax statistics in the
GEMM epilogue, ce_finalize -> loss; backward turns the bf16 logits into dlogits in place.

Backward mirrors it with dgrad GEMMs (B operand MN-major, ReLU mask fused in the epilogue),
wgrad GEMMs (both operands MN-major, fp32 straight into the flat gradient buffer), flash
attention backward and the fused LN-backward (+ residual-gradient add + bias-gradient column
sums).  See DESIGN.md for the buffer plan.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

from .. import ops
from ..utils import nvtx

BF16, F32 = torch.bfloat16, torch.float32
# B200_DETERMINISTIC=1: every gradient is summed in a fixed order - the token-embedding gradient by sorting the tokens,
# the bias gradients by one CTA per column strip - instead of with fp32 atomics whose arrival order varies from run to
# run (the other kernels - GEMMs without split-K, attention, LayerNorm - are order-deterministic already)
DETERMINISTIC = os.environ.get("B200_DETERMINISTIC", "0") == "1"
WGRAD_STREAM = os.environ.get("B200_WGRAD_STREAM", "1") == "1"


class Workspace:
    """All activations and scratch for one (B, T) shape."""

    def __init__(self, eng: "Engine", B: int, T: int, train: bool):
        lay = eng.layout
        d, H, hs, L, V = lay.d, lay.H, lay.hs, lay.L, lay.V
        M = B * T
        dev = eng.device
        self.B, self.T, self.M, self.train = B, T, M, train
        Ls = L if train else 1                      # inference reuses slot 0 for every layer
        e = lambda *s, dt=BF16: torch.empty(*s, device=dev, dtype=dt)
        self.resid_a = e(Ls + 1, M, d, dt=F32)      # [l] = input of block l; [Ls] = input of final LN
        self.resid_b = e(Ls, M, d, dt=F32)
        self.h1, self.h2 = e(Ls, M, d), e(Ls, M, d)
        self.qkv = e(Ls, M, 3 * d)
        self.att = e(Ls, M, d)
        self.act = e(Ls, M, 4 * d)
        # sign bits of the ReLU outputs (one int64 per row and 64 columns): what the MLP dgrad epilogue masks with
        self.relu_bits = torch.empty(Ls, M, (4 * d + 63) // 64, device=dev, dtype=torch.int64)
        self.mlp_out = e(M, d)
        self.lse = e(Ls, B, H, T, dt=F32)
        self.ln_stats = e(2 * Ls + 1, 2, M, dt=F32)  # (mean, rstd) for ln1/ln2 of each layer + final
        self.hf = e(M, d)
        self.Vp = (V + 7) // 8 * 8
        self.logits = e(M, self.Vp)
        self.nt = 2 * ((V + 255) // 256)            # CE statistics entries per row: one per 256-column tile and epilogue group
        self.ce_stats = e(self.nt, M, 2, dt=F32)
        self.ce_lse = e(M, dt=F32)
        self.loss_rows = e(M, dt=F32)
        self.loss = torch.zeros(1, device=dev, dtype=F32)
        self.idx = torch.zeros(B, T, device=dev, dtype=torch.int64)
        self.tgt = torch.zeros(B, T, device=dev, dtype=torch.int64)
        if train:
            self.dx = e(M, d, dt=F32)
            # gradients read by BOTH a dgrad GEMM (main stream) and a wgrad GEMM (weight-gradient stream) are double
            # buffered by layer parity, so the next layer can overwrite one copy while the other is still being read
            self.dxa = [e(M, d), e(M, d), e(M, d)]      # bf16 gradient at a block's output (input of its MLP backward);
                                                        # three: layer l writes layer l-1's copy while layer l+1's wgrad may still read its own
            self.dxb = e(M, d)                          # bf16 gradient after the MLP branch (attention's dout)
            self.dh = [e(M, 4 * d), e(M, 4 * d)]
            self.dqkv = [e(M, 3 * d), e(M, 3 * d)]
            self.dtmp = e(M, d)
            self.delta = e(B, H, T, dt=F32)
            self.dq_acc = torch.zeros(M, d, device=dev, dtype=F32) if ops.ATTN_BWD_IMPL != "tc" else None
            self.ln_partial = e(max(ops.ln_bwd_nparts(M, d), 1) * 3 * d, dt=F32)
            self.ln_partial2 = torch.empty_like(self.ln_partial)     # second LayerNorm of a block (one finalize for both)
        self.graph: Optional[torch.cuda.CUDAGraph] = None


class Engine:
    def __init__(self, model):
        self.model = model
        self.layout = model.layout
        self.flat: torch.Tensor = model._flat
        assert self.flat.is_cuda and self.flat.dtype == F32
        self.device = self.flat.device
        if not ops.lib.load().b200_ln_supported(self.layout.d):
            raise ValueError(f"n_embed={self.layout.d} is not supported by the LayerNorm kernels")
        if self.layout.hs not in (64, 128):
            raise ValueError(f"head size {self.layout.hs} is not supported by the attention kernels (64 or 128)")
        n = self.layout.total
        self.wb = torch.empty(n, device=self.device, dtype=BF16)     # bf16 shadow weights
        self._wb_version = -1
        self.grad: Optional[torch.Tensor] = None
        self.exp_avg: Optional[torch.Tensor] = None
        self.exp_avg_sq: Optional[torch.Tensor] = None
        self.step_count = 0
        self.hyper = torch.zeros(8, device=self.device, dtype=F32)
        self.betas, self.eps, self.weight_decay = (0.9, 0.999), 1e-8, 0.01   # torch.optim.AdamW defaults
        self.ws: Dict[Tuple[int, int, bool], Workspace] = {}
        # weight gradients run on their own stream: nothing in the backward chain waits for them, so their CTAs fill the
        # SMs the dgrad GEMMs' last partial wave leaves idle and overlap the memory-bound LayerNorm / attention kernels
        self.wstream = torch.cuda.Stream(device=self.device) if WGRAD_STREAM else None
        self.dp = None                     # parallel.dp.DataParallel / parallel.zero.ZeroDataParallel
        self.store = None                  # sharded mode: the object that serves P / W / G instead of the flat buffers
        self.launches_per_step = 0
        self._gen = None

    # ------------------------------------------------------------------ parameters
    def W(self, name: str) -> torch.Tensor:       # bf16 weights the GEMMs read (flat shadow, or a gathered window)
        return self.store.W(name) if self.store is not None else self.layout.view(self.wb, name)

    def P(self, name: str) -> torch.Tensor:       # fp32 parameters consumed in fp32 (LayerNorm, biases, embeddings)
        return self.store.P(name) if self.store is not None else self.layout.view(self.flat, name)

    def G(self, name: str) -> torch.Tensor:       # fp32 gradient view (flat buffer, or a rolling window)
        return self.store.G(name) if self.store is not None else self.layout.view(self.grad, name)

    def params_changed(self) -> None:
        self._wb_version = -1

    def sync_weights(self) -> None:
        """Refresh the bf16 shadow if the fp32 masters were modified by anyone but our AdamW."""
        if self.store is not None:
            return                         # sharded: the reduce kernel keeps the bf16 shards in step with the masters
        v = self.flat._version
        if v != self._wb_version:
            ops.cast_bf16(self.flat, self.wb)
            self._wb_version = v

    def ensure_grad(self) -> None:
        if self.grad is None and self.store is None:
            self.grad = torch.zeros_like(self.flat)

    def ensure_opt_state(self) -> None:
        if self.exp_avg is None:
            self.exp_avg = torch.zeros_like(self.flat)
            self.exp_avg_sq = torch.zeros_like(self.flat)

    def release_full_opt_state(self) -> None:
        """Drop the full-size moment buffers (8 B/parameter).  With data parallelism attached they are only
        materialised for a checkpoint or a resume; the live moments are the owner slices in dp.m / dp.v."""
        self.exp_avg = None
        self.exp_avg_sq = None

    def workspace(self, B: int, T: int, train: bool) -> Workspace:
        key = (B, T, train)
        w = self.ws.get(key)
        if w is None:
            if len(self.ws) >= 4:          # keep memory bounded when shapes vary (generation prefill)
                for k in [k for k in self.ws if not k[2]][:2]:
                    del self.ws[k]
            w = self.ws[key] = Workspace(self, B, T, train)
        return w

    # ------------------------------------------------------------------ forward
    def _blocks_forward(self, w: Workspace, kv=None, pos0: int = 0) -> None:
        lay = self.layout
        B, T, L, H, hs = w.B, w.T, lay.L, lay.H, lay.hs
        z = self.store
        if z is not None:
            z.begin_pass()                 # sharded weights: a new all-gather generation
        for l in range(L):
            s = l if w.train else 0
            pre = f"attn_blocks.{l}."
            ra_in = w.resid_a[s]
            st1, st2 = w.ln_stats[2 * s], w.ln_stats[2 * s + 1]
            if l == 0:
                ops.ln_fwd(ra_in, self.P(pre + "ln1.weight"), self.P(pre + "ln1.bias"), out=w.h1[s],
                           mean=st1[0], rstd=st1[1])
            else:
                sp = (l - 1) if w.train else 0
                ops.ln_fwd(w.resid_b[sp], self.P(pre + "ln1.weight"), self.P(pre + "ln1.bias"), branch=w.mlp_out,
                           x_out=ra_in, out=w.h1[s], mean=st1[0], rstd=st1[1])
            if z is not None:  # sharded weights: this GEMM also carries the all-gather of the next layer
                z.gemm_fwd(l, w.h1[s], w.qkv[s])
            else:
                ops.gemm(w.h1[s], self.W(pre + "attn.wqkv"), out=w.qkv[s])
            if kv is not None:
                ops.kv_append(w.qkv[s], kv[0][l], kv[1][l], B, T, H, hs, kv[0].shape[3], pos0=pos0)
            ops.attn_fwd(w.qkv[s], B, T, H, hs, out=w.att[s], lse=w.lse[s])
            ops.ln_fwd(ra_in, self.P(pre + "ln2.weight"), self.P(pre + "ln2.bias"), branch=w.att[s],
                       x_out=w.resid_b[s], out=w.h2[s], mean=st2[0], rstd=st2[1])
            ops.gemm(w.h2[s], self.W(pre + "mlp.hidden.weight"), out=w.act[s],
                     bias=self.P(pre + "mlp.hidden.bias"), relu=True, relu_bits_out=w.relu_bits[s] if w.train else None)
            ops.gemm(w.act[s], self.W(pre + "mlp.proj.weight"), out=w.mlp_out, bias=self.P(pre + "mlp.proj.bias"))
        sl = (L - 1) if w.train else 0
        sf = L if w.train else 1
        stf = w.ln_stats[2 * (L if w.train else 1)]
        ops.ln_fwd(w.resid_b[sl], self.P("layer_norm.weight"), self.P("layer_norm.bias"), branch=w.mlp_out,
                   x_out=w.resid_a[sf], out=w.hf, mean=stf[0], rstd=stf[1])

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None, *, train: bool = False,
                loss_scale: float = 1.0):
        """Run the model.  Returns (workspace, logits view (B,T,V) bf16, loss 0-dim fp32 or None).
        ``train=True`` keeps every activation for ``backward``."""
        B, T = idx.shape
        if T > self.layout.ctx:
            raise IndexError(f"sequence length {T} exceeds context_length {self.layout.ctx}")
        if train and self.dp is not None and getattr(self.dp, "full_flat", None) is not None:
            self.dp.release_full()
        w = self.workspace(B, T, train)
        self.sync_weights()
        if idx.data_ptr() != w.idx.data_ptr():
            w.idx.copy_(idx)
        if targets is not None and targets.data_ptr() != w.tgt.data_ptr():
            w.tgt.copy_(targets.reshape(B, T))
        self._forward_ws(w, targets is not None, loss_scale)
        V = self.layout.V
        logits = w.logits[:, :V].view(B, T, V)
        return w, logits, (w.loss[0] if targets is not None else None)

    def _forward_ws(self, w: Workspace, with_loss: bool, loss_scale: float = 1.0) -> None:
        lay = self.layout
        if self.store is not None:         # sharded fp32 tables: rows come from their owners over NVLink
            self.store.embed_fwd(w.idx, w.resid_a[0])
        elif self.dp is not None and getattr(self.dp, "shard_embed", False):
            self.dp.embed_fwd(w.idx, w.resid_a[0])     # replicated mode: the tables alone stay with their owners
        else:
            ops.embed_fwd(w.idx, self.P("token_embed.weight"), self.P("position_embed.weight"), out=w.resid_a[0])
        self._blocks_forward(w)
        V = lay.V
        if with_loss and w.train and self.dp is not None and self.dp.vp is not None:
            self.dp.vp.forward(w, loss_scale)      # vocab-parallel LM head + CE across the ranks
            return
        ops.gemm(w.hf, self.W("lm_head.weight"), out=w.logits[:, :V], bias=self.P("lm_head.bias"),
                 stats=w.ce_stats if with_loss else None, bn=256)
        if with_loss:
            ops.zero_(w.loss)
            ops.ce_finalize(w.ce_stats, w.nt, w.logits, w.tgt.view(-1), w.ce_lse, w.loss_rows, w.loss, V,
                            sum_scale=loss_scale / w.M)

    # ------------------------------------------------------------------ backward
    def backward(self, w: Workspace, *, grad_scale: float = 1.0, accumulate: bool = False,
                 dp_sync: bool = True) -> None:
        """Gradients of (grad_scale * mean CE) w.r.t. every parameter -> the flat fp32 grad buffer.
        With data parallelism attached, each finished bucket is handed to the fused
        reduce+AdamW kernel on the communication stream (``dp_sync=False`` on all but the last
        micro-step of a gradient accumulation)."""
        dp = self.dp if dp_sync else None
        assert w.train, "backward needs a training workspace"
        self.ensure_grad()
        lay = self.layout
        B, T, M, L, H, hs, V = w.B, w.T, w.M, lay.L, lay.H, lay.hs, lay.V
        fm = 1 if accumulate else 0
        z = self.store
        if z is not None:
            # sharded mode: the gradient windows hold ONE micro-step; each micro-step is reduce-scattered and the owners
            # accumulate their slices (ZeroDataParallel.begin_backward), so locally nothing is ever accumulated
            z.begin_backward(accumulate, dp_sync)
            dp = self.dp
            fm = 0
            accumulate = False
        if not accumulate:
            ops.zero_(self.G("token_embed.weight"))  # scatter-add target (memset node, not a fill kernel)
        main = torch.cuda.current_stream()
        ws = self.wstream
        pending = []                                 # one event per layer: its weight gradients are complete
        used_w = [False]                             # an event of a stream that never joined a capture must not be waited on

        def on_w(fn):
            """Run fn on the weight-gradient stream after everything enqueued on the main stream so far."""
            if ws is None:
                fn()
                return
            ev = torch.cuda.Event()
            ev.record(main)
            ws.wait_event(ev)
            used_w[0] = True
            with torch.cuda.stream(ws):
                fn()

        def w_done():
            if ws is None or not used_w[0]:
                return None
            ev = torch.cuda.Event()
            ev.record(ws)
            return ev

        # LM head + CE
        if self.dp is not None and self.dp.vp is not None:
            self.dp.vp.backward(w, grad_scale, accumulate)
        else:
            dlog = w.logits[:, :V]
            if DETERMINISTIC:      # bias gradient by the fixed-order column-sum kernel instead of per-CTA fp32 atomics
                ops.ce_grad(dlog, w.ce_lse, w.tgt.view(-1), None, V, grad_scale / M, accumulate=accumulate)
                ops.colsum(dlog, out=self.G("lm_head.bias"), accumulate=accumulate)
            else:
                ops.ce_grad(dlog, w.ce_lse, w.tgt.view(-1), self.G("lm_head.bias"), V, grad_scale / M,
                            accumulate=accumulate)
            on_w(lambda: ops.gemm(dlog, w.hf, out=self.G("lm_head.weight"), a_mn=True, b_mn=True, f32_mode=fm))
            ops.gemm(dlog, self.W("lm_head.weight"), out=w.dtmp, b_mn=True)
        stf = w.ln_stats[2 * L]
        ops.ln_bwd(w.dtmp, w.resid_a[L], stf[0], stf[1], self.P("layer_norm.weight"), self.G("layer_norm.weight"),
                   self.G("layer_norm.bias"), dx_out=w.dx, dx_bf16=w.dxa[(L - 1) % 3], partial=w.ln_partial,
                   accumulate=accumulate)
        head_done = w_done()
        if dp is not None:
            dp.bucket_ready("head", head_done)
        for l in reversed(range(L)):
            pre = f"attn_blocks.{l}."
            st1, st2 = w.ln_stats[2 * l], w.ln_stats[2 * l + 1]
            par = l & 1
            dxa, dh, dqkv = w.dxa[l % 3], w.dh[par], w.dqkv[par]
            if ws is not None and len(pending) >= 2:
                main.wait_event(pending[-2])   # layer l+2's weight gradients have finished reading this parity's buffers
            if z is not None:
                z.before_layer_bwd(l)          # the gradient window of layer l+2 must have been reduced by every rank
            # ---- MLP
            on_w(lambda: ops.gemm(dxa, w.act[l], out=self.G(pre + "mlp.proj.weight"), a_mn=True, b_mn=True, f32_mode=fm))
            fused_bias = not DETERMINISTIC        # bias gradient of the MLP-up Linear: column sums of dh, taken from the
            if fused_bias:                         # dgrad GEMM's staged tiles (fp32 atomics) instead of a second pass
                if not accumulate:
                    ops.zero_(self.G(pre + "mlp.hidden.bias"))
                ops.gemm(dxa, self.W(pre + "mlp.proj.weight"), out=dh, b_mn=True, relu_bits=w.relu_bits[l],
                         colsum_out=self.G(pre + "mlp.hidden.bias"))
            else:
                ops.gemm(dxa, self.W(pre + "mlp.proj.weight"), out=dh, b_mn=True, relu_bits=w.relu_bits[l])

            def mlp_up_wgrad():
                if not fused_bias:
                    ops.colsum(dh, out=self.G(pre + "mlp.hidden.bias"), accumulate=accumulate)
                ops.gemm(dh, w.h2[l], out=self.G(pre + "mlp.hidden.weight"), a_mn=True, b_mn=True, f32_mode=fm)
            on_w(mlp_up_wgrad)
            if z is not None:                  # carries the all-gather of layer l-1's weights for its backward
                z.gemm_bwd(l, dh, w.dtmp)
            else:
                ops.gemm(dh, self.W(pre + "mlp.hidden.weight"), out=w.dtmp, b_mn=True)
            # LayerNorm backward in two steps: the partial sums of ln2 and ln1 are reduced by ONE launch at the end of the block
            ops.ln_bwd_partial(w.dtmp, w.resid_b[l], st2[0], st2[1], self.P(pre + "ln2.weight"), w.ln_partial,
                               dx_in=w.dx, dx_out=w.dx, dx_bf16=w.dxb, with_colsum=True)
            # ---- attention
            ops.attn_bwd(w.qkv[l], w.att[l], w.dxb, w.lse[l], B, T, H, hs, dqkv=dqkv, delta=w.delta,
                         dq_acc=w.dq_acc)
            on_w(lambda: ops.gemm(dqkv, w.h1[l], out=self.G(pre + "attn.wqkv"), a_mn=True, b_mn=True, f32_mode=fm))
            ops.gemm(dqkv, self.W(pre + "attn.wqkv"), out=w.dtmp, b_mn=True)
            ops.ln_bwd_partial(w.dtmp, w.resid_a[l], st1[0], st1[1], self.P(pre + "ln1.weight"), w.ln_partial2,
                               dx_in=w.dx, dx_out=w.dx, dx_bf16=w.dxa[(l - 1) % 3])
            ops.ln_bwd_finalize2(w.M, lay.d, w.ln_partial, self.G(pre + "ln2.weight"), self.G(pre + "ln2.bias"),
                                 self.G(pre + "mlp.proj.bias"), w.ln_partial2, self.G(pre + "ln1.weight"),
                                 self.G(pre + "ln1.bias"), None, accumulate=accumulate)
            done = w_done()
            if done is not None:
                pending.append(done)
            if dp is not None:
                dp.bucket_ready(f"block{l}", done)
        for ev in pending[-2:] + ([head_done] if head_done is not None else []):
            main.wait_event(ev)                # join: the optimizer (and the end of a graph capture) follow on the main stream
        if DETERMINISTIC:
            # stable sort of the token ids (torch.sort, capturable) -> runs of equal tokens summed in position order
            order = torch.sort(w.idx.view(-1), stable=True).indices
            ops.embed_bwd_sorted(w.idx, order, w.dx, self.G("token_embed.weight"), self.G("position_embed.weight")[:T],
                                 accumulate=accumulate)
        else:
            ops.embed_bwd(w.idx, w.dx, self.G("token_embed.weight"), self.G("position_embed.weight")[:T],
                          accumulate=accumulate)
        if not accumulate and T < lay.ctx:
            ops.zero_(self.G("position_embed.weight")[T:])
        if z is not None and not dp_sync:
            z.reduce_and_step()                # micro-step of an accumulation: flush the remaining buckets into the
                                               # accumulators before the next micro-step overwrites the windows

    # ------------------------------------------------------------------ optimizer
    def set_hyper(self, lr: float, grad_scale: float = 1.0) -> None:
        """Stage AdamW scalars for the NEXT optimizer launch (also read by graph replays)."""
        self.step_count += 1
        if self.dp is not None and self.dp.vp is not None:
            self.dp.vp.set_hyper(ops.adamw_hyper(lr, self.step_count, self.betas[0], self.betas[1], self.eps,
                                                 self.weight_decay, grad_scale))
        if self.dp is not None:
            grad_scale = grad_scale / self.dp.world       # average of the per-rank mean losses
        vals = ops.adamw_hyper(lr, self.step_count, self.betas[0], self.betas[1], self.eps, self.weight_decay,
                               grad_scale)
        # pageable source: the driver stages the 32 bytes at call time, so later host writes cannot race
        self.hyper.copy_(torch.tensor(vals, dtype=F32))
        self.lr = lr

    def optimizer_step(self) -> None:
        """Fused AdamW over the whole flat buffer (fp32 masters + bf16 shadow) using staged scalars."""
        if self.dp is not None:
            self.dp.reduce_and_step()
            if self.dp.vp is not None:
                self.dp.vp.step()
        else:
            self.ensure_opt_state()
            ops.adamw(self.flat, self.grad, self.exp_avg, self.exp_avg_sq, self.wb, self.hyper)
        # our kernel wrote the masters and the shadow together: keep them marked in sync
        if self.flat is not None:
            self._wb_version = self.flat._version

    def zero_grad(self) -> None:
        if self.grad is not None and self.store is None:
            self.grad.zero_()

    # ------------------------------------------------------------------ whole training step
    def train_step(self, idx: torch.Tensor, targets: torch.Tensor, lr: float, *, use_graph: bool = True
                   ) -> torch.Tensor:
        """forward + backward + (gradient reduction) + AdamW.  Returns the loss as a 0-dim device
        tensor (no host sync).  From the second call with the same shape the step is one CUDA
        graph replay."""
        B, T = idx.shape
        if self.dp is not None and getattr(self.dp, "full_flat", None) is not None:
            self.dp.release_full()         # sharded mode was materialised (generation / checkpoint): back to the shards
        w = self.workspace(B, T, True)
        self.ensure_grad()
        if self.dp is None:
            self.ensure_opt_state()        # with data parallelism the moments live sharded in dp.m / dp.v
        self.sync_weights()
        w.idx.copy_(idx, non_blocking=True)
        w.tgt.copy_(targets.reshape(B, T), non_blocking=True)
        self.set_hyper(lr)
        if use_graph and w.graph is not None:
            w.graph.replay()
        elif use_graph and getattr(w, "warm", False):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._step_body(w)
            w.graph = g
            g.replay()
        else:
            self._step_body(w)
            w.warm = True
        return w.loss[0]

    def _step_body(self, w: Workspace) -> None:
        if self.dp is not None:
            self.dp.begin_step()
        with nvtx.range("forward"):
            self._forward_ws(w, True)
        with nvtx.range("backward"):
            self.backward(w)
        with nvtx.range("optimizer"):
            self.optimizer_step()

    # ------------------------------------------------------------------ nn.Module facade path
    def module_forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor]):
        grad_mode = torch.is_grad_enabled() and targets is not None and self.model.training
        if grad_mode and self.store is not None:
            raise RuntimeError("the autograd (torch-optimizer) path needs full parameters; with dp_mode='zero' use "
                               "Engine.train_step / Trainer")
        idx = idx.contiguous()
        if idx.dtype != torch.int64:
            idx = idx.long()
        if targets is not None:
            targets = targets.contiguous().long()
        if not grad_mode:
            _, logits, loss = self.forward(idx, targets, train=False)
            # the workspace is reused by the next call of the same shape: hand out copies, not views
            return logits.clone(), (loss.clone() if loss is not None else None)
        params = [p for _, p in self.model.named_parameters()]
        loss, logits = _EngineLoss.apply(self, idx, targets, *params)
        return logits, loss

    def forward_embedding(self, idx: torch.Tensor):
        """(post-ReLU hidden of the last block (B,T,4d), its residual input (B,T,d)) on the native kernels - the
        reference's ``Transformer.forward_embedding`` (src/models/transformer.py:80-94), fp32 copies of the bf16 / fp32
        workspace tensors (the inference workspace reuses one slot per layer, so after the sweep it holds the LAST
        block's activations)."""
        if self.store is not None:
            raise RuntimeError("forward_embedding needs full parameters; with dp_mode='zero' call engine.dp.materialize() first")
        idx = idx.contiguous().long()
        B, T = idx.shape
        if T > self.layout.ctx:
            raise IndexError(f"sequence length {T} exceeds context_length {self.layout.ctx}")
        w = self.workspace(B, T, False)
        self.sync_weights()
        w.idx.copy_(idx)
        ops.embed_fwd(w.idx, self.P("token_embed.weight"), self.P("position_embed.weight"), out=w.resid_a[0])
        self._blocks_forward(w)
        d = self.layout.d
        return w.act[0].view(B, T, 4 * d).float(), w.resid_b[0].view(B, T, d).clone()

    # ------------------------------------------------------------------ generation
    def generate(self, idx, max_new_tokens, **kw):
        from .generation import Generator
        if self.dp is not None:
            # vocab-parallel / sharded weights -> a complete local copy for the small-shape decode kernels
            getattr(self.dp, "prepare_generation", self.dp.gather_weights)()
        if self._gen is None:
            self._gen = Generator(self)
        return self._gen.generate(idx, max_new_tokens, **kw)


class _EngineLoss(torch.autograd.Function):
    """One autograd node for the whole model: forward runs the engine's forward schedule, backward
    runs its backward schedule and hands each Parameter a view of the flat gradient buffer."""

    @staticmethod
    def forward(ctx, eng: Engine, idx, targets, *params):
        w, logits, loss = eng.forward(idx, targets, train=True)
        ctx.eng, ctx.w = eng, w
        logits = logits.clone()       # backward turns the workspace logits into dlogits in place
        ctx.mark_non_differentiable(logits)
        return loss.clone(), logits

    @staticmethod
    def backward(ctx, gloss, _glogits):
        eng: Engine = ctx.eng
        eng.backward(ctx.w, grad_scale=float(gloss))
        names = [n for n, _ in eng.model.named_parameters()]
        # Copies, not views of the shared gradient buffer: autograd's AccumulateGrad keeps the tensor it is given as
        # p.grad, and a second backward() without zero_grad (ordinary torch gradient accumulation) would first
        # overwrite the buffer and then add it to itself.  One clone of the flat buffer, sliced per parameter.
        g = eng.grad.clone()
        return (None, None, None) + tuple(eng.layout.view(g, n) for n in names)
