"""Sharded data parallelism (``dp_mode="zero"``): parameters, gradients and optimizer state are sharded for real.

What every rank keeps (P = parameters, W = ranks):

    fp32 masters, AdamW moments   owned 1/W slice of every bucket          12 P / W bytes
    bf16 weights                  owned 1/W slice (the "shard")              2 P / W bytes
    gathered bf16 weights         a rolling window of TWO layers + the LM head
    fp32 gradients                a rolling window of TWO layers + LM-head and embedding gradients
    LayerNorm weights / biases    a small replicated fp32 store (they are consumed in fp32 by every rank)

and nothing full-size: no flat fp32 master buffer, no flat bf16 shadow, no flat gradient buffer (the replicated
mode of ``parallel/dp.py`` holds 10 P bytes of those per rank).

Data movement, all inside our own kernels over peer-mapped / multicast memory (``parallel/symm.py``):

  * **all-gather fused into the GEMMs** (``b200_gemm_ag``, csrc/kernels/gemm2.cu): the first GEMM of a layer carries
    the gather of the NEXT layer's weights in four extra warps per CTA - forward: layer l's QKV GEMM pulls layer l+1
    (the last layer's pulls the LM head); backward: layer l's MLP-up dgrad GEMM pulls layer l-1 - so the transfer
    runs under that GEMM's math and is complete when the next layer starts.  Layer 0's forward GEMM also pulls its own
    weights and lets the tiles whose rows are already local run first.  The windows of layers L-1 and L-2 and the LM
    head still hold the forward's copy when backward starts, so they are not gathered twice.
  * **reduce-scatter + AdamW fused** (``b200_zero_reduce_adamw``, csrc/comm/zero.cu): one kernel per bucket on the
    communication stream as soon as backward has produced the bucket; the pull is a ``multimem.ld_reduce`` (summed in
    the NVSwitch) when the window is bound to a multicast object, W peer loads otherwise; the result never exists
    outside the owner's slice.
  * **embedding rows** are read from their owners' fp32 master shards over NVLink (``b200_embed_fwd_sharded``).

Replaces torch DDP's replicated parameters + Reducer + ncclAllReduce + separate AdamW (reference
scripts/train_transformer.py:122-123,92,126).
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from .. import ops
from ..models.layout import ParamLayout
from ..ops import lib
from .symm import SymmBuffer

F32, BF16 = torch.float32, torch.bfloat16
_NOT_SMALL = -(1 << 62)
_BIG_FP32 = ("token_embed.weight", "position_embed.weight")          # fp32-consumed but sharded (read from the owner)
_MATRIX = ("attn.wqkv", "mlp.hidden.weight", "mlp.proj.weight", "lm_head.weight")


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class ZeroPlan:
    """Static ownership map of the sharded mode (pure Python, unit-tested on CPU).

    Buckets follow backward order: ``head`` (final LayerNorm + LM head), ``block{L-1}`` .. ``block0``, ``embed``.
    Bucket b covers flat elements [start, end); rank r owns [start + r*chunk, start + (r+1)*chunk) (clipped), and its
    slice lives at element ``soff[b]`` of every per-rank shard buffer (the same offset on every rank, so a peer's
    chunk is ``peer_shard_base + soff[b]``)."""

    def __init__(self, layout: ParamLayout, world: int):
        self.layout, self.world = layout, world
        self.names: List[str] = []
        self.start: List[int] = []
        self.end: List[int] = []
        self.chunk: List[int] = []
        self.soff: List[int] = []
        off = 0
        for name, s, e in layout.buckets():
            c = _round_up((e - s + world - 1) // world, 128)
            self.names.append(name)
            self.start.append(s)
            self.end.append(e)
            self.chunk.append(c)
            self.soff.append(off)
            off += c
        self.shard_elems = off
        self.index: Dict[str, int] = {n: i for i, n in enumerate(self.names)}
        self.nb = len(self.names)
        # small replicated fp32 store: every fp32-consumed parameter except the embedding tables
        self.small_off: Dict[str, int] = {}
        cur = 0
        for sl in layout.slots.values():
            if sl.name.endswith(_MATRIX) or sl.name in _BIG_FP32:
                continue
            self.small_off[sl.name] = cur
            cur = _round_up(cur + sl.numel, 128)
        self.small_elems = max(cur, 128)
        self.max_layer = max((e - s) for s, e in layout.layer_ranges) if layout.layer_ranges else 0

    def size(self, b: int) -> int:
        return self.end[b] - self.start[b]

    def owned(self, b: int, rank: int) -> Tuple[int, int]:
        """(offset inside the bucket, element count) of rank's slice."""
        lo = min(self.size(b), rank * self.chunk[b])
        hi = min(self.size(b), lo + self.chunk[b])
        return lo, hi - lo

    def block_bucket(self, layer: int) -> int:
        return self.index[f"block{layer}"]

    def slot_tables(self, b: int):
        """(end offsets relative to the bucket, small-store offset minus slot start, or _NOT_SMALL) of the bucket's slots."""
        s, e = self.start[b], self.end[b]
        inside = [sl for sl in self.layout.slots.values() if s <= sl.offset < e]
        ends, smalls = [], []
        for i, sl in enumerate(inside):
            nxt = inside[i + 1].offset if i + 1 < len(inside) else e
            ends.append(nxt - s)
            smalls.append(self.small_off[sl.name] - (sl.offset - s) if sl.name in self.small_off else _NOT_SMALL)
        return ends, smalls

    def bytes_per_rank(self) -> Dict[str, int]:
        """Parameter / gradient / optimizer bytes one rank holds (the memory claim of the mode)."""
        lay = self.layout
        head = lay.head_range[1] - lay.head_range[0]
        emb = lay.embed_range[1] - lay.embed_range[0]
        return {"master+moments": 12 * self.shard_elems, "bf16_shard": 2 * self.shard_elems,
                "weight_windows": 2 * 2 * self.max_layer + 2 * head, "grad_windows": 2 * 4 * self.max_layer,
                "head_grad": 4 * head, "embed_grad": 4 * emb, "small_fp32": 4 * self.small_elems}


class ZeroDataParallel:
    zero = True
    vp = None

    def __init__(self, eng, ctx, ctas: int = 0):
        self.eng, self.ctx = eng, ctx
        self.world, self.rank = ctx.world, ctx.rank
        if self.world not in (1, 2, 4, 8):
            raise ValueError("data-parallel world size must be 1, 2, 4 or 8 (one NVSwitch domain)")
        lay: ParamLayout = eng.layout
        if lay.L < 2:
            raise ValueError("sharded mode needs at least two transformer blocks (two-layer rolling windows)")
        self.mode = "zero"
        self.ctas = ctas or int(os.environ.get("B200_DP_CTAS", "148"))
        dev = eng.device
        W, r = self.world, self.rank
        self.plan = plan = ZeroPlan(lay, W)
        self.nb, self.index = plan.nb, plan.index
        sb = lambda nbytes, mc=False: SymmBuffer(nbytes, r, W, dev, multicast=mc)
        S = plan.shard_elems
        # ---- persistent, sharded
        self.sym_master = sb(S * 4)                    # peers read embedding rows out of it
        self.sym_shard16 = sb(S * 2)                   # peers' all-gather reads it
        self.master = self.sym_master.view(F32)[:S]
        self.shard16 = self.sym_shard16.view(BF16)[:S]
        self.m = torch.zeros(S, device=dev, dtype=F32)
        self.v = torch.zeros(S, device=dev, dtype=F32)
        # ---- rolling windows
        head = plan.size(plan.index["head"])
        emb = plan.size(plan.index["embed"])
        self.win_w = torch.empty(2, plan.max_layer, device=dev, dtype=BF16)
        self.head_w = torch.empty(head, device=dev, dtype=BF16)
        self.sym_gwin = [sb(plan.max_layer * 4, True), sb(plan.max_layer * 4, True)]
        self.sym_head_g = sb(head * 4, True)
        self.sym_embed_g = sb(emb * 4, True)
        self.sym_small = sb(plan.small_elems * 4, True)
        self.small = self.sym_small.view(F32)[:plan.small_elems]
        self.sym_sig = sb(2 * W * self.nb * 4)
        mcs = [b.mc_ptr for b in (*self.sym_gwin, self.sym_head_g, self.sym_embed_g, self.sym_small)]
        want = os.environ.get("B200_DP_MULTIMEM", "auto")      # same policy as the replicated mode: on from 4 ranks
        self.multicast = all(mcs) and (want == "1" or (want == "auto" and W >= 4))
        # ---- device-side protocol state
        z = lambda *s: torch.zeros(*s, device=dev, dtype=torch.int32)
        self.flags, self.counters = z(self.nb, W), z(self.nb, W)
        self.gen_dev, self.epoch_dev, self.err, self.cta_counter = z(1), z(1), z(1), z(self.nb)
        self.comm_stream = torch.cuda.Stream(device=dev)
        self._launched = set()
        self._slot_arrays = []
        for b in range(self.nb):
            ends, smalls = plan.slot_tables(b)
            k = len(ends)
            self._slot_arrays.append(((ctypes.c_longlong * k)(*ends), (ctypes.c_longlong * k)(*smalls), k))
        self._src_arrays = [self._peer_array(self.sym_shard16, plan.soff[b] * 2) for b in range(self.nb)]
        eb = plan.index["embed"]
        self._emb_array = self._peer_array(self.sym_master, plan.soff[eb] * 4)
        self._small_array = self.sym_small.ptr_array
        self.full_flat: Optional[torch.Tensor] = None     # only while materialised (checkpoint, generation)
        self.gacc: Optional[torch.Tensor] = None          # fp32 accumulator over micro-steps (1/W of the model), on demand
        self._accum_only = False
        # ---- take over the parameters: identical replicas from rank 0, then keep only the owned slices
        if W > 1:
            dist.broadcast(eng.flat, src=0)
        self._shard_from_full(eng.flat)
        self._drop_full()
        eng.store = self
        if eng.exp_avg is not None:
            self.state_loaded()
        if W > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ helpers
    def _peer_array(self, sbuf: SymmBuffer, byte_off: int):
        return (ctypes.c_void_p * self.world)(*[p + byte_off for p in sbuf.ptrs])

    def _grad_buf(self, b: int) -> SymmBuffer:
        name = self.plan.names[b]
        if name == "head":
            return self.sym_head_g
        if name == "embed":
            return self.sym_embed_g
        return self.sym_gwin[int(name[5:]) % 2]

    def _shard_from_full(self, flat: torch.Tensor) -> None:
        plan = self.plan
        for b in range(self.nb):
            lo, cnt = plan.owned(b, self.rank)
            s = plan.start[b] + lo
            o = plan.soff[b]
            self.master[o:o + cnt].copy_(flat[s:s + cnt])
            self.shard16[o:o + cnt].copy_(flat[s:s + cnt])
        lay = self.eng.layout
        for name, off in plan.small_off.items():
            sl = lay.slots[name]
            self.small[off:off + sl.numel].copy_(flat[sl.offset:sl.offset + sl.numel])

    def _drop_full(self) -> None:
        """Release every full-size buffer: the module's Parameters become empty placeholders."""
        eng = self.eng
        with torch.no_grad():
            for _, p in eng.model.named_parameters():
                p.data = torch.empty(0, device=eng.device, dtype=F32)
        eng.model._flat = None
        eng.flat = None
        eng.wb = None
        eng.grad = None
        self.full_flat = None
        eng.store = self

    # ------------------------------------------------------------------ parameter store face (Engine.P / W / G)
    def P(self, name: str) -> torch.Tensor:
        off = self.plan.small_off.get(name)
        if off is None:
            raise KeyError(f"{name} is sharded in zero mode; it has no full fp32 view")
        sl = self.eng.layout.slots[name]
        return self.small[off:off + sl.numel].view(sl.shape)

    def _win_view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        lay = self.eng.layout
        sl = lay.slots[name]
        if sl.layer == lay.L:
            base = lay.head_range[0]
        elif sl.layer < 0:
            base = lay.embed_range[0]
        else:
            base = lay.layer_ranges[sl.layer][0]
        o = sl.offset - base
        return buf[o:o + sl.numel].view(sl.shape)

    def W(self, name: str) -> torch.Tensor:
        sl = self.eng.layout.slots[name]
        if sl.layer == self.eng.layout.L:
            return self._win_view(self.head_w, name)
        return self._win_view(self.win_w[sl.layer % 2], name)

    def G(self, name: str) -> torch.Tensor:
        sl = self.eng.layout.slots[name]
        lay = self.eng.layout
        if sl.layer == lay.L:
            buf = self.sym_head_g
        elif sl.layer < 0:
            buf = self.sym_embed_g
        else:
            buf = self.sym_gwin[sl.layer % 2]
        return self._win_view(buf.view(F32), name)

    # ------------------------------------------------------------------ per-pass / per-layer protocol
    def begin_step(self) -> None:
        ops.incr_int(self.epoch_dev, 1)
        self._launched = set()

    def begin_backward(self, accumulate: bool, dp_sync: bool) -> None:
        """Gradient accumulation: the rolling windows cannot hold a running sum, so every micro-step is reduce-scattered
        as usual and the owner adds its slice to a 1/W-sized fp32 accumulator; the last micro-step (dp_sync) picks the
        accumulator up and runs AdamW.  Each micro-step is its own hand-shake epoch."""
        if accumulate:
            ops.incr_int(self.epoch_dev, 1)
            self._launched = set()
        self._accum_only = not dp_sync
        if (accumulate or not dp_sync) and self.gacc is None:
            self.gacc = torch.zeros(self.plan.shard_elems, device=self.eng.device, dtype=F32)
        self.begin_pass()

    def begin_pass(self) -> None:
        """Start of a forward or a backward pass: a new gather generation."""
        ops.incr_int(self.gen_dev, 1)

    def _job(self, b: int, dst: torch.Tensor):
        plan = self.plan
        return (self._src_arrays[b], dst.data_ptr(), plan.chunk[b], plan.size(b), self.flags[b].data_ptr(),
                self.counters[b].data_ptr())

    _NOJOB = (None, None, 0, 0, None, None)

    def _gemm_ag(self, a, bw, out, b_mn, jobs, wait=None, bias=None, relu=False, relu_bits_out=None) -> None:
        if b_mn:
            K, N = bw.shape
        else:
            N, K = bw.shape
        M = a.shape[0]
        j0 = jobs[0] if len(jobs) > 0 else self._NOJOB
        j1 = jobs[1] if len(jobs) > 1 else self._NOJOB
        wf, wc, wo = wait if wait is not None else (None, 0, 0)
        rc = lib.load().b200_gemm_ag(a.data_ptr(), bw.data_ptr(), out.data_ptr(), M, N, K, a.stride(0), bw.stride(0),
                                     out.stride(0), int(b_mn), None if bias is None else bias.data_ptr(), int(relu),
                                     None if relu_bits_out is None else relu_bits_out.data_ptr(), None,
                                     wf, wc, wo, self.world, self.rank, self.gen_dev.data_ptr(), self.err.data_ptr(),
                                     len(jobs), *j0, *j1, torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "gemm_ag")
        ops.LAUNCHES += 1

    def gemm_fwd(self, layer: int, a: torch.Tensor, out: torch.Tensor) -> None:
        """QKV GEMM of `layer` + the all-gather of the NEXT layer's weights (and, for layer 0, of its own)."""
        lay = self.eng.layout
        plan = self.plan
        name = f"attn_blocks.{layer}.attn.wqkv"
        bw = self.W(name)
        jobs, wait = [], None
        b = plan.block_bucket(layer)
        if layer == 0:
            jobs.append(self._job(b, self.win_w[0]))
            wait = (self.flags[b].data_ptr(), plan.chunk[b], lay.slots[name].offset - lay.layer_ranges[0][0])
        if layer + 1 < lay.L:
            jobs.append(self._job(plan.block_bucket(layer + 1), self.win_w[(layer + 1) % 2]))
        else:
            jobs.append(self._job(plan.index["head"], self.head_w))
        self._gemm_ag(a, bw, out, False, jobs, wait)

    def gemm_bwd(self, layer: int, a: torch.Tensor, out: torch.Tensor) -> None:
        """MLP-up dgrad GEMM of `layer` (dx = dh @ W1) + the all-gather of layer-1's weights for ITS backward.
        Layers L-1 and L-2 still sit in the windows from the forward pass."""
        lay = self.eng.layout
        bw = self.W(f"attn_blocks.{layer}.mlp.hidden.weight")
        if 1 <= layer <= lay.L - 2:
            jobs = [self._job(self.plan.block_bucket(layer - 1), self.win_w[(layer - 1) % 2])]
            self._gemm_ag(a, bw, out, True, jobs)
        else:
            ops.gemm(a, bw, out=out, b_mn=True)

    def embed_fwd(self, idx: torch.Tensor, out: torch.Tensor) -> None:
        lay = self.eng.layout
        B, T = idx.shape
        eb = self.plan.index["embed"]
        base = lay.embed_range[0]
        rc = lib.load().b200_embed_fwd_sharded(idx.data_ptr(), self._emb_array, self.world, self.plan.chunk[eb],
                                               lay.slots["token_embed.weight"].offset - base,
                                               lay.slots["position_embed.weight"].offset - base, out.data_ptr(), B, T,
                                               lay.d, lay.V, torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "embed_fwd_sharded")
        ops.LAUNCHES += 1

    def before_layer_bwd(self, layer: int) -> None:
        """Layer `layer` is about to overwrite the gradient window that layer+2 used: every rank's reduce of that
        bucket must have finished reading it."""
        if layer + 2 <= self.eng.layout.L - 1:
            b = self.plan.block_bucket(layer + 2)
            self._wait(b, b + 1)

    def _wait(self, b_lo: int, b_hi: int) -> None:
        rc = lib.load().b200_zero_wait(self.sym_sig.view(torch.int32).data_ptr(), self.world, self.nb, b_lo, b_hi,
                                       self.epoch_dev.data_ptr(), 0, self.err.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "zero_wait")
        ops.LAUNCHES += 1

    def bucket_ready(self, name: str, also=None) -> None:
        b = self.index[name]
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        if also is not None:                   # the bucket's weight gradients are produced on their own stream
            self.comm_stream.wait_event(also)
        self._launch(b)

    def _launch(self, b: int) -> None:
        plan = self.plan
        lo, cnt = plan.owned(b, self.rank)
        o = plan.soff[b]
        gb = self._grad_buf(b)
        ends, smalls, k = self._slot_arrays[b]
        mc = self.multicast
        rc = lib.load().b200_zero_reduce_adamw(
            gb.ptr_array, ctypes.c_void_p(gb.mc_ptr) if mc else None, self._small_array,
            ctypes.c_void_p(self.sym_small.mc_ptr) if mc else None, self.sym_sig.ptr_array, self.world, self.rank, b,
            self.nb, self.epoch_dev.data_ptr(), lo, cnt, self.master[o:].data_ptr(), self.m[o:].data_ptr(),
            self.v[o:].data_ptr(), self.shard16[o:].data_ptr(), self.eng.hyper.data_ptr(), ends, smalls, k,
            self.cta_counter[b:].data_ptr(), self.err.data_ptr(), self.ctas,
            None if self.gacc is None else self.gacc[o:].data_ptr(), int(self._accum_only), self.comm_stream.cuda_stream)
        lib.check(rc, "zero_reduce_adamw")
        ops.LAUNCHES += 1
        self._launched.add(b)

    def reduce_and_step(self) -> None:
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        for b in range(self.nb):
            if b not in self._launched:
                self._launch(b)
        done = torch.cuda.Event()
        done.record(self.comm_stream)
        main.wait_event(done)
        self._wait(0, self.nb)
        self._launched = set()

    def check_errors(self) -> None:
        if int(self.err.item()) != 0:
            raise RuntimeError("data-parallel kernel timed out waiting for a peer (a rank died or diverged)")

    # ------------------------------------------------------------------ full state (checkpoints, generation, resume)
    def _gather_into(self, local: torch.Tensor, out_flat: torch.Tensor) -> None:
        plan = self.plan
        for b in range(self.nb):
            c, o = plan.chunk[b], plan.soff[b]
            if self.world > 1:
                parts = torch.empty(self.world * c, device=local.device, dtype=local.dtype)
                dist.all_gather_into_tensor(parts, local[o:o + c].contiguous())
            else:
                parts = local[o:o + c]
            n = plan.size(b)
            out_flat[plan.start[b]:plan.end[b]].copy_(parts[:n])

    def materialize(self, with_moments: bool = False) -> None:
        """Collective.  Rebuild the full flat fp32 parameters (and moments) on every rank and point the nn.Module at
        them: checkpoints, generation and evaluation with small shapes run on the replicated code path meanwhile.
        ``release_full()`` drops the copies again; the shards stay authoritative."""
        eng = self.eng
        torch.cuda.synchronize()
        lay = eng.layout
        if self.full_flat is None:
            full = torch.zeros(lay.total, device=eng.device, dtype=F32)
            self._gather_into(self.master, full)
            self.full_flat = full
            with torch.no_grad():
                for name, p in eng.model.named_parameters():
                    p.data = lay.view(full, name)
            eng.model._flat = full
            eng.flat = full
            eng.wb = torch.empty(lay.total, device=eng.device, dtype=BF16)
            eng.store = None
            eng.params_changed()
            eng.sync_weights()
        if with_moments:
            eng.ensure_opt_state()
            self._gather_into(self.m, eng.exp_avg)
            self._gather_into(self.v, eng.exp_avg_sq)

    def release_full(self) -> None:
        if self.full_flat is not None:
            self._drop_full()

    def gather_full_state(self) -> None:
        self.materialize(with_moments=True)

    def gather_weights(self) -> None:
        """Evaluation through Engine.forward runs sharded; nothing to do."""

    def prepare_generation(self) -> None:
        self.materialize()

    def state_loaded(self) -> None:
        """A checkpoint was loaded into the (materialised) full buffers: re-shard and drop them."""
        eng = self.eng
        if self.full_flat is not None:
            self._shard_from_full(self.full_flat)
        if eng.exp_avg is not None:
            plan = self.plan
            for b in range(self.nb):
                lo, cnt = plan.owned(b, self.rank)
                s, o = plan.start[b] + lo, plan.soff[b]
                self.m[o:o + cnt].copy_(eng.exp_avg[s:s + cnt])
                self.v[o:o + cnt].copy_(eng.exp_avg_sq[s:s + cnt])
            eng.release_full_opt_state()
        self.release_full()
        if self.world > 1:                 # the peers' all-gathers read these shards: nobody runs ahead of the re-shard
            torch.cuda.synchronize()
            dist.barrier()

    def memory_report(self) -> Dict[str, float]:
        rep = {k: v / 2 ** 30 for k, v in self.plan.bytes_per_rank().items()}
        rep["total_GiB"] = sum(rep.values())
        return rep
