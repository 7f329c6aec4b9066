"""Data parallelism with the collective fused into the optimizer kernel.

What torch DDP does for the reference (``DDP(model)`` at reference
``scripts/train_transformer.py:122-123``: bucketed fp32 ``ncclAllReduce`` from autograd hooks, then a
separate ``foreach`` AdamW over every replica's full parameter set) is done here by ONE kernel per
gradient bucket (``csrc/comm/dp.cu``), launched on a side stream the moment backward has finished
the bucket:

    reduce-scatter   each rank pulls its 1/W slice of the bucket from every peer's gradient
                     buffer with P2P loads over NVLink and sums it (fp32, scaled by 1/W)
    AdamW            on that slice only - fp32 master weights and the moments are sharded
                     across ranks (ZeRO-1 style: optimizer state is 1/W per GPU)
    all-gather       the new bf16 weights (fp32 too for tensors consumed in fp32: embeddings,
                     LayerNorm, biases) are pushed into every peer's weight buffers with P2P stores

Cross-rank ordering uses system-scope flags in peer-mapped signal pads; the next forward is gated by
``b200_dp_wait_done``.  Buckets follow backward order: LM head, blocks L-1..0, embeddings.  Every
rank launches the same kernels in the same order; spin loops time out into an error flag.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

from ..models.layout import ParamLayout
from ..ops import lib
from .. import ops

F32, BF16 = torch.float32, torch.bfloat16
_BF16_ONLY = ("attn.wqkv", "mlp.hidden.weight", "mlp.proj.weight", "lm_head.weight")


@dataclass
class BucketPlan:
    name: str
    start: int            # element offset of the bucket in the flat buffer
    end: int
    chunk: int            # elements owned by each rank (last rank may own fewer)
    slot_end: List[int]   # ascending element END offsets of the slots in the bucket
    slot_fp32: List[int]  # 0 bf16 push only | 1 consumers read fp32: push fp32 too | 2 no push (rows read from the owner)

    def slice_of(self, rank: int) -> Tuple[int, int]:
        s = min(self.end, self.start + rank * self.chunk)
        e = min(self.end, s + self.chunk)
        return s, e - s


_OWNER_READ = ("token_embed.weight", "position_embed.weight")


def plan_buckets(layout: ParamLayout, world: int, vocab_parallel: bool = False, shard_embed: bool = False
                 ) -> List[BucketPlan]:
    """Static ownership map: bucket b's elements [start + r*chunk, start + (r+1)*chunk) belong to rank r.
    With a vocab-parallel LM head, lm_head.weight / lm_head.bias are owned by vocabulary slice instead
    (parallel/vp.py) and leave the head bucket."""
    slots = list(layout.slots.values())
    plans = []
    for name, start, end in layout.buckets():
        if vocab_parallel and name == "head":
            end = layout.slots["lm_head.weight"].offset
        n = end - start
        chunk = (n + world - 1) // world
        chunk = (chunk + 3) // 4 * 4
        inside = [s for s in slots if start <= s.offset < end]
        slot_end, slot_fp32 = [], []
        for i, s in enumerate(inside):
            nxt = inside[i + 1].offset if i + 1 < len(inside) else end
            slot_end.append(nxt)
            slot_fp32.append(2 if (shard_embed and s.name in _OWNER_READ) else (0 if s.name.endswith(_BF16_ONLY) else 1))
        plans.append(BucketPlan(name, start, end, chunk, slot_end, slot_fp32))
    return plans


class DataParallel:
    def __init__(self, eng, ctx, mode: str = "ddp", ctas: int = 0, ctas_tail: int = 0, vocab_parallel: bool = False):
        from .symm import SymmBuffer
        if mode != "ddp":
            raise ValueError(f"DataParallel is the replicated mode; dp_mode={mode!r} is parallel.zero.ZeroDataParallel "
                             "(use parallel.make_data_parallel)")
        import os
        # CTAs per reduce kernel.  Measured at W=2 (bench.py, ms/step): 8 -> 23.2, 16 -> 22.0, 32 -> 21.1, 148 -> 20.4:
        # the kernel is bandwidth/latency bound and co-resides with the GEMM CTAs without slowing them much
        self.ctas = ctas or int(os.environ.get("B200_DP_CTAS", "148"))
        self.ctas_tail = ctas_tail or int(os.environ.get("B200_DP_CTAS_TAIL", "148"))
        self.eng, self.ctx, self.mode = eng, ctx, mode
        self.world, self.rank = ctx.world, ctx.rank
        if self.world not in (1, 2, 4, 8):
            raise ValueError("data-parallel world size must be 1, 2, 4 or 8 (one NVSwitch domain)")
        lay: ParamLayout = eng.layout
        dev = eng.device
        n = lay.total
        # embedding tables stay with their owners (embed_fwd reads rows over NVLink) unless B200_DP_SHARD_EMBED=0
        self.shard_embed = self.world > 1 and os.environ.get("B200_DP_SHARD_EMBED", "1") == "1"
        self.plans = plan_buckets(lay, self.world, vocab_parallel, self.shard_embed)
        self.nb = len(self.plans)
        self.index: Dict[str, int] = {p.name: i for i, p in enumerate(self.plans)}
        # ---- symmetric buffers: gradients, bf16 weights, fp32 masters, signal pad
        self.sym_grad = SymmBuffer(n * 4, self.rank, self.world, dev)
        self.sym_wb = SymmBuffer(n * 2, self.rank, self.world, dev)
        self.sym_flat = SymmBuffer(n * 4, self.rank, self.world, dev)
        self.sym_sig = SymmBuffer(2 * self.world * self.nb * 4, self.rank, self.world, dev, multicast=False)
        # NVSwitch multicast mappings of the three buffers (all or none): the reduce kernel then pulls with
        # multimem.ld_reduce and pushes with multimem.st
        # Measured on 8 x B200 (GPT-2-small-shaped step): 16.88 ms with multimem vs 17.35 ms with per-peer loads/stores;
        # at 2 GPUs the switch adds latency for nothing to reduce (17.03 vs 16.64 ms): default on from 4 ranks.
        mcs = (self.sym_grad.mc_ptr, self.sym_wb.mc_ptr, self.sym_flat.mc_ptr)
        want = os.environ.get("B200_DP_MULTIMEM", "auto")
        self.multicast = all(mcs) and (want == "1" or (want == "auto" and self.world >= 4))
        self.mc_ptrs = tuple(ctypes.c_void_p(p) for p in mcs) if self.multicast else (None, None, None)
        new_flat = self.sym_flat.view(F32)[:n]
        new_flat.copy_(eng.flat)
        if self.world > 1:
            dist.broadcast(new_flat, src=0)               # identical replicas to start from
        model = eng.model
        with torch.no_grad():
            for name, p in model.named_parameters():
                p.data = lay.view(new_flat, name)
        model._flat = new_flat
        eng.flat = new_flat
        eng.wb = self.sym_wb.view(BF16)[:n]
        eng.grad = self.sym_grad.view(F32)[:n]
        eng.params_changed()
        eng.sync_weights()
        # ---- sharded optimizer state: only the owned slices
        self.owned = [p.slice_of(self.rank) for p in self.plans]
        total_owned = sum(cnt for _, cnt in self.owned)
        self.m = torch.zeros(max(total_owned, 4), device=dev, dtype=F32)
        self.v = torch.zeros(max(total_owned, 4), device=dev, dtype=F32)
        self.opt_off = []
        off = 0
        for _, cnt in self.owned:
            self.opt_off.append(off)
            off += cnt
        self.zero = False
        self.counter = torch.zeros(self.nb, device=dev, dtype=torch.int32)
        self.err = torch.zeros(1, device=dev, dtype=torch.int32)
        self.epoch_dev = torch.zeros(1, device=dev, dtype=torch.int32)   # incremented on the device each step
        self.comm_stream = torch.cuda.Stream(device=dev)
        self._launched = set()
        self._slot_arrays = []
        for p in self.plans:
            k = len(p.slot_end)
            self._slot_arrays.append(((ctypes.c_longlong * k)(*p.slot_end), (ctypes.c_int * k)(*p.slot_fp32), k))
        # rank r's fp32 master slice of the embedding bucket, as every rank addresses it
        pe = self.plans[self.index["embed"]]
        self._emb_array = (ctypes.c_void_p * self.world)(*[self.sym_flat.ptrs[r] + 4 * (pe.start + r * pe.chunk)
                                                            for r in range(self.world)])
        self.vp = None
        if vocab_parallel and self.world > 1:
            from .vp import VocabParallelHead
            self.vp = VocabParallelHead(eng, self)
        if eng.exp_avg is not None:        # moments loaded before the ranks were attached (resume): keep the owned slices
            self.state_loaded()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ per-step protocol
    def bucket_ready(self, name: str, also=None) -> None:
        """Called by Engine.backward right after the bucket's last gradient kernel was enqueued."""
        b = self.index[name]
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        if also is not None:                   # the bucket's weight gradients are produced on their own stream
            self.comm_stream.wait_event(also)
        self._launch(b)

    def _launch(self, b: int, tail: bool = False) -> None:
        start, cnt = self.owned[b]
        end_arr, fp_arr, k = self._slot_arrays[b]
        off = self.opt_off[b]
        L = lib.load()
        rc = L.b200_dp_reduce_adamw(self.sym_grad.ptr_array, self.sym_wb.ptr_array, self.sym_flat.ptr_array,
                                    self.sym_sig.ptr_array, self.world, self.rank, b, self.nb, self.epoch_dev.data_ptr(), start, cnt,
                                    self.m[off:].data_ptr(), self.v[off:].data_ptr(), self.eng.hyper.data_ptr(),
                                    end_arr, fp_arr, k, self.counter[b:].data_ptr(), self.err.data_ptr(),
                                    self.ctas_tail if tail else self.ctas,
                                    *self.mc_ptrs, self.comm_stream.cuda_stream)
        lib.check(rc, "dp_reduce_adamw")
        ops.LAUNCHES += 1
        self._launched.add(b)

    def reduce_and_step(self) -> None:
        """End of the step: reduce whatever has not been launched yet (the embedding bucket), then gate
        the main stream on every owner having delivered every bucket."""
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        for b in range(self.nb):
            if b not in self._launched:
                self._launch(b, tail=True)
        done = torch.cuda.Event()
        done.record(self.comm_stream)
        main.wait_event(done)
        rc = lib.load().b200_dp_wait_done(self.sym_sig.view(torch.int32).data_ptr(), self.world, self.nb,
                                          self.epoch_dev.data_ptr(), self.err.data_ptr(), main.cuda_stream)
        lib.check(rc, "dp_wait_done")
        ops.LAUNCHES += 1
        self._launched = set()

    def gather_weights(self) -> None:
        """Make the local weights complete: vocab-parallel LM head slices, and the fp32 embedding tables when they are
        kept with their owners (off the hot path: before generation, evaluation with the plain kernels, checkpoints)."""
        if self.vp is not None:
            self.vp.gather_full(with_moments=False)
        if self.shard_embed:
            torch.cuda.synchronize()
            b = self.index["embed"]
            self._gather_bucket(b, self.eng.flat[self.owned[b][0]:self.owned[b][0] + self.owned[b][1]], self.eng.flat)

    def embed_fwd(self, idx: torch.Tensor, out: torch.Tensor) -> None:
        """Token + position embedding with every fp32 row read from the rank that owns (and updates) it."""
        lay = self.eng.layout
        B, T = idx.shape
        p = self.plans[self.index["embed"]]
        rc = lib.load().b200_embed_fwd_sharded(idx.data_ptr(), self._emb_array, self.world, p.chunk,
                                               lay.slots["token_embed.weight"].offset - p.start,
                                               lay.slots["position_embed.weight"].offset - p.start, out.data_ptr(), B, T,
                                               lay.d, lay.V, torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "embed_fwd_sharded")
        ops.LAUNCHES += 1

    def begin_step(self) -> None:
        """First thing of a step (main stream): advance the device-side epoch."""
        ops.incr_int(self.epoch_dev, 1)
        self._launched = set()

    def check_errors(self) -> None:
        if int(self.err.item()) != 0:
            raise RuntimeError("data-parallel kernel timed out waiting for a peer (a rank died or diverged)")

    # ------------------------------------------------------------------ full state for checkpoints
    def _gather_bucket(self, b: int, loc: torch.Tensor, out_flat: torch.Tensor) -> None:
        p = self.plans[b]
        pad = torch.zeros(p.chunk, device=loc.device, dtype=loc.dtype)
        pad[:loc.numel()] = loc
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        if self.world > 1:
            dist.all_gather(parts, pad)
        else:
            parts = [pad]
        for r, part in enumerate(parts):
            s, cnt = p.slice_of(r)
            out_flat[s:s + cnt] = part[:cnt]

    def _gather(self, local_slices: List[torch.Tensor], out_flat: torch.Tensor) -> None:
        for b, loc in enumerate(local_slices):
            self._gather_bucket(b, loc, out_flat)

    def gather_full_state(self) -> None:
        """Make fp32 masters complete on every rank and materialise full moments in eng.exp_avg(_sq).
        Collective: every rank must call it."""
        eng = self.eng
        torch.cuda.synchronize()
        self._gather([eng.flat[s:s + c] for s, c in self.owned], eng.flat)
        eng.ensure_opt_state()
        self._gather([self.m[o:o + c] for o, (_, c) in zip(self.opt_off, self.owned)], eng.exp_avg)
        self._gather([self.v[o:o + c] for o, (_, c) in zip(self.opt_off, self.owned)], eng.exp_avg_sq)
        if self.vp is not None:
            self.vp.gather_full(with_moments=True)
        eng._wb_version = eng.flat._version

    def state_loaded(self) -> None:
        """Full moments were loaded into eng.exp_avg(_sq) (resume): keep the owned slices."""
        eng = self.eng
        if getattr(self, "vp", None) is not None:
            self.vp.state_loaded()
        for o, (s, c) in zip(self.opt_off, self.owned):
            self.m[o:o + c] = eng.exp_avg[s:s + c]
            self.v[o:o + c] = eng.exp_avg_sq[s:s + c]
        eng.release_full_opt_state()
