"""Vocab-parallel LM head + cross-entropy across the data-parallel ranks (csrc/comm/vp.cu).

Rank r owns vocabulary rows ``[r*Vs, min(V, (r+1)*Vs))`` of ``lm_head.weight`` / ``lm_head.bias``: it computes
the logits of ALL ranks' tokens against its slice with the tcgen05 GEMM (softmax statistics in the epilogue),
the ranks exchange the per-row statistics / target logits / log-sum-exps over peer-mapped memory, and the LM-head
weight gradient comes out complete per slice - the largest gradient tensor of the model (a quarter of the
GPT-2-small-shaped parameters) is never all-reduced, and its AdamW state is sharded by construction.
"""
from __future__ import annotations

import ctypes
from typing import Dict

import torch
import torch.distributed as dist

from .. import ops
from ..ops import lib
from .symm import SymmBuffer

BF16, F32 = torch.bfloat16, torch.float32


class _VpWorkspace:
    def __init__(self, vp: "VocabParallelHead", M: int):
        eng, W, r = vp.eng, vp.world, vp.rank
        d, dev = eng.layout.d, eng.device
        Mw = M * W
        self.M, self.Mw = M, Mw
        sb = lambda nbytes: SymmBuffer(nbytes, r, W, dev)
        self.hf = sb(M * d * 2)                                   # peers read my final hidden states
        self.tgt = sb(M * 8)
        self.logits = sb(Mw * vp.ldl * 2)                         # peers read target logits of my slice
        self.stats = sb(vp.nt * Mw * 8)
        self.lse_all = sb(Mw * 4)                                 # peers write the lse of their rows
        self.partial = sb(Mw * d * 2)                             # peers read the partial dhf of their rows
        self.hf_all = torch.empty(Mw, d, device=dev, dtype=BF16)
        self.tgt_all = torch.empty(Mw, device=dev, dtype=torch.int64)
        self.stats.view(F32).fill_(float("-inf"))                 # tiles a short slice never writes read as "empty"
        self.stats.view(F32).view(-1, 2)[:, 1] = 0.0
        torch.cuda.synchronize()
        dist.barrier()


class VocabParallelHead:
    def __init__(self, eng, dp):
        self.eng, self.dp = eng, dp
        self.world, self.rank = dp.world, dp.rank
        lay = eng.layout
        V, W = lay.V, self.world
        self.Vs = ((V + W - 1) // W + 7) // 8 * 8
        self.v0 = min(V, self.rank * self.Vs)
        self.n_local = max(0, min(V, self.v0 + self.Vs) - self.v0)
        assert self.n_local > 0, "vocabulary too small for this many ranks"
        self.ldl = self.Vs
        self.nt = 2 * ((self.Vs + 255) // 256)      # CE statistics entries per row (256-column tile x epilogue group)
        self.sig = SymmBuffer(4 * W * 4, self.rank, W, eng.device)
        self.ws: Dict[int, _VpWorkspace] = {}
        d = lay.d
        n = self.n_local * d
        self.m_w = torch.zeros(n, device=eng.device, dtype=F32)
        self.v_w = torch.zeros(n, device=eng.device, dtype=F32)
        self.m_b = torch.zeros((self.n_local + 3) // 4 * 4, device=eng.device, dtype=F32)
        self.v_b = torch.zeros_like(self.m_b)
        self.hyper = torch.zeros(8, device=eng.device, dtype=F32)
        # hand-shake generation: one per forward pass (NOT the optimizer-step epoch - with gradient accumulation several
        # forward/backward passes share a step, and a flag that already equals the epoch would let a rank run ahead)
        self.pass_dev = torch.zeros(1, device=eng.device, dtype=torch.int32)

    # views of the owned slice
    def _w(self, flat):
        return self.eng.layout.view(flat, "lm_head.weight")[self.v0:self.v0 + self.n_local]

    def _b(self, flat):
        return self.eng.layout.view(flat, "lm_head.bias")[self.v0:self.v0 + self.n_local]

    def workspace(self, M: int) -> _VpWorkspace:
        w = self.ws.get(M)
        if w is None:
            w = self.ws[M] = _VpWorkspace(self, M)
        return w

    def _arr(self, sb: SymmBuffer):
        return sb.ptr_array

    # ------------------------------------------------------------------ forward: logits slice + loss
    def forward(self, w, loss_scale: float = 1.0) -> None:
        """w: engine Workspace (w.hf, w.tgt, w.loss filled / to fill)."""
        eng, L = self.eng, lib.load()
        vw = self.workspace(w.M)
        d, V, W, r = eng.layout.d, eng.layout.V, self.world, self.rank
        st = torch.cuda.current_stream().cuda_stream
        ops.incr_int(self.pass_dev, 1)
        ep, err = self.pass_dev.data_ptr(), self.dp.err.data_ptr()
        vw.hf.view(BF16)[:w.M * d].view(w.M, d).copy_(w.hf)
        vw.tgt.view(torch.int64)[:w.M].copy_(w.tgt.view(-1))
        lib.check(L.b200_vp_gather_rows(self._arr(vw.hf), self._arr(vw.tgt), vw.hf_all.data_ptr(), vw.tgt_all.data_ptr(),
                                        w.M, d, W, r, self.v0, self._arr(self.sig), 0, ep, err, st), "vp_gather_rows")
        logits = vw.logits.view(BF16)[:vw.Mw * self.ldl].view(vw.Mw, self.ldl)
        stats = vw.stats.view(F32)
        ops.gemm(vw.hf_all, self._w(eng.wb), out=logits[:, :self.n_local], bias=self._b(eng.flat), stats=stats, bn=256)
        w.loss.zero_()
        lib.check(L.b200_vp_ce_finalize(self._arr(vw.stats), self._arr(vw.logits), self._arr(vw.lse_all),
                                        w.tgt.data_ptr(), w.loss.data_ptr(), w.loss_rows.data_ptr(), w.M, W, r, V, self.Vs,
                                        self.nt, self.ldl, loss_scale / w.M, self._arr(self.sig), 1, ep, err, st),
                  "vp_ce_finalize")
        ops.LAUNCHES += 2

    # ------------------------------------------------------------------ backward: dlogits, dW slice, dhf
    def backward(self, w, grad_scale: float = 1.0, accumulate: bool = False) -> None:
        eng, L = self.eng, lib.load()
        vw = self.workspace(w.M)
        d, W, r = eng.layout.d, self.world, self.rank
        st = torch.cuda.current_stream().cuda_stream
        ep, err = self.pass_dev.data_ptr(), self.dp.err.data_ptr()
        lib.check(L.b200_vp_barrier(self._arr(self.sig), W, r, 2, ep, err, st), "vp_barrier")   # lse_all complete
        logits = vw.logits.view(BF16)[:vw.Mw * self.ldl].view(vw.Mw, self.ldl)[:, :self.n_local]
        lse_all = vw.lse_all.view(F32)[:vw.Mw]
        ops.ce_grad(logits, lse_all, vw.tgt_all, self._b(eng.grad), self.n_local, grad_scale / vw.Mw, accumulate=accumulate)
        ops.gemm(logits, vw.hf_all, out=self._w(eng.grad), a_mn=True, b_mn=True, f32_mode=1 if accumulate else 0)
        partial = vw.partial.view(BF16)[:vw.Mw * d].view(vw.Mw, d)
        ops.gemm(logits, self._w(eng.wb), out=partial, b_mn=True)
        lib.check(L.b200_vp_reduce_rows(self._arr(vw.partial), w.dtmp.data_ptr(), w.M, d, W, r, float(W),
                                        self._arr(self.sig), 3, ep, err, st), "vp_reduce_rows")
        ops.LAUNCHES += 2

    # ------------------------------------------------------------------ optimizer on the owned slice
    def set_hyper(self, vals) -> None:
        v = list(vals)
        v[7] = 1.0                                   # the slice gradient is already the global-mean gradient
        self.hyper.copy_(torch.tensor(v, dtype=F32))

    def step(self) -> None:
        eng = self.eng
        ops.adamw(self._w(eng.flat), self._w(eng.grad), self.m_w, self.v_w, self._w(eng.wb), self.hyper)
        nb = self.n_local
        ops.adamw(self._b(eng.flat), self._b(eng.grad), self.m_b[:nb], self.v_b[:nb], self._b(eng.wb), self.hyper)

    # ------------------------------------------------------------------ full head for checkpoints / generation
    def gather_full(self, with_moments: bool = True) -> None:
        """All ranks end up with the complete lm_head (fp32 masters, bf16 shadow, and moments in eng.exp_avg/_sq)."""
        eng = self.eng
        lay = eng.layout
        torch.cuda.synchronize()

        def ag(local: torch.Tensor, full_view: torch.Tensor, width: int):
            pad = torch.zeros(self.Vs * width, device=eng.device, dtype=local.dtype)
            pad[:local.numel()] = local.reshape(-1)
            parts = [torch.empty_like(pad) for _ in range(self.world)]
            dist.all_gather(parts, pad)
            for rr, part in enumerate(parts):
                v0 = min(lay.V, rr * self.Vs)
                n = max(0, min(lay.V, v0 + self.Vs) - v0)
                if n:
                    full_view[v0:v0 + n] = part[:n * width].view(n, width) if width > 1 else part[:n]

        ag(self._w(eng.flat), lay.view(eng.flat, "lm_head.weight"), lay.d)
        ag(self._b(eng.flat), lay.view(eng.flat, "lm_head.bias"), 1)
        lay.view(eng.wb, "lm_head.weight").copy_(lay.view(eng.flat, "lm_head.weight"))
        lay.view(eng.wb, "lm_head.bias").copy_(lay.view(eng.flat, "lm_head.bias"))
        if with_moments:
            eng.ensure_opt_state()
            ag(self.m_w, lay.view(eng.exp_avg, "lm_head.weight"), lay.d)
            ag(self.v_w, lay.view(eng.exp_avg_sq, "lm_head.weight"), lay.d)
            ag(self.m_b[:self.n_local], lay.view(eng.exp_avg, "lm_head.bias"), 1)
            ag(self.v_b[:self.n_local], lay.view(eng.exp_avg_sq, "lm_head.bias"), 1)
        eng._wb_version = eng.flat._version

    def state_loaded(self) -> None:
        eng, lay = self.eng, self.eng.layout
        self.m_w.copy_(lay.view(eng.exp_avg, "lm_head.weight")[self.v0:self.v0 + self.n_local].reshape(-1))
        self.v_w.copy_(lay.view(eng.exp_avg_sq, "lm_head.weight")[self.v0:self.v0 + self.n_local].reshape(-1))
        self.m_b[:self.n_local].copy_(lay.view(eng.exp_avg, "lm_head.bias")[self.v0:self.v0 + self.n_local])
        self.v_b[:self.n_local].copy_(lay.view(eng.exp_avg_sq, "lm_head.bias")[self.v0:self.v0 + self.n_local])
