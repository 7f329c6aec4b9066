"""Python face of the native kernels: thin, allocation-light wrappers around the C ABI.

Every wrapper enqueues on ``torch.cuda.current_stream()``, does no host synchronisation and
raises if the native library reports an error.  Output buffers may be passed in (the engine
preallocates everything so a training step is CUDA-graph capturable); when omitted they are
allocated with torch.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch

from . import lib
from ..utils import nvtx as _nvtx

BF16 = torch.bfloat16
F32 = torch.float32


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


LAUNCHES = 0   # kernels of this library enqueued so far (bench.py reports the count per timed region)

_KERNELS_PER_CALL = {"embed_bwd": 2, "ln_bwd": 2, "attn_bwd": 3, "attn_bwd_tc": 3}


def _chk(rc: int, what: str) -> None:
    global LAUNCHES
    if rc != 0:
        lib.check(rc, what)
    LAUNCHES += _KERNELS_PER_CALL.get(what, 1)
    if _nvtx.ENABLED:              # B200_NVTX=1: one NVTX marker per launch, named after the kernel family
        torch.cuda.nvtx.mark(what)


# ------------------------------------------------------------------------------------------ GEMM
GEMM_2CTA = os.environ.get("B200_GEMM_2CTA", "1") == "1"   # CTA-pair kernel for large problems (validated on B200)


def gemm(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, *, a_mn: bool = False,
         b_mn: bool = False, out_dtype=BF16, f32_mode: int = 0, alpha: float = 1.0,
         bias: Optional[torch.Tensor] = None, relu: bool = False, mask: Optional[torch.Tensor] = None,
         stats: Optional[torch.Tensor] = None, bn: int = 0, split_k: int = 0,
         two_cta: Optional[bool] = None, relu_bits_out: Optional[torch.Tensor] = None,
         relu_bits: Optional[torch.Tensor] = None, colsum_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C[M,N] = alpha * opA @ opB^T on the tcgen05 kernel.

    a: [M,K] (a_mn=False) or [K,M] (a_mn=True); b: [N,K] (b_mn=False) or [K,N] (b_mn=True); bf16,
    last dim contiguous.  out bf16 (epilogue: +bias, ReLU, *(mask>0), softmax stats) or fp32
    (f32_mode 0 store / 1 accumulate / 2 atomic+split-K).
    relu_bits_out / relu_bits: int64 [M, ceil(N/64)] sign bits of the ReLU output, written by the forward GEMM and
    applied by the matching dgrad GEMM (the fast replacement of mask=activations).
    colsum_out: fp32 [N] that receives += the column sums of the bf16 output (the bias gradient of the Linear whose output
    gradient a dgrad GEMM produces).  Fused into the epilogue of the CTA-pair kernel for the dgrad form (a K-major, b_mn,
    bf16 out); any other form falls back to the separate column-sum kernel."""
    assert a.dtype == BF16 and b.dtype == BF16 and a.stride(-1) == 1 and b.stride(-1) == 1
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"inner dimensions differ: {K} vs {Kb}"
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
        if out_dtype == F32 and f32_mode != 0:
            out.zero_()
    assert out.shape == (M, N) and out.stride(-1) == 1
    out_f32 = out.dtype == F32
    assert out_f32 or out.dtype == BF16
    if stats is not None:
        assert bn in (0, 256), "the CE-statistics epilogue is used with 256-column tiles"
        assert stats.dtype == F32 and stats.is_contiguous() and stats.numel() >= gemm_stats_entries(N) * M * 2, \
            f"stats needs {gemm_stats_entries(N)} x {M} x 2 floats (see ops.gemm_stats_entries)"
    for t in (relu_bits_out, relu_bits):
        assert t is None or (t.dtype == torch.int64 and t.is_contiguous() and t.shape == (M, (N + 63) // 64))
    L = lib.load()
    if two_cta is None:
        two_cta = GEMM_2CTA and M >= 512 and N >= 256 and bn in (0, 128, 256)
    if colsum_out is not None:
        assert colsum_out.dtype == F32 and colsum_out.is_contiguous() and colsum_out.numel() == N and not out_f32
        fused = (two_cta and not a_mn and b_mn and bn in (0, 256) and mask is None and stats is None and FUSED_COLSUM)
        if fused:
            rc = L.b200_gemm_2cta_colsum(_p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), out.stride(0), 0, 1,
                                         float(alpha), _p(bias), int(relu), _p(relu_bits_out), _p(relu_bits), 256,
                                         _p(colsum_out), _st())
            _chk(rc, "gemm")
            return out
    fn = L.b200_gemm_2cta if two_cta else L.b200_gemm
    if two_cta and bn == 0:
        bn = 256
    if two_cta and out_f32 and split_k == 0 and f32_mode in (0, 1) and bn == 256:
        # small weight-gradient outputs with a huge K (tokens): too few 256x256 tiles for the 74 CTA pairs.  256x128
        # tiles fill the machine without split-K, i.e. without a zero-fill launch and without fp32 atomics (deterministic
        # gradients); measured equal to split-K = 2 + fill on B200 (tools/sweep_wgrad.py).
        tiles = ((M + 255) // 256) * ((N + bn - 1) // bn)
        if tiles <= 48 and K >= 2048:
            bn = 128
    rc = fn(_p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), out.stride(0),
            int(a_mn), int(b_mn), int(out_f32), f32_mode, float(alpha), _p(bias), int(relu),
            _p(mask), 0 if mask is None else mask.stride(0), _p(stats), _p(relu_bits_out), _p(relu_bits), bn, split_k,
            _st())
    _chk(rc, "gemm")
    if colsum_out is not None:
        colsum(out, out=colsum_out, accumulate=True)
    return out


FUSED_COLSUM = os.environ.get("B200_FUSED_COLSUM", "1") == "1"


def gemm_stats_entries(N: int, bn: int = 256) -> int:
    """Rows of the `stats` tensor ([entries, M, 2] fp32) the GEMM epilogue fills: one (max, sum exp) pair per n-tile of width
    `bn` and per epilogue warp group (two groups for tiles of 128 or 256 columns)."""
    return (2 if bn >= 128 else 1) * ((N + bn - 1) // bn)


def gemm_skinny(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, *, bias=None, relu: bool = False,
                split_k: int = 0) -> torch.Tensor:
    """out[M,N] = act(x[M,K] @ w[N,K]^T + bias) for M <= 128 (decode): the weight-streaming tcgen05 kernel
    (K split across a thread-block cluster, partials reduced through distributed shared memory)."""
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, device=x.device, dtype=BF16)
    _chk(lib.load().b200_gemm_skinny(_p(x), _p(w), _p(out), M, N, K, x.stride(0), w.stride(0), out.stride(0), _p(bias),
                                     int(relu), split_k, _st()), "gemm_skinny")
    return out


def gemm_pick_bn(M: int, N: int, K: int, out_f32: bool = False) -> int:
    return lib.load().b200_gemm_pick_bn(M, N, K, int(out_f32))


# ------------------------------------------------------------------------------------------ elementwise
def cast_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert x.dtype == F32 and x.is_contiguous()
    if out is None:
        out = torch.empty_like(x, dtype=BF16)
    _chk(lib.load().b200_cast_f32_bf16(_p(x), _p(out), x.numel(), _st()), "cast")
    return out


def colsum(a: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    M, N = a.shape
    if out is None:
        out = torch.zeros(N, device=a.device, dtype=F32)
    _chk(lib.load().b200_colsum_bf16(_p(a), _p(out), M, N, a.stride(0), int(accumulate), _st()), "colsum")
    return out


def embed_fwd(idx, wte, wpe, out=None, pos0: int = 0, pos_dev=None):
    B, T = idx.shape
    V, d = wte.shape
    assert idx.dtype == torch.int64 and idx.is_contiguous()
    if out is None:
        out = torch.empty(B * T, d, device=wte.device, dtype=F32)
    _chk(lib.load().b200_embed_fwd(_p(idx), _p(wte), _p(wpe), _p(out), B, T, d, V, pos0, _p(pos_dev), _st()),
         "embed_fwd")
    return out


def embed_bwd(idx, dx, dwte, dwpe, accumulate: bool = False):
    B, T = idx.shape
    V, d = dwte.shape
    _chk(lib.load().b200_embed_bwd(_p(idx), _p(dx), _p(dwte), _p(dwpe), B, T, d, V, int(accumulate), _st()),
         "embed_bwd")


def embed_bwd_sorted(idx, order, dx, dwte, dwpe, accumulate: bool = False):
    """Deterministic embedding backward: `order` is a stable argsort of idx.view(-1)."""
    B, T = idx.shape
    V, d = dwte.shape
    _chk(lib.load().b200_embed_bwd_sorted(_p(idx), _p(order), _p(dx), _p(dwte), _p(dwpe), B, T, d, V, int(accumulate), _st()),
         "embed_bwd")


def zero_(t: torch.Tensor) -> torch.Tensor:
    """Zero a contiguous tensor with a memset node (no kernel launch; capturable)."""
    assert t.is_contiguous()
    rc = lib.load().b200_memset_zero(_p(t), t.numel() * t.element_size(), _st())
    if rc != 0:
        lib.check(rc, "memset_zero")
    return t


def ln_fwd(x, gamma, beta, *, branch=None, x_out=None, out=None, out_f32=None, mean=None, rstd=None,
           eps: float = 1e-5, want_stats: bool = True):
    """r = x (+ branch -> stored to x_out, default in place) -> bf16 LN(r) (+ mean, rstd)."""
    if branch is not None and x_out is None:
        x_out = x
    M, d = x.shape
    if out is None and out_f32 is None:
        out = torch.empty(M, d, device=x.device, dtype=BF16)
    if want_stats and mean is None:
        mean = torch.empty(M, device=x.device, dtype=F32)
        rstd = torch.empty(M, device=x.device, dtype=F32)
    _chk(lib.load().b200_ln_fwd(_p(x), _p(branch), _p(x_out), _p(gamma), _p(beta), _p(out), _p(out_f32), _p(mean), _p(rstd),
                                M, d, eps, _st()), "ln_fwd")
    return out if out is not None else out_f32, mean, rstd


def ln_bwd_nparts(M: int, d: int) -> int:
    return lib.load().b200_ln_bwd_nparts(M, d)


def ln_bwd(dy, x, mean, rstd, gamma, dgamma, dbeta, *, dx_in=None, dx_out=None, dx_bf16=None, colsum_out=None,
           partial=None, accumulate: bool = False):
    M, d = x.shape
    if dx_out is None:
        dx_out = torch.empty(M, d, device=x.device, dtype=F32)
    if partial is None:
        partial = torch.empty(ln_bwd_nparts(M, d) * 3 * d, device=x.device, dtype=F32)
    _chk(lib.load().b200_ln_bwd(_p(dy), _p(x), _p(mean), _p(rstd), _p(gamma), _p(dx_in), _p(dx_out), _p(dx_bf16),
                                _p(partial), _p(dgamma), _p(dbeta), _p(colsum_out), M, d, int(accumulate), _st()),
         "ln_bwd")
    return dx_out


def ln_bwd_partial(dy, x, mean, rstd, gamma, partial, *, dx_in=None, dx_out=None, dx_bf16=None, with_colsum: bool = False):
    """Main kernel of ln_bwd only: dx (+ residual gradient, + bf16 copy); the parameter-gradient partial sums stay in
    `partial` until ln_bwd_finalize2 reduces them (two LayerNorms per launch)."""
    M, d = x.shape
    _chk(lib.load().b200_ln_bwd_partial(_p(dy), _p(x), _p(mean), _p(rstd), _p(gamma), _p(dx_in), _p(dx_out), _p(dx_bf16),
                                        _p(partial), int(with_colsum), M, d, _st()), "ln_bwd_partial")
    return dx_out


def ln_bwd_finalize2(M, d, partial_a, dgamma_a, dbeta_a, colsum_a, partial_b, dgamma_b, dbeta_b, colsum_b,
                     accumulate: bool = False):
    _chk(lib.load().b200_ln_bwd_finalize2(_p(partial_a), _p(dgamma_a), _p(dbeta_a), _p(colsum_a), _p(partial_b), _p(dgamma_b),
                                          _p(dbeta_b), _p(colsum_b), M, d, int(accumulate), _st()), "ln_bwd_finalize2")


def adamw(p, g, m, v, p_bf16, hyper):
    _chk(lib.load().b200_adamw(_p(p), _p(g), _p(m), _p(v), _p(p_bf16), _p(hyper), p.numel(), _st()), "adamw")


def adamw_hyper(lr: float, step: int, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, grad_scale=1.0):
    """The 8 floats b200_adamw reads (same derivations as torch.optim.AdamW's single step)."""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    return [1.0 - lr * weight_decay, beta2, 1.0 - beta1, 1.0 - beta2, lr / bc1, math.sqrt(bc2), eps, grad_scale]


# ------------------------------------------------------------------------------------------ CE
def ce_finalize(stats, n_tiles, logits, targets, lse, loss_rows, loss_sum, V, sum_scale: float = 1.0):
    M = logits.shape[0]
    _chk(lib.load().b200_ce_finalize(_p(stats), n_tiles, _p(logits), logits.stride(0), _p(targets), _p(lse),
                                     _p(loss_rows), _p(loss_sum), M, V, float(sum_scale), _st()), "ce_finalize")


def ce_grad(logits, lse, targets, dbias, V, scale: float, accumulate: bool = False):
    M = logits.shape[0]
    _chk(lib.load().b200_ce_grad(_p(logits), logits.stride(0), _p(lse), _p(targets), _p(dbias), M, V, scale,
                                 int(accumulate), _st()), "ce_grad")


# ------------------------------------------------------------------------------------------ attention
ATTN_FWD_IMPL = os.environ.get("B200_ATTN_FWD", "tc")   # "tc" = tcgen05/TMEM kernel, "mma" = mma.sync kernel


def attn_fwd(qkv, B, T, H, hs, out=None, lse=None, impl: Optional[str] = None):
    if out is None:
        out = torch.empty(B * T, H * hs, device=qkv.device, dtype=BF16)
    if lse is None:
        lse = torch.empty(B, H, T, device=qkv.device, dtype=F32)
    L = lib.load()
    which = impl or ATTN_FWD_IMPL
    if which in ("tc1", "tc2"):      # tcgen05 kernel with 1 / 2 query tiles per CTA forced (tests, benchmarks)
        _chk(L.b200_attn_fwd_tc_tiles(_p(qkv), _p(out), _p(lse), B, T, H, hs, 1.0 / math.sqrt(hs), int(which[2]), _st()),
             "attn_fwd")
        return out, lse
    fn = L.b200_attn_fwd_tc if which == "tc" else L.b200_attn_fwd
    _chk(fn(_p(qkv), _p(out), _p(lse), B, T, H, hs, 1.0 / math.sqrt(hs), _st()), "attn_fwd")
    return out, lse


ATTN_BWD_IMPL = os.environ.get("B200_ATTN_BWD", "tc")


def attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=None, delta=None, dq_acc=None, counters=None,
             impl: Optional[str] = None):
    """Default: the tcgen05 two-pass backward (only `delta` scratch).  impl="mma": the mma.sync kernel, which
    also needs dq_acc (fp32 [B*T, H*hs], zero on entry, left zero)."""
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    if delta is None:
        delta = torch.empty(B, H, T, device=qkv.device, dtype=F32)
    if (impl or ATTN_BWD_IMPL) == "tc":
        _chk(lib.load().b200_attn_bwd_tc(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dqkv), B, T, H, hs,
                                         1.0 / math.sqrt(hs), _st()), "attn_bwd_tc")
    else:
        if dq_acc is None:
            dq_acc = torch.zeros(B * T, H * hs, device=qkv.device, dtype=F32)
        _chk(lib.load().b200_attn_bwd(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dq_acc), _p(dqkv), B, T, H,
                                      hs, 1.0 / math.sqrt(hs), _st()), "attn_bwd")
    return dqkv


def kv_append(qkv, kcache, vcache, B, t, H, hs, Tmax, pos0=0, pos_dev=None):
    _chk(lib.load().b200_kv_append(_p(qkv), _p(kcache), _p(vcache), B, t, H, hs, Tmax, pos0, _p(pos_dev), _st()),
         "kv_append")


def attn_decode(qkv, kcache, vcache, out, B, H, hs, Tmax, length=1, pos_dev=None, append: bool = False):
    _chk(lib.load().b200_attn_decode(_p(qkv), _p(kcache), _p(vcache), _p(out), B, H, hs, Tmax, length,
                                     1.0 / math.sqrt(hs), _p(pos_dev), int(append), _st()), "attn_decode")


def attn_decode_split(qkv, kcache, vcache, out, B, H, hs, Tmax, scratch, tickets, splits: int, length=1, pos_dev=None,
                      append: bool = False):
    """Split-KV decode attention: scratch fp32 [B*H*splits*(hs+2)], tickets int32 [B*H] (zeroed once)."""
    assert splits == 1 or (scratch.numel() >= B * H * splits * (hs + 2) and tickets.numel() >= B * H)
    _chk(lib.load().b200_attn_decode_split(_p(qkv), _p(kcache), _p(vcache), _p(out), _p(scratch), _p(tickets), B, H, hs, Tmax,
                                           length, 1.0 / math.sqrt(hs), _p(pos_dev), int(append), splits, _st()),
         "attn_decode")


def sample(logits, next_tok, V, temperature=1.0, seed=0, step=0, pos_dev=None):
    B = logits.shape[0]
    _chk(lib.load().b200_sample(_p(logits), logits.stride(0), _p(next_tok), B, V, float(temperature), 0, seed, step,
                                _p(pos_dev), _st()), "sample")


def incr_int(t: torch.Tensor, delta: int = 1):
    _chk(lib.load().b200_incr_int(_p(t), delta, _st()), "incr_int")


def l2_flush(buf: torch.Tensor):
    _chk(lib.load().b200_l2_flush(_p(buf), buf.numel() * buf.element_size(), _st()), "l2_flush")


def scatter_tokens(next_tok, seq, col=0, pos_dev=None):
    B = next_tok.shape[0]
    _chk(lib.load().b200_scatter_tokens(_p(next_tok), _p(seq), B, seq.stride(0), col, _p(pos_dev), _st()),
         "scatter_tokens")


def gather_last_rows(x, out, B, T):
    _chk(lib.load().b200_gather_last_rows(_p(x), _p(out), B, T, x.shape[-1] * x.element_size(), _st()),
         "gather_last_rows")
